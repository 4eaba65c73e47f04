#!/bin/bash
# round-2 GPU call H: full -m gpu suite (new: streamed key load, multi-device C ABI), full default bench line
mkdir -p gpurun_out
timeout 1300 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -16 gpurun_out/r2h_pytest.log
timeout 1500 python bench.py > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2h_bench.err
head -c 6000 gpurun_out/r2h_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2h_bench_ref.json 2> gpurun_out/r2h_bench_ref.err; echo "ref rc=$?"
head -c 1500 gpurun_out/r2h_bench_ref.json
