#!/bin/bash
# round-2 GPU call A: full -m gpu suite on the current build, then a short bench line (2^22 proof + standalone MSM/NTT)
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpu.txt; nproc >> gpurun_out/r2a_gpu.txt
timeout 1300 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -40 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -c 800 gpurun_out/r2a_bench.err
head -c 3000 gpurun_out/r2a_bench.json
