#!/bin/bash
# round-2 GPU call 2: tree accumulation parity + sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm or prove_vs_oracle or prove_golden or tree" > gpurun_out/r2_2_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2_2_pytest.log
timeout 600 python tools/tree_sweep.py 22 1 > gpurun_out/r2_2_sweep_g1.jsonl 2> gpurun_out/r2_2_sweep_g1.err; echo "sweep g1 rc=$?"; tail -3 gpurun_out/r2_2_sweep_g1.err
cat gpurun_out/r2_2_sweep_g1.jsonl
timeout 600 python tools/tree_sweep.py 20 2 > gpurun_out/r2_2_sweep_g2.jsonl 2> gpurun_out/r2_2_sweep_g2.err; echo "sweep g2 rc=$?"; tail -3 gpurun_out/r2_2_sweep_g2.err
cat gpurun_out/r2_2_sweep_g2.jsonl
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-standalone --skip-cpu > gpurun_out/r2_2_bench.json 2> gpurun_out/r2_2_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_2_bench.err
head -c 700 gpurun_out/r2_2_bench.json
