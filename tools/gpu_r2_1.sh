#!/bin/bash
# round-2 GPU call 1: new parity tests, full bench line, ncu full of the accumulate + NTT kernels (baseline evidence)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_1_gpu.txt; nproc >> gpurun_out/r2_1_gpu.txt
timeout 1100 python -m pytest tests -m gpu -x -q > gpurun_out/r2_1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_1_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_1_bench.json 2> gpurun_out/r2_1_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2_1_bench.err
timeout 120 python tools/prof_one.py 22 24 > gpurun_out/r2_1_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"msm_accumulate|ntt_pass" -c 8 -o gpurun_out/r2_1_prof python tools/prof_one.py 22 24 > gpurun_out/r2_1_ncu.log 2>&1
echo "ncu rc=$?"
tail -5 gpurun_out/r2_1_pytest.log
head -c 1500 gpurun_out/r2_1_bench.json
