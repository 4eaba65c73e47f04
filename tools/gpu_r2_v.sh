#!/bin/bash
# round-2 GPU call V: inlined doubling + Euclidean inversion in the single-thread tail of the standalone multiexp
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_verify.py tests/test_cpp_shim.py -m gpu -x -q -k "msm or sum or verify or shim or prove_vs_oracle" > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2v_pytest.log
for g in 1 2; do for lg in 16 22; do timeout 100 python tools/msm_trace.py $lg $g 2>/dev/null | head -2; done; done
