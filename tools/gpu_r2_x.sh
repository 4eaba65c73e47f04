#!/bin/bash
# round-2 GPU call X: prove_batch with the witness upload overlapped (parity, priority test, timing)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_verify.py tests/test_multi_c_abi.py tests/test_cpp_shim.py -m gpu -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2x_pytest.log
timeout 300 python tools/batch_time.py 22 4 | tee gpurun_out/r2x_batch_time.json
