#!/bin/bash
# round-2 GPU call R: the no-allocation entry point -- its test alone in a fresh process, then the whole -m gpu suite (-x) with it last
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2r_pytest_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r2r_pytest_alone.log
timeout 1300 python -m pytest tests -m gpu -x -q > gpurun_out/r2r_pytest_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/r2r_pytest_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
