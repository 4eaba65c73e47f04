#!/bin/bash
# round-2 GPU call AB: G2 tree addition pass with out-of-line wide product / reduction (smaller loop body) against the inlined form
echo "== inlined (default)"
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 2 2> /tmp/a.err | head -1; python tools/trace_summary.py /tmp/a.err | grep -E "total|down0|up0"
timeout 200 python tools/prof_prove.py 22 5 2>&1 | grep "^prove" | tail -2
echo "== out-of-line pieces"
cp crypto3-fil-proofs_b200/lib/libb2g16.so /tmp/keep.so; cp crypto3-fil-proofs_b200/lib_alt/libb2g16_g2calls.so crypto3-fil-proofs_b200/lib/libb2g16.so
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm_g2 or tree or prove_vs_oracle" 2>&1 | tail -1
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 2 2> /tmp/b.err | head -1; python tools/trace_summary.py /tmp/b.err | grep -E "total|down0|up0"
timeout 200 python tools/prof_prove.py 22 5 2>&1 | grep "^prove" | tail -2
cp /tmp/keep.so crypto3-fil-proofs_b200/lib/libb2g16.so
