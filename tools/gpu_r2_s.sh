#!/bin/bash
# round-2 GPU call S: halves squaring in the G1 tree addition pass (A/B against the previous object), launch list of a 2^22 proof on the final build
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm or prove_vs_oracle or prove_golden or tree" > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2s_pytest.log
echo "== with halves squaring"
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 1 2> gpurun_out/r2s_trace_sqr.err | head -1; python tools/trace_summary.py gpurun_out/r2s_trace_sqr.err | grep -E "total|down0|up0"
timeout 200 python tools/prof_prove.py 22 6 2>&1 | grep "^prove" | tail -3
echo "== previous squaring"
cp crypto3-fil-proofs_b200/lib/libb2g16.so /tmp/keep.so; cp crypto3-fil-proofs_b200/lib_alt/libb2g16_nosqr.so crypto3-fil-proofs_b200/lib/libb2g16.so
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 1 2> gpurun_out/r2s_trace_nosqr.err | head -1; python tools/trace_summary.py gpurun_out/r2s_trace_nosqr.err | grep -E "total|down0|up0"
timeout 200 python tools/prof_prove.py 22 6 2>&1 | grep "^prove" | tail -3
cp /tmp/keep.so crypto3-fil-proofs_b200/lib/libb2g16.so
timeout 200 python tools/prof_prove.py 22 3 > gpurun_out/r2s_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2s_launches_prove_2p22.csv python tools/prof_prove.py 22 2 > gpurun_out/r2s_ncu1.log 2>&1
echo "ncu launches rc=$?"
L=$(grep -o "launches [0-9]*" gpurun_out/r2s_ncu1.log | tail -1 | cut -d' ' -f2)
python tools/launch_summary.py gpurun_out/r2s_launches_prove_2p22.csv $L | tee gpurun_out/r2s_launches_prove_2p22.txt
gzip -f gpurun_out/r2s_launches_prove_2p22.csv
