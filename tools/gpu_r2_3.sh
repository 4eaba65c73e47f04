#!/bin/bash
# round-2 GPU call 3: NTT with combined twiddles + Fr reduction (parity, time); launch list + ncu of the tree kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ntt or h_from_abc or field or fr32 or prove_vs_oracle or prove_golden" > gpurun_out/r2_3_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2_3_pytest.log
timeout 300 python tools/probe.py 27 0 > gpurun_out/r2_3_probe.jsonl 2> gpurun_out/r2_3_probe.err; echo "probe rc=$?"; tail -3 gpurun_out/r2_3_probe.err; cat gpurun_out/r2_3_probe.jsonl | cut -c1-300
B2G16_NTT_NO_TW_TABLES=1 timeout 300 python tools/probe.py 27 0 > gpurun_out/r2_3_probe_notw.jsonl 2>/dev/null; tail -2 gpurun_out/r2_3_probe_notw.jsonl | cut -c1-300
timeout 120 python tools/prof_one.py 22 0 > gpurun_out/r2_3_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_3_launches_msm.csv python tools/prof_one.py 22 0 > gpurun_out/r2_3_ncu1.log 2>&1
echo "ncu1 rc=$?"
python tools/launch_summary.py gpurun_out/r2_3_launches_msm.csv 70
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tree_down0|tree_up0" -s 2 -c 4 -o gpurun_out/r2_3_prof_tree python tools/prof_one.py 22 0 > gpurun_out/r2_3_ncu2.log 2>&1
echo "ncu2 rc=$?"
