#!/bin/bash
# round-2 GPU call M: TMA tile staging in the NTT (parity + A/B timing + launch list / ncu of the TMA pass)
mkdir -p gpurun_out
B2G16_NTT_TMA=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ntt or h_from_abc or prove_vs_oracle or prove_golden" > gpurun_out/r2m_pytest_tma.log 2>&1; echo "pytest(tma) rc=$?"; tail -4 gpurun_out/r2m_pytest_tma.log
for t in 0 1; do
echo "== B2G16_NTT_TMA=$t"
B2G16_NTT_TMA=$t timeout 300 python tools/probe.py 27 0 2>/dev/null | grep ntt_log_n | tail -4
done | tee gpurun_out/r2m_ntt_tma_ab.txt
for t in 0 1; do
echo "== proofs B2G16_NTT_TMA=$t"; B2G16_NTT_TMA=$t timeout 200 python tools/prof_prove.py 22 5 2>&1 | grep "^prove" | tail -3
done | tee -a gpurun_out/r2m_ntt_tma_ab.txt
for t in 0 1; do
B2G16_NTT_TMA=$t timeout 300 ncu --set full --clock-control none -k regex:ntt_pass -s 6 -c 3 -o /tmp/r2m_ntt_tma$t python tools/prof_one.py 0 24 > gpurun_out/r2m_ncu_tma$t.log 2>&1; echo "ncu tma=$t rc=$?"
ncu -i /tmp/r2m_ntt_tma$t.ncu-rep --page raw --csv > gpurun_out/r2m_ntt_tma${t}_raw.csv 2>/dev/null
python tools/ncu_stage_summary.py gpurun_out/r2m_ntt_tma${t}_raw.csv gpurun_out/r2m_ncu_full_ntt_tma$t.json "ncu --set full of the three ntt_pass_kernel launches of one 2^24 transform, B2G16_NTT_TMA=$t"
python - <<PY
import csv
rows=list(csv.reader(open("gpurun_out/r2m_ntt_tma${t}_raw.csv")))
h=rows[0]; ix={k:i for i,k in enumerate(h)}
for r in rows[2:]:
    print(r[ix["Kernel Name"]][:60], r[ix["gpu__time_duration.sum"]], r[ix["sm__throughput.avg.pct_of_peak_sustained_elapsed"]], r[ix["launch__registers_per_thread"]], r[ix["dram__bytes_read.sum"]])
PY
gzip -f gpurun_out/r2m_ntt_tma${t}_raw.csv
done
