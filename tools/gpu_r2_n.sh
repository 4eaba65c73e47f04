#!/bin/bash
# round-2 GPU call N: final build -- full -m gpu suite, default bench line, reference arm, ncu A/B of the TMA / non-TMA contiguous pass
mkdir -p gpurun_out
timeout 1300 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -10 gpurun_out/r2n_pytest.log
timeout 1500 python bench.py > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r2n_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2n_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "util", d["roofline"]["pipe_utilisation_actual"], "traffic", d["roofline"]["traffic"])
print("parity", d["parity"]); print("cpu", d.get("cpu_baseline", {}).get("value"))
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    print(k, (d.get(k) or {}).get("prove_s"), (d.get(k) or {}).get("hbm_used_GB"))
PY
for t in 0 1; do
timeout 120 env B2G16_NTT_TMA=$t python tools/prof_one.py 0 24 > gpurun_out/r2n_plain_tma$t.log 2>&1 && \
B2G16_NTT_TMA=$t timeout 300 ncu --set full --clock-control none -k regex:ntt_pass -s 3 -c 3 -o /tmp/r2n_ntt_tma$t python tools/prof_one.py 0 24 > gpurun_out/r2n_ncu_tma$t.log 2>&1; echo "ncu tma=$t rc=$?"
ncu -i /tmp/r2n_ntt_tma$t.ncu-rep --page raw --csv > gpurun_out/r2n_ntt_tma${t}_raw.csv 2>/dev/null
python tools/ncu_stage_summary.py gpurun_out/r2n_ntt_tma${t}_raw.csv gpurun_out/r2n_ncu_full_ntt_2p24_tma$t.json "ncu --set full of the three ntt_pass_kernel launches of one 2^24 transform, B2G16_NTT_TMA=$t"
gzip -f gpurun_out/r2n_ntt_tma${t}_raw.csv
done
