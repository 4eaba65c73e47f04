#!/bin/bash
# round-2 GPU call G: NTT launch shapes (stage-by-stage vs register rounds with twiddle prefetch; with / without streamed twiddle
# tables), C-ABI multi-device scheduling test, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_c_abi.py tests/test_gpu_parity.py -m gpu -x -q -k "multi or ntt or h_from_abc or prove_vs_oracle" > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2g_pytest.log
for tw in "" 1; do for v in 0 1 2 3; do
echo "== NTT variant $v no_tw_tables=$tw"
if [ -n "$tw" ]; then export B2G16_NTT_NO_TW_TABLES=1; else unset B2G16_NTT_NO_TW_TABLES; fi
B2G16_NTT_VARIANT=$v timeout 300 python tools/probe.py 27 0 2>/dev/null | grep ntt_log_n | tail -3
done; done | tee gpurun_out/r2g_ntt_variants.txt
unset B2G16_NTT_NO_TW_TABLES
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"
tail -c 500 gpurun_out/r2g_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2g_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
PY
