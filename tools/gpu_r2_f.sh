#!/bin/bash
# round-2 GPU call F: register-round NTT (three launch shapes), G2 tree kernels at 3 / 4 CTAs per SM
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ntt or h_from_abc or prove_vs_oracle or prove_golden or msm" > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2f_pytest.log
for v in 0 1 2; do
echo "== NTT variant $v"
B2G16_NTT_VARIANT=$v timeout 300 python tools/probe.py 27 0 2>/dev/null | grep ntt_log_n | tail -4
done | tee gpurun_out/r2f_ntt_variants.txt
echo "== G2 tree kernels, 3 CTAs per SM"
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 2 2> gpurun_out/r2f_trace_g2_mb3.err; python tools/trace_summary.py gpurun_out/r2f_trace_g2_mb3.err
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 1 2> gpurun_out/r2f_trace_g1.err; python tools/trace_summary.py gpurun_out/r2f_trace_g1.err; head -3 gpurun_out/r2f_trace_g1.err
echo "== G2 tree kernels, 4 CTAs per SM"
cp crypto3-fil-proofs_b200/lib/libb2g16.so /tmp/libb2g16.so.keep
cp crypto3-fil-proofs_b200/lib_alt/libb2g16_mb4.so crypto3-fil-proofs_b200/lib/libb2g16.so
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 2 2> gpurun_out/r2f_trace_g2_mb4.err; python tools/trace_summary.py gpurun_out/r2f_trace_g2_mb4.err
cp /tmp/libb2g16.so.keep crypto3-fil-proofs_b200/lib/libb2g16.so
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
tail -c 500 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2f_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
PY
