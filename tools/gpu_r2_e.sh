#!/bin/bash
# round-2 GPU call E: lazy-reduction Fq2 product, x-only up0 loads, bit-decomposition reduce tail: parity, step trace, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_verify.py -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2e_pytest.log
for g in 1 2; do
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 $g 2> gpurun_out/r2e_trace_g$g.err | tee gpurun_out/r2e_trace_g$g.txt
python tools/trace_summary.py gpurun_out/r2e_trace_g$g.err | tee -a gpurun_out/r2e_trace_g$g.txt
done
B2G16_RED_BITS=0 timeout 200 python tools/msm_trace.py 22 1 2>/dev/null
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench rc=$?"
tail -c 500 gpurun_out/r2e_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2e_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
PY
