#!/bin/bash
# round-2 GPU call P (2 GPUs): fix-up kernel fast path + blinding terms on a host thread (parity, trace, proof time); multi-device C ABI with page-locked witnesses
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_multi_c_abi.py tests/test_gpu_sharded_single.py -m gpu -x -q -k "msm or prove or tree or multi or sharded" > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2p_pytest.log
for g in 2 1; do
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 $g 2> gpurun_out/r2p_trace_g$g.err | tee gpurun_out/r2p_trace_g$g.txt
python tools/trace_summary.py gpurun_out/r2p_trace_g$g.err | grep -E "total|fix_pieces|xyzz|reduce"
done
timeout 200 python tools/prof_prove.py 22 6 2>&1 | grep "^prove" | tail -4
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu --skip-standalone > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2p_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
PY
timeout 300 python tools/multi_c_abi.py 22 39 > gpurun_out/r2p_multi_c_abi_n2.json 2> gpurun_out/r2p_multi_c_abi.err; echo "multi_c_abi rc=$?"; tail -2 gpurun_out/r2p_multi_c_abi.err; cat gpurun_out/r2p_multi_c_abi_n2.json
