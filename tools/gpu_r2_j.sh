#!/bin/bash
# round-2 GPU call J: L2 prefetch in the tree kernels, priority gate; proofs with the multiexps concurrent vs serialised; streamed key load
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_multi_c_abi.py tests/test_gpu_verify.py -m gpu -x -q -k "msm or prove_vs_oracle or prove_golden or tree or multi or priority or streamed" > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2j_pytest.log
for g in 2 1; do
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 $g 2> gpurun_out/r2j_trace_g$g.err | tee gpurun_out/r2j_trace_g$g.txt
python tools/trace_summary.py gpurun_out/r2j_trace_g$g.err | tee -a gpurun_out/r2j_trace_g$g.txt
done
echo "== proofs, multiexps concurrent (default)"; timeout 200 python tools/prof_prove.py 22 6 2>&1 | grep "^prove" | tail -4
echo "== proofs, multiexps serialised"; B2G16_SERIAL_LOG=0 timeout 200 python tools/prof_prove.py 22 6 2>&1 | grep "^prove" | tail -4
for i in 1 2; do
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu --skip-standalone > gpurun_out/r2j_bench_$i.json 2> gpurun_out/r2j_bench_$i.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r2j_bench_$i.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
PY
done
timeout 600 python tools/key_cache_time.py 22 /tmp > gpurun_out/r2j_key_cache_time_2p22.json 2> gpurun_out/r2j_key_cache.err; echo "key cache rc=$?"; tail -3 gpurun_out/r2j_key_cache.err; cat gpurun_out/r2j_key_cache_time_2p22.json
