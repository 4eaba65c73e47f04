#!/bin/bash
# round-2 GPU call I: inlined lazy Fq2 products in the G2 tree fast path, G1 tree kernels pinned to 128 registers;
# ncu --set full of every accumulation-stage kernel of one 2^22 multiexp (G1 and G2)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "msm or prove_vs_oracle or prove_golden or tree" > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2i_pytest.log
for g in 2 1; do
B2G16_MSM_TRACE=1 timeout 200 python tools/msm_trace.py 22 $g 2> gpurun_out/r2i_trace_g$g.err | tee gpurun_out/r2i_trace_g$g.txt
python tools/trace_summary.py gpurun_out/r2i_trace_g$g.err | tee -a gpurun_out/r2i_trace_g$g.txt
done
timeout 600 python bench.py --steps 5 --warmup 3 --skip-32gib --skip-cpu > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?"
tail -c 300 gpurun_out/r2i_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2i_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "acc", [(x["group"], round(x["ms"], 2)) for x in d["accumulate_launches_in_proof"]])
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
PY
for g in 1 2; do
timeout 120 python tools/prof_msm_once.py 22 $g > gpurun_out/r2i_once_g$g.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:"tree_|msm_accumulate|msm_fix|scan_" -o /tmp/r2i_msm_g$g python tools/prof_msm_once.py 22 $g > gpurun_out/r2i_ncu_g$g.log 2>&1
echo "ncu g$g rc=$?"
ncu -i /tmp/r2i_msm_g$g.ncu-rep --page raw --csv > gpurun_out/r2i_msm_g${g}_raw.csv 2>/dev/null
python tools/ncu_stage_summary.py gpurun_out/r2i_msm_g${g}_raw.csv gpurun_out/r2i_ncu_full_msm_g${g}_2p22.json "ncu --set full --clock-control none of every accumulation-stage launch of ONE group-$g multiexp, 2^22 dense bases, uniform scalars (tools/prof_msm_once.py 22 $g)"
gzip -f gpurun_out/r2i_msm_g${g}_raw.csv
done
ls -la gpurun_out | tail -12
