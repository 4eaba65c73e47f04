#!/bin/bash
# round-2 GPU call W: final build -- full -m gpu suite, smoke, default bench line
mkdir -p gpurun_out
timeout 1300 python -m pytest tests -m gpu -x -q > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2w_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1500 python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r2w_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2w_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "util", d["roofline"]["pipe_utilisation_actual"], "launches", d["gpu_launches"])
print("parity", d["parity"]); print("cpu", d.get("cpu_baseline", {}).get("value"), "clocks", d["clocks"])
print("standalone", {k: v.get("ms") for k, v in d.get("standalone", {}).items()})
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    print(k, (d.get(k) or {}).get("prove_s"), (d.get(k) or {}).get("hbm_used_GB"))
PY
