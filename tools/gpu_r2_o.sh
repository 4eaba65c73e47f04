#!/bin/bash
# round-2 GPU call O (4 GPUs): shim value arithmetic (C++ consumer), multi-GPU parity under torchrun, bench --gpus 4
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cpp_shim.py tests/test_gpu_multi.py tests/test_multi_c_abi.py -m gpu -x -q > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2o_pytest.log
timeout 300 python tools/multi_c_abi.py 22 39 > gpurun_out/r2o_multi_c_abi_n4.json 2> gpurun_out/r2o_multi_c_abi.err; echo "multi_c_abi rc=$?"; cat gpurun_out/r2o_multi_c_abi_n4.json
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2o_bench_n4.json 2> gpurun_out/r2o_bench_n4.err; echo "bench rc=$?"
grep -v "^W1105\|^\*\*\*\*\|OMP_NUM" gpurun_out/r2o_bench_n4.err | tail -c 800
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2o_bench_n4.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"])
print("one_proof", d["one_proof_sharded"]["prove_s"], (d["one_proof_sharded"].get("witness_map_sharded_too") or {}).get("prove_s"))
print("msm", d["msm_g1_sharded"]["ms"], "ntt", {k: v["ms"] for k, v in d["ntt_fr_sharded"].items() if isinstance(v, dict) and "ms" in v})
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    v = d.get(k) or {}; sh = v.get("sharded_over_all_gpus") or {}
    print(k, v.get("prove_s"), sh.get("prove_s"), (sh.get("witness_map_sharded_too") or {}).get("prove_s"))
PY
