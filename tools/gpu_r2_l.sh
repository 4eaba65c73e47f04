#!/bin/bash
# round-2 GPU call L (8 GPUs): the multi-device C ABI from one process, then bench --gpus 8 (default flags: incl. the 2^27 proofs, the
# sharded forms and the 10-partition 32 GiB multi-proof)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python tools/multi_c_abi.py 22 39 > gpurun_out/r2l_multi_c_abi_n8.json 2> gpurun_out/r2l_multi_c_abi.err; echo "multi_c_abi rc=$?"; tail -3 gpurun_out/r2l_multi_c_abi.err; cat gpurun_out/r2l_multi_c_abi_n8.json
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2l_bench_n8.json 2> gpurun_out/r2l_bench_n8.err; echo "bench rc=$?"
grep -v "^W1105\|^\*\*\*\*\|OMP_NUM" gpurun_out/r2l_bench_n8.err | tail -c 1500
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2l_bench_n8.json").read().strip().splitlines()[-1])
for k in ("value", "e2e", "one_proof_sharded", "msm_g1_sharded", "ntt_fr_sharded"):
    print(k, json.dumps(d.get(k))[:900])
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    v = d.get(k) or {}
    print(k, v.get("prove_s"), v.get("hbm_used_GB"), json.dumps(v.get("sharded_over_all_gpus"))[:700])
    print(json.dumps(v.get("porep_c2_32GiB_all_partitions"))[:900])
PY
