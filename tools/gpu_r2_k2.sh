#!/bin/bash
# round-2 GPU call K2 (2 GPUs): bench --gpus 2 after the memory fixes (torch cache emptied before the 2^27 legs, tree scratch retry), streamed key load
mkdir -p gpurun_out
timeout 300 python tools/key_cache_time.py 22 /tmp > gpurun_out/r2k_key_cache_time_2p22.json 2> gpurun_out/r2k_key_cache.err; echo "key cache rc=$?"; tail -2 gpurun_out/r2k_key_cache.err; cut -c1-420 gpurun_out/r2k_key_cache_time_2p22.json
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2k_bench_n2.json 2> gpurun_out/r2k_bench_n2.err; echo "bench rc=$?"
grep -v "^W1105\|^\*\*\*\*\|OMP_NUM" gpurun_out/r2k_bench_n2.err | tail -c 1500
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2k_bench_n2.json").read().strip().splitlines()[-1])
for k in ("value", "e2e", "one_proof_sharded", "msm_g1_sharded", "ntt_fr_sharded"):
    print(k, json.dumps(d.get(k))[:900])
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    v = d.get(k) or {}
    print(k, v.get("prove_s"), v.get("hbm_used_GB"), json.dumps(v.get("sharded_over_all_gpus"))[:700])
PY
