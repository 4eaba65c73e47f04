#!/bin/bash
# round-2 GPU call K (2 GPUs): multi-GPU parity under torchrun, the multi-device C ABI from one process, bench --gpus 2
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_multi_c_abi.py -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2k_pytest.log
timeout 600 python tools/multi_c_abi.py 22 39 > gpurun_out/r2k_multi_c_abi_n2.json 2> gpurun_out/r2k_multi_c_abi.err; echo "multi_c_abi rc=$?"; tail -3 gpurun_out/r2k_multi_c_abi.err; cat gpurun_out/r2k_multi_c_abi_n2.json
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2k_bench_n2.json 2> gpurun_out/r2k_bench_n2.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2k_bench_n2.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2k_bench_n2.json").read().strip().splitlines()[-1])
for k in ("value", "e2e", "one_proof_sharded", "msm_g1_sharded", "ntt_fr_sharded"):
    print(k, json.dumps(d.get(k))[:700])
for k in ("porep_c2_32GiB_partition", "window_post_32GiB"):
    v = d.get(k) or {}
    print(k, {a: v.get(a) for a in ("prove_s", "sharded", "sharded_full", "all_partitions")}, json.dumps(v)[:200])
PY
