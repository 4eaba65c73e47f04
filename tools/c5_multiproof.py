"""BASELINE configs[4]: PoRep C2 for the 32 GiB sector shape, 10 partitions of ~2^27 constraints under one key, on
all GPUs of the box:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/c5_multiproof.py [workload] [partitions]
Schedule (b2g16.dist.partition_schedule): full waves one partition per GPU, the remainder spread over sub-groups
(10 on 8 GPUs: 8 + 2 x 4-GPU proofs).  Witnesses are generated on the device beforehand (synthesis is outside the
path); rank 0 batch-verifies the 192 x P bytes natively and prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import numpy as np
import torch
import torch.distributed as dist
from b2g16 import codec, dist as bd, lib, shapes


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "porep-c2-32GiB"
    n_part = int(sys.argv[2]) if len(sys.argv) > 2 else shapes.POREP_PARTITIONS[32 << 30]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    ctx = lib.Context(local, ts.cuda_stream)
    log_m, n_in, n_free, seed = bench.WORKLOADS[workload][0], bench.n_in_of(workload), 8, 2026
    toxic, _, _ = bench.toxic_from_seed(seed)
    t0 = time.perf_counter()
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
    nv = pk.shape()["n_vars"]
    sched = bd.PartitionScheduler(n_part, ctx=ctx, log_m=log_m)
    heads, dzs = {}, {}
    for k in range(n_part):
        heads[k] = codec.random_fr(1 + n_in + n_free, np.random.Generator(np.random.PCG64(5000 + k)))
        heads[k][0] = 0; heads[k][0, 0] = 1
    for k, ranks in sched.plan:
        if rank in ranks:                                  # a witness per partition this rank takes part in
            dz = torch.zeros((nv, 32), dtype=torch.uint8, device=dev)
            dz[: heads[k].shape[0]] = torch.from_numpy(heads[k]).to(dev)
            ctx.synth_witness(pk, dz.data_ptr(), n_free)
            dzs[k] = dz
    rs = codec.random_fr(n_part, np.random.Generator(np.random.PCG64(77)))
    ss = codec.random_fr(n_part, np.random.Generator(np.random.PCG64(78)))
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    times, multi = [], None
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); dist.barrier(); e0.record()
        m = sched.prove_all(ctx, pk, lambda k: dzs[k].data_ptr(), rs, ss)
        e1.record(); torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()) / 1e3)
        assert multi is None or m == multi
        multi = m
    if rank == 0:
        pvk = ctx.vk_prepare(vk)
        pubs = np.stack([heads[k][1:1 + n_in] for k in range(n_part)])
        t0 = time.perf_counter()
        ok = ctx.verify_batch(pvk, pubs, multi)
        verify_s = time.perf_counter() - t0
        bad = pubs.copy(); bad[n_part - 1, 0, 0] ^= 1
        rejected = not ctx.verify_batch(pvk, bad, multi)
        pvk.free()
        print(json.dumps({"workload": workload, "partitions": n_part, "n_gpus": world, "log_m": log_m, "n_in": n_in, "multi_proof_bytes": len(multi),
                          "prove_all_s": min(times), "all_s": times, "setup_s": setup_s,
                          "schedule": [[k, ranks] for k, ranks in sched.plan], "batch_verified_natively": bool(ok), "perturbed_input_rejected": bool(rejected),
                          "verify_batch_s": verify_s}), flush=True)
    sched.close()
    pk.free(); ctx.close()
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
