"""One large proof over all GPUs of the box (latency mode; BASELINE configs[3] / configs[4]):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/latency_proof.py [workload] [reps]
Every rank generates the same synthetic key (device setup from the same seed), rank 0's witness is broadcast,
the proof's five multiexps are range-sharded (b2g16.dist.prove_sharded); rank 0 verifies the 192 bytes with the
native verifier (b2g16_verify) and prints one JSON line.  With N = 1 it times b2g16_prove_device for the same instance."""
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
from b2g16 import codec, dist as bd, lib


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "porep-c2-32GiB"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    mode = sys.argv[3] if len(sys.argv) > 3 else "msm"        # "msm": multiexps sharded, witness map repeated; "full": witness map sharded too
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    ctx = lib.Context(local, ts.cuda_stream)
    log_m, n_in, n_free, seed = bench.WORKLOADS[workload][0], bench.n_in_of(workload), 8, 2026
    toxic, r, sr = bench.toxic_from_seed(seed)
    t0 = time.perf_counter()
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
    shape = pk.shape()
    head = codec.random_fr(1 + n_in + n_free, np.random.Generator(np.random.PCG64(4242)))
    head[0] = 0; head[0, 0] = 1
    dz = torch.zeros((shape["n_vars"], 32), dtype=torch.uint8, device=dev)
    dz[: head.shape[0]] = torch.from_numpy(head).to(dev)
    ctx.synth_witness(pk, dz.data_ptr(), n_free)          # same seed and head on every rank: identical witnesses, no broadcast needed
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    times, proof = [], None
    peer = bd.PeerNtt(ctx, log_m) if (world > 1 and mode == "full") else None
    for _ in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        if world == 1:
            p = ctx.prove_device(pk, dz.data_ptr(), r, sr)
        elif peer is not None:
            p = bd.prove_sharded_full(ctx, pk, peer, dz.data_ptr(), r, sr)
        else:
            p = bd.prove_sharded(ctx, pk, dz.data_ptr(), r, sr)
        e1.record(); barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        times.append(ms / 1e3)
        assert proof is None or p == proof
        proof = p
    if rank == 0:
        pvk = ctx.vk_prepare(vk)
        t0 = time.perf_counter()
        ok = ctx.verify(pvk, head[1:1 + n_in], proof)
        verify_s = time.perf_counter() - t0
        bad = head[1:1 + n_in].copy(); bad[0, 0] ^= 1
        rejected = not ctx.verify(pvk, bad, proof)
        pvk.free()
        print(json.dumps({"workload": workload, "describe": bench.WORKLOADS[workload][2], "log_m": log_m, "n_in": n_in, "n_cons": shape["n_cons"], "n_gpus": world,
                          "prove_s": min(times[1:]), "all_s": times, "first_s_includes_table_setup": times[0], "setup_s": setup_s, "proof_head": proof[:8].hex(),
                          "verified_natively": bool(ok), "perturbed_input_rejected": bool(rejected), "verify_s": verify_s,
                          "mode": "single GPU" if world == 1 else ("five multiexps range-sharded over the ranks, witness map sharded (fused peer-memory transforms)" if peer is not None
                                                                     else "five multiexps range-sharded over the ranks, witness map on every rank")}), flush=True)
    if peer is not None:
        peer.close()
    pk.free(); ctx.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
