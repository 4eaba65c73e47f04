"""BASELINE configs[2]: standalone G1 / G2 multiexp sweep 2^16 .. 2^26 and Fr transform sweep 2^16 .. 2^27 at 1 / 2 / 4 / 8 GPUs.
One process: single-GPU sweep (G1, G2, NTT; per-size time, throughput, roofline fractions; the multiexps checked by the
linearity identity).  Under torchrun: the sharded G1 multiexp and the sharded transform of bench.py at every size.
Usage: python tools/sweep_c3.py            |  torchrun --nproc-per-node N tools/sweep_c3.py      -> JSON lines"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from b2g16 import codec, lib

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(local, ts.cuda_stream)
hbm_peak, peak_src = bench.peaks()
mad_peak, _ = ctx.mad32_peak()


def timed(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def emit(d):
    if rank == 0:
        print(json.dumps(d), flush=True)


if world == 1:
    for group, top, per_pt in ((1, 26, bench.MAD32_PER_G1_POINT), (2, 26, 3 * bench.MAD32_PER_G1_POINT)):
        n_top = 1 << top
        kk = bench.dev_random_fr(torch, n_top, 99, dev)
        bases = ctx.bases_generate(group, kk.data_ptr(), n_top)
        sz = 96 * group
        res = torch.zeros(192, dtype=torch.uint8, device=dev)
        for log_n in range(16, top + 1, 2):
            n = 1 << log_n
            sc = bench.dev_random_fr(torch, n, 7 + log_n, dev)
            ctx.msm_resident_async(bases, sc.data_ptr(), n, res.data_ptr())
            ms = timed(lambda: ctx.msm_resident_async(bases, sc.data_ptr(), n, res.data_ptr()))
            # linearity identity: sum_i s_i (k_i G) = (sum_i s_i k_i) G
            tot = ctx.fr_dot_device(kk.data_ptr(), sc.data_ptr(), n)
            one = torch.from_numpy(np.ascontiguousarray(tot).reshape(1, 32)).to(dev)
            b1 = ctx.bases_generate(group, one.data_ptr(), 1)
            ok = bytes(b1.download()[0]) == bytes(res.cpu().numpy()[:sz])
            b1.free()
            assert ok, "linearity identity violated"
            ach = n * per_pt / (ms * 1e-3)
            emit({"op": "msm_g%d" % group, "n_gpus": 1, "log_n": log_n, "ms": round(ms, 4), "Mpts_per_s": round(n / ms / 1e3, 2), "parity_checked": True,
                  "roofline_frac_int32": round(ach / mad_peak, 4)})
            del sc
        bases.free()
        del kk
        torch.cuda.empty_cache()
    for log_n in range(16, 28):
        x = bench.dev_random_fr(torch, 1 << log_n, log_n, dev)
        ctx.ntt_device(x.data_ptr(), log_n)
        ms = timed(lambda: ctx.ntt_device(x.data_ptr(), log_n))
        gbps = 2 * 32 * (1 << log_n) / (ms * 1e-3) / 1e9
        emit({"op": "ntt_fr", "n_gpus": 1, "log_n": log_n, "ms": round(ms, 4), "alg_GBps": round(gbps, 1), "roofline_frac_hbm": round(gbps / hbm_peak, 4),
              "roofline_frac_int32": round((1 << log_n) * log_n / 2 * 136.0 / (ms * 1e-3) / mad_peak, 4)})
        del x
else:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize(); dist.barrier()

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for log_n in range(16, 27, 2):
        args = types.SimpleNamespace(log_msm=log_n, log_ntt=log_n)
        d = bench.msm_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, mad_peak)
        emit({"op": "msm_g1_sharded", "n_gpus": world, "log_n": log_n, "ms": round(d["ms"], 4), "Mpts_per_s": round(d["Mpts_per_s"], 2), "parity_checked": True,
              "roofline_frac_int32": round(d["roofline"]["frac"], 4)})
        torch.cuda.empty_cache()
    for log_n in range(17, 28, 2):
        args = types.SimpleNamespace(log_msm=log_n, log_ntt=log_n)
        d = bench.ntt_sharded_bench(ctx, torch, dist, dev, args, rank, world, barrier, max_over_ranks, hbm_peak)
        emit({"op": "ntt_fr_sharded", "n_gpus": world, "log_n": log_n, "natural_ms": round(d["natural"]["ms"], 4), "cyclic_ms": round(d["cyclic"]["ms"], 4),
              "cyclic_p2p_fused_ms": round(d["cyclic_p2p_fused"]["ms"], 4) if "ms" in d["cyclic_p2p_fused"] else None, "round_trip_checked": True})
        torch.cuda.empty_cache()
    dist.barrier()
    dist.destroy_process_group()
ctx.close()
