"""Quick device timings (not the bench contract): MAD32 peak, NTT sweep, MSM sweep with
per-stage times.  Usage: python tools/probe.py [max_log_ntt] [max_log_msm]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
import numpy as np
import torch
from b2g16 import codec, lib


def dev_random_fr(n, seed):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    x = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device="cuda", generator=g)
    x[:, 31] &= 0x3F                      # < 2^254 < r : canonical, near-uniform
    return x


def timed(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    max_ntt = int(sys.argv[1]) if len(sys.argv) > 1 else 26
    max_msm = int(sys.argv[2]) if len(sys.argv) > 2 else 24
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)             # all torch work and the library share one stream
    ctx = lib.Context(0, ts.cuda_stream)
    peak, ms = ctx.mad32_peak()
    print(json.dumps({"mad32_peak_per_s": peak, "ms": ms}), flush=True)
    for log_n in range(16, max_ntt + 1, 2 if max_ntt > 20 else 1):
        x = dev_random_fr(1 << log_n, log_n)
        ctx.ntt_device(x.data_ptr(), log_n)                 # warm-up + tables
        t = timed(lambda: ctx.ntt_device(x.data_ptr(), log_n))
        gbs = 2 * 32 * (1 << log_n) / (t * 1e-3) / 1e9
        print(json.dumps({"ntt_log_n": log_n, "ms": round(t, 4), "alg_GBps": round(gbs, 1)}), flush=True)
        del x
    for group in (1, 2):
        top = max_msm if group == 1 else max_msm - 2
        k = dev_random_fr(1 << top, 99)
        t0 = time.time()
        bases = ctx.bases_generate(group, k.data_ptr(), 1 << top)
        torch.cuda.synchronize()
        print(json.dumps({"gen_bases_group": group, "log_n": top, "s": round(time.time() - t0, 2)}), flush=True)
        for log_n in range(16, top + 1, 2):
            s = dev_random_fr(1 << log_n, 7 + log_n)
            out = torch.zeros(192, dtype=torch.uint8, device="cuda")
            for c in ([0] if log_n < 20 else [0, 14, 18, 20]):
                ctx.set_msm_window(c)
                ctx.set_msm_timing(False)
                ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
                t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr()), reps=2)
                ctx.set_msm_timing(True)
                ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
                st = ctx.last_msm_timing()
                print(json.dumps({"msm_group": group, "log_n": log_n, "c": c, "ms": round(t, 3), "Mpts_s": round((1 << log_n) / t / 1e3, 2),
                                  "stages_ms": {k2: round(v, 3) for k2, v in st.items()}}), flush=True)
            ctx.set_msm_window(0); ctx.set_msm_timing(False)
        bases.free()
    ctx.close()


if __name__ == "__main__":
    main()
