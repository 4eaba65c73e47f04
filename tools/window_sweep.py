"""Window-width sweep of the G1 / G2 multiexp on device-resident data: for every size, time the MSM at
each window width c and print the per-stage split.  Feeds msm_pick_window (csrc/msm_impl.cuh).
Usage: python tools/window_sweep.py [group] [min_log] [max_log]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from b2g16 import lib
from probe import dev_random_fr, timed

group = int(sys.argv[1]) if len(sys.argv) > 1 else 1
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 18
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 26
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
k = dev_random_fr(1 << hi, 99)
bases = ctx.bases_generate(group, k.data_ptr(), 1 << hi)
del k
out = torch.zeros(192, dtype=torch.uint8, device="cuda")
for log_n in range(lo, hi + 1):
    s = dev_random_fr(1 << log_n, 7 + log_n)
    best = None
    for c in range(max(10, log_n - 8), min(23, log_n - 1) + 1):
        if (1 << log_n) * ((255 + c) // c + 1) >= 4.0e9:
            continue
        ctx.set_msm_window(c)
        ctx.set_msm_timing(False)
        try:
            ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
            t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr()), reps=2)
        except lib.B2g16Error as e:
            print(json.dumps({"group": group, "log_n": log_n, "c": c, "error": str(e)}), flush=True)
            continue
        ctx.set_msm_timing(True)
        ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr())
        st = ctx.last_msm_timing()
        print(json.dumps({"group": group, "log_n": log_n, "c": c, "ms": round(t, 3), "stages_ms": {k2: round(v, 2) for k2, v in st.items()}}), flush=True)
        if best is None or t < best[1]:
            best = (c, t)
    ctx.set_msm_window(0); ctx.set_msm_timing(False)
    t0 = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_n, out.data_ptr()), reps=2)
    print(json.dumps({"group": group, "log_n": log_n, "best_c": best[0], "best_ms": round(best[1], 3), "default_ms": round(t0, 3)}), flush=True)
    del s
bases.free()
ctx.close()
