"""Step-by-step device times of one multiexp (B2G16_MSM_TRACE=1: CUDA events after every step of the accumulation and
reduction), aggregated by step.  Usage: B2G16_MSM_TRACE=1 python tools/msm_trace.py [log_n=22] [group=1] 2> trace.txt"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from b2g16 import lib
from probe import dev_random_fr, timed

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
group = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
n = 1 << log_n
k, s = dev_random_fr(n, 99), dev_random_fr(n, 7)
bases = ctx.bases_generate(group, k.data_ptr(), n)
out = torch.zeros(192, dtype=torch.uint8, device="cuda")
t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr()), reps=3)
print("msm group", group, "2^%d" % log_n, round(t, 3), "ms", flush=True)
ctx.set_msm_timing(True)
sys.stderr.write("TRACE_BEGIN\n"); sys.stderr.flush()
ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr()); torch.cuda.synchronize()
print("stages", {a: round(b, 3) for a, b in ctx.last_msm_timing().items()}, flush=True)
ctx.close()
