"""One MSM (G1, 2^log_msm) and one NTT (2^log_ntt) on device-resident data -- the command ncu wraps.
Usage: python tools/prof_one.py [log_msm] [log_ntt] [group]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from b2g16 import lib
from probe import dev_random_fr, timed

log_msm = int(sys.argv[1]) if len(sys.argv) > 1 else 22
log_ntt = int(sys.argv[2]) if len(sys.argv) > 2 else 24
group = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
if log_msm:
    k, s = dev_random_fr(1 << log_msm, 99), dev_random_fr(1 << log_msm, 7)
    bases = ctx.bases_generate(group, k.data_ptr(), 1 << log_msm)
    out = torch.zeros(192, dtype=torch.uint8, device="cuda")
    t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), 1 << log_msm, out.data_ptr()), reps=2)
    print("msm group", group, "2^%d" % log_msm, round(t, 3), "ms")
if log_ntt:
    x = dev_random_fr(1 << log_ntt, 5)
    t = timed(lambda: ctx.ntt_device(x.data_ptr(), log_ntt), reps=2)
    print("ntt 2^%d" % log_ntt, round(t, 3), "ms")
ctx.close()
