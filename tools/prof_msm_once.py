"""Exactly one multiexp (no warm-up) on generated bases -- the command `ncu --set full -k regex:...` wraps so that the
capture holds every accumulate-stage kernel of ONE multiexp.  Usage: python tools/prof_msm_once.py [log_n=22] [group=1]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from b2g16 import lib
from probe import dev_random_fr

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 22
group = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
n = 1 << log_n
k, s = dev_random_fr(n, 99), dev_random_fr(n, 7)
bases = ctx.bases_generate(group, k.data_ptr(), n)
out = torch.zeros(192, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr())
torch.cuda.synchronize()
print("one msm done, group", group, "2^%d" % log_n, bytes(out.cpu().numpy()[:8]).hex())
ctx.close()
