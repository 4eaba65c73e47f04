"""The cross-rank step of the sharded transform (local form) on one GPU -- the command ncu wraps.
Usage: python tools/prof_cross.py [log_n] [log_g]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from b2g16 import lib
from probe import dev_random_fr, timed

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 27
log_g = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
M = (1 << log_n) >> log_g
x = dev_random_fr(M, 3)
y = torch.empty_like(x)
t = timed(lambda: ctx.ntt_cross_device(x.data_ptr(), y.data_ptr(), log_n, log_g, 1), reps=3)
print("ntt_cross 2^%d / %d ranks: %d elements per rank, %.3f ms, %.1f GB/s" % (log_n, 1 << log_g, M, t, 2 * M * 32 / t / 1e6))
ctx.close()
