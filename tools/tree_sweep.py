"""MSM time and stage split against the number of batched-affine tree levels / outputs per thread / window width.
Usage: python tools/tree_sweep.py [log_n=22] [group=1]   -> one JSON line per configuration"""
import json
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
cs = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0]
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
n = 1 << log_n
k, s = dev_random_fr(n, 99), dev_random_fr(n, 7)
bases = ctx.bases_generate(group, k.data_ptr(), n)
del k
out = torch.zeros(192, dtype=torch.uint8, device="cuda")
ref = None
for c in cs:
    ctx.set_msm_window(c)
    for levels, K in [(0, 16), (2, 16), (3, 16), (4, 16), (5, 16), (6, 16), (7, 16), (-1, 8), (-1, 32)]:
        ctx.set_tuning("tree_levels", levels); ctx.set_tuning("tree_k", K)
        ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr())
        torch.cuda.synchronize()
        res = bytes(out.cpu().numpy()[:96 * group])
        assert ref is None or res == ref, "result changed with the tree configuration"
        ref = res
        t = timed(lambda: ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr()), reps=3)
        ctx.set_msm_timing(True)
        ctx.msm_resident_async(bases, s.data_ptr(), n, out.data_ptr()); torch.cuda.synchronize()
        st = ctx.last_msm_timing()
        ctx.set_msm_timing(False)
        print(json.dumps({"log_n": log_n, "group": group, "c": c, "tree_levels": levels, "tree_k": K, "ms": round(t, 3), "stages": {a: round(b, 3) for a, b in st.items()}}), flush=True)
ctx.close()
