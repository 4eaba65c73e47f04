"""One synthetic Groth16 instance (domain 2^log_m), device setup, then `reps` device-resident proofs --
the command ncu wraps for the per-kernel launch list of a proof.  Usage: python tools/prof_prove.py [log_m] [reps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import numpy as np
import torch
from b2g16 import lib, synth

log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
s = synth.SynthR1cs(log_m, 39, 2026, n_free=8)
toxic, r, sr = bench.toxic_from_seed(2026)
pk = ctx.pk_generate(bench._Cs(s), toxic)
dz = torch.zeros((s.n_vars, 32), dtype=torch.uint8, device="cuda")
dz[: s.z_head.shape[0]] = torch.from_numpy(s.z_head).cuda()
ctx.synth_witness(pk, dz.data_ptr(), s.n_free)
torch.cuda.synchronize()
print("PROVE_BEGIN launches", ctx.launches, flush=True)
for i in range(reps):
    t0 = time.perf_counter()
    l0 = ctx.launches
    proof = ctx.prove_device(pk, dz.data_ptr(), r, sr)
    print("prove", i, round((time.perf_counter() - t0) * 1e3, 2), "ms", "launches", ctx.launches - l0, proof[:8].hex(), flush=True)
pk.free()
ctx.close()
