"""Device-generated PoRep-shaped instance at domain 2^log_m: setup, witness, timed proofs, pairing check.
Usage: python tools/large_proof.py log_m [n_in] [reps]   (BASELINE configs 4/5: log_m = 27, n_in = 327)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch
from b2g16 import lib, synth
import util
from oracle import coracle

log_m = int(sys.argv[1])
n_in = int(sys.argv[2]) if len(sys.argv) > 2 else 327
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
n_free, seed = 8, 2026
ts = torch.cuda.Stream()
torch.cuda.set_stream(ts)
ctx = lib.Context(0, ts.cuda_stream)
toxic, r, sr = util.toxic_from_seed(seed)
t0 = time.perf_counter()
pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic, want_vk=True)
torch.cuda.synchronize()
t_setup = time.perf_counter() - t0
shape = pk.shape()
head = synth._hashed_fr(synth._stream(seed, 5), 1 + n_in + n_free).copy()
head[0] = 0; head[0, 0] = 1
dz = torch.zeros((shape["n_vars"], 32), dtype=torch.uint8, device="cuda")
dz[: head.shape[0]] = torch.from_numpy(head).cuda()
t0 = time.perf_counter()
ctx.synth_witness(pk, dz.data_ptr(), n_free)
torch.cuda.synchronize()
t_wit = time.perf_counter() - t0
times = []
for i in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    proof = ctx.prove_device(pk, dz.data_ptr(), r, sr)
    e1.record(); torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1) / 1e3)
free_b, total_b = torch.cuda.mem_get_info()
coracle.build()
t0 = time.perf_counter()
ok = util.pairing_verify(coracle, vk, head[1:1 + n_in], proof)
print(json.dumps({"log_m": log_m, "shape": shape, "setup_s": round(t_setup, 2), "witness_s": round(t_wit, 2), "prove_s": [round(t, 3) for t in times],
                  "pairing_verify": bool(ok), "verify_s": round(time.perf_counter() - t0, 1), "hbm_used_GB": round((total_b - free_b) / 1e9, 1),
                  "proof": proof.hex()[:32] + "..."}), flush=True)
pk.free()
ctx.close()
