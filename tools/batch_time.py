"""b2g16_prove_batch with pageable host witnesses (what a C++ caller holds) against the same proofs one b2g16_prove at a
time: the batch uploads partition k + 1 while partition k is proved.  Usage: python tools/batch_time.py [log_m=22] [P=4]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from b2g16 import lib

log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
P = int(sys.argv[2]) if len(sys.argv) > 2 else 4
n_in, n_free, seed = 39, bench.BENCH_N_FREE, bench.BENCH_SEED
toxic, r, sr = bench.toxic_from_seed(seed)
ctx = lib.Context(0)
pk = ctx.pk_generate_synth(log_m, n_in, n_free, seed, toxic)
nv = pk.shape()["n_vars"]
dz = torch.zeros((nv, 32), dtype=torch.uint8, device="cuda:0")
head = bench.witness_head(n_in, 0)
dz[: head.shape[0]] = torch.from_numpy(head).cuda()
ctx.synth_witness(pk, dz.data_ptr(), n_free)
z = dz.cpu().numpy()
del dz
zs = np.ascontiguousarray(np.broadcast_to(z, (P,) + z.shape))          # pageable
rs, ss = np.stack([r] * P), np.stack([sr] * P)
want = ctx.prove(pk, z, r, sr)
ctx.prove_batch(pk, zs, rs, ss)
t0 = time.perf_counter(); one = [ctx.prove(pk, zs[k], r, sr) for k in range(P)]; t_one = time.perf_counter() - t0
t0 = time.perf_counter(); got = ctx.prove_batch(pk, zs, rs, ss); t_batch = time.perf_counter() - t0
assert got == want * P and one == [want] * P
print(json.dumps({"log_m": log_m, "partitions": P, "host_buffers": "pageable", "one_call_per_partition_s": round(t_one, 4), "prove_batch_s": round(t_batch, 4),
                  "per_partition_s": [round(t_one / P, 4), round(t_batch / P, 4)], "proofs_identical": True}))
pk.free(); ctx.close()
