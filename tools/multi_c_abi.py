"""Several GPUs from ONE process through the C ABI (b2g16_init_multi, csrc/multi.cu): times
  * a batch of P partitions (P = devices, and P = devices + 2 so that the last wave runs in latency mode),
  * one proof spread over all devices (b2g16_prove_multi),
against the same proofs on one device, and checks that the bytes are identical.
Usage: python tools/multi_c_abi.py [log_m=22] [n_in=39]      (uses every visible device)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
import numpy as np
import torch
import bench
from b2g16 import codec, lib

log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
n_in = int(sys.argv[2]) if len(sys.argv) > 2 else 39
n_dev = torch.cuda.device_count()
n_free, seed = bench.BENCH_N_FREE, bench.BENCH_SEED
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
ctx.prove(pk, z, r, sr)
t0 = time.perf_counter(); want = ctx.prove(pk, z, r, sr); t_one = time.perf_counter() - t0
pk.free(); ctx.close()

m = lib.MultiContext(list(range(n_dev)))
k = m.key_generate_synth(log_m, n_in, n_free, seed, toxic)
out = {"devices": n_dev, "log_m": log_m, "n_in": n_in, "one_device_proof_s": round(t_one, 4), "host_buffers": "one-device proof: pageable numpy array; multi-device calls: page-locked"}
for P in (n_dev, n_dev + 2 if n_dev > 2 else n_dev + 1, 4 * n_dev):
    zs_t = torch.empty((P,) + z.shape, dtype=torch.uint8, pin_memory=True)      # page-locked host witnesses: the copies run at PCIe speed
    zs = zs_t.numpy()
    zs[:] = z
    rs, ss = np.stack([r] * P), np.stack([sr] * P)
    m.prove_batch(k, zs, rs, ss)                                    # warm-up
    times = []
    for _ in range(4):
        t0 = time.perf_counter(); got = m.prove_batch(k, zs, rs, ss); times.append(time.perf_counter() - t0)
        assert got == [want] * P, "multi-device batch differs from the single-device proof"
    dt = min(times)
    out["batch_%d_partitions_s" % P] = round(dt, 4)
    out["batch_%d_partitions_all_s" % P] = [round(t, 4) for t in times]
    out["batch_%d_partitions_per_s" % P] = round(P / dt, 3)
z_t = torch.empty(z.shape, dtype=torch.uint8, pin_memory=True); zp = z_t.numpy(); zp[:] = z
m.prove(k, zp, r, sr)
t0 = time.perf_counter(); got = m.prove(k, zp, r, sr); dt = time.perf_counter() - t0
assert got == want, "latency-mode proof differs"
out["one_proof_over_all_devices_s"] = round(dt, 4)
out["parity"] = "every proof byte-identical to b2g16_prove on one device"
print(json.dumps(out))
m.key_free(k); m.close()
