"""Cost of a parameter-cache miss (F/caches.hpp:48-67 lookup_groth_params -> S/core/parameter_cache.hpp:125-129
read_cached_params): a `.params` file of the 2 KiB-shaped partition (domain 2^22) written from a resident key,
read back (bellman-style density-filtered a / b queries expanded) and loaded onto the device; the reloaded key
must export the same queries.  One JSON line.    python tools/key_cache_time.py [log_m] [dir]"""
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "crypto3-fil-proofs_b200", "python")]
import numpy as np

from b2g16 import codec, lib


def main():
    log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
    d = sys.argv[2] if len(sys.argv) > 2 else "/tmp"
    n_in, n_free = 39, 64
    ctx = lib.Context(0)
    toxic = codec.random_fr(5, np.random.Generator(np.random.PCG64(9)))
    t = time.time()
    pk, vk = ctx.pk_generate_synth(log_m, n_in, n_free, 1, toxic, want_vk=True)
    t_gen = time.time() - t
    shp = pk.shape()
    nv, m = shp["n_vars"], 1 << log_m
    params = {"gamma_g2": vk[288:480], "ic": vk[672:].reshape(-1, 96)}
    t = time.time()
    for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", shp["n_aux"]), (4, "h", m - 1)):
        params[n] = pk.export(which, cnt)
    for which, n in ((5, "alpha_g1"), (6, "beta_g1"), (7, "delta_g1"), (8, "beta_g2"), (9, "delta_g2")):
        params[n] = pk.export(which, 1)[0]
    t_export = time.time() - t
    digest = {n: hashlib.sha256(params[n].tobytes()).hexdigest() for n in ("a", "b_g1", "b_g2", "l", "h")}
    pk.free()
    dens = {}
    t = time.time()
    for n, w in (("a", 96), ("b_g1", 96), ("b_g2", 192)):
        params[n], dens[n] = lib.query_filter(params[n], w)
    t_filter = time.time() - t
    path = os.path.join(d, "v28-key-cache-time.params")
    t = time.time()
    lib.params_write(path, params)
    t_write = time.time() - t
    size = os.path.getsize(path)
    del params
    t = time.time()
    back = lib.params_read(path)
    t_read = time.time() - t
    t = time.time()
    for n, w in (("a", 96), ("b_g1", 96), ("b_g2", 192)):
        back[n] = lib.query_expand(back[n], dens[n], w)
    t_expand = time.time() - t

    class Key:
        pass
    key = Key()
    for n in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h"):
        setattr(key, n, back[n])

    class ShapeOnly:
        n_cons, n_aux, mats = shp["n_cons"], shp["n_aux"], None
    ShapeOnly.log_m, ShapeOnly.n_in = log_m, n_in
    t = time.time()
    pk2 = ctx.pk_load(key, ShapeOnly)
    t_load = time.time() - t
    same = all(hashlib.sha256(pk2.export(which, cnt).tobytes()).hexdigest() == digest[n]
               for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", shp["n_aux"]), (4, "h", m - 1)))
    os.remove(path)
    pk2.free()
    # ---- the streamed loader (b2g16_pk_load_file): dense file -> pinned ring -> device conversion, no host copy of the key
    for n, w in (("a", 96), ("b_g1", 96), ("b_g2", 192)):
        pass                                                    # back[n] is dense again (expanded above)
    dense_path = os.path.join(d, "v28-key-cache-time-dense.params")
    back["gamma_g2"], back["ic"] = vk[288:480], vk[672:].reshape(-1, 96)
    lib.params_write(dense_path, back)
    dense_size = os.path.getsize(dense_path)
    del back, key
    streamed = {}
    for chunk in (8, 32, 128):
        ctx.set_tuning("key_chunk_mib", chunk)
        t = time.time()
        pk3 = ctx.pk_load_file(dense_path, ShapeOnly)
        dt = time.time() - t
        streamed["chunk_%d_MiB" % chunk] = {"s": round(dt, 3), "GB_per_s": round(dense_size / dt / 1e9, 2)}
        same3 = all(hashlib.sha256(pk3.export(which, cnt).tobytes()).hexdigest() == digest[n]
                    for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", shp["n_aux"]), (4, "h", m - 1)))
        assert same3, "streamed key differs"
        pk3.free()
    ctx.set_tuning("key_chunk_mib", 32)
    os.remove(dense_path)
    pk2 = None
    print(json.dumps({"what": "parameter-cache miss, domain 2^%d" % log_m, "file_GB": round(size / 1e9, 3),
                      "streamed_load": {"what": "b2g16_pk_load_file on the dense file (page-cache warm: written just before)", "file_GB": round(dense_size / 1e9, 3), **streamed, "key_identical": True},
                      "density": {n: round(float(v.mean()), 3) for n, v in dens.items()},
                      "s": {"generate_on_device": round(t_gen, 3), "export_to_host": round(t_export, 3), "filter": round(t_filter, 3),
                            "params_write": round(t_write, 3), "params_read": round(t_read, 3), "expand": round(t_expand, 3),
                            "pk_load": round(t_load, 3)},
                      "reloaded_key_identical": bool(same)}))
    assert same
    ctx.close()


if __name__ == "__main__":
    main()
