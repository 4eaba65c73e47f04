"""Generates tests/golden/*.npz from the slow Python oracle (oracle/pyref).

Run from the repo root:  python tests/golden/make_golden.py
The reference holds no golden vectors for this path, so these are
produced by the obviously-correct big-int oracle: affine chord-and-tangent group
law, O(n^2) DFT cross-check, pairing-verified Groth16 proofs.  The reference KATs
that do exist (Fr canonical encodings, libs/storage/test/core/fr32.cpp:51-85 and
libs/filecoin/test/api/mod.cpp:35-43) are embedded verbatim as `fr32_kat`.
"""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "crypto3-fil-proofs_b200", "python"))
from oracle.pyref import bls12_381 as c, groth16 as g   # noqa: E402
from b2g16 import codec, synth                            # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def fields():
    rng = random.Random(0xF1E1D)
    d = {}
    for name, mod, size in (("fr", c.R, 32), ("fq", c.P, 48)):
        edge = [0, 1, 2, mod - 1, mod - 2, (1 << (8 * size - 8 * (size % 16 == 0))) % mod,
                pow(2, 32 * (size // 4), mod), (mod - 1) // 2, (mod + 1) // 2]
        a = edge + [rng.randrange(mod) for _ in range(64)]
        b = list(reversed(edge)) + [rng.randrange(mod) for _ in range(64)]
        enc = lambda v: np.frombuffer(b"".join(int(x).to_bytes(size, "little") for x in v), dtype=np.uint8).reshape(-1, size)
        d[name + "_a"], d[name + "_b"] = enc(a), enc(b)
        d[name + "_add"] = enc([(x + y) % mod for x, y in zip(a, b)])
        d[name + "_sub"] = enc([(x - y) % mod for x, y in zip(a, b)])
        d[name + "_mul"] = enc([(x * y) % mod for x, y in zip(a, b)])
        d[name + "_sqr"] = enc([(x * x) % mod for x in a])
        d[name + "_neg"] = enc([(-x) % mod for x in a])
        d[name + "_inv"] = enc([pow(x, mod - 2, mod) for x in a])
    # reference KATs: (bytes, is_canonical)
    kat = [(bytes(range(32)), 1),
           (b"\xff" * 31 + b"\x73", 0), (b"\xff" * 31 + b"\x72", 1),
           (b"\xff" * 30 + b"\xec\x73", 1), (b"\xff" * 30 + b"\xed\x73", 0),
           (b"\x00" * 32, 1), (b"\xff" * 32, 0)]
    d["fr32_kat_bytes"] = np.frombuffer(b"".join(k for k, _ in kat), dtype=np.uint8).reshape(-1, 32)
    d["fr32_kat_ok"] = np.array([v for _, v in kat], dtype=np.uint8)
    for (k, ok) in kat:
        assert (c.fr_from_bytes_le(k) is not None) == bool(ok)
    np.savez_compressed(os.path.join(OUT, "fields.npz"), **d)


def ntts():
    rng = random.Random(0x17717)
    d = {}
    for ln in range(0, 11):
        v = [rng.randrange(c.R) for _ in range(1 << ln)]
        if ln <= 6:
            assert g.ntt(v) == g.dft_naive(v) and g.ntt(v, True) == g.dft_naive(v, True)
        d[f"in_{ln}"] = codec.fr_from_ints(v)
        for inv in (0, 1):
            for cos in (0, 1):
                d[f"out_{ln}_{inv}{cos}"] = codec.fr_from_ints(g.transform(v, bool(inv), bool(cos)))
    np.savez_compressed(os.path.join(OUT, "ntt.npz"), **d)


def msms():
    rng = random.Random(0x35A)
    d = {}
    for n in (1, 2, 3, 17, 256):
        ks = [rng.randrange(1, c.R) for _ in range(n)]
        sc = [rng.randrange(c.R) for _ in range(n)]
        # edge scalars: 0, 1, r-1, window boundaries for several c
        edge = [0, 1, c.R - 1, 2, (1 << 15), (1 << 16) - 1, (1 << 16), (1 << 20) - 1, (1 << 21),
                (1 << 254) % c.R, c.R - 2, (1 << 13) - 1, 1 << 12, 0xFFFF0000FFFF, 1 << 128]
        for i, e in enumerate(edge[: max(0, n - 2)]):
            sc[i] = e
        g1 = [c.g1_gen_mul(k) for k in ks]
        g2 = [c.g2_gen_mul(k) for k in ks]
        if n >= 17:
            g1[3], g2[3] = None, None                      # identity base
            g1[5], g2[5] = g1[4], g2[4]                    # duplicate
            g1[7], g2[7] = c.g1_neg(g1[6]), c.g2_neg(g2[6])  # negated neighbour
            sc[7] = sc[6]                                  # so they cancel exactly
            sc[5] = sc[4]
        d[f"g1_bases_{n}"], d[f"g2_bases_{n}"] = codec.g1_from_points(g1), codec.g2_from_points(g2)
        d[f"scalars_{n}"] = codec.fr_from_ints(sc)
        d[f"g1_out_{n}"] = codec.g1_from_points([c.msm_g1(g1, sc)])[0]
        d[f"g2_out_{n}"] = codec.g2_from_points([c.msm_g2(g2, sc)])[0]
        print("msm", n, flush=True)
    # all-zero scalars -> identity; bases all identity -> identity
    np.savez_compressed(os.path.join(OUT, "msm.npz"), **d)


def to_py_r1cs(s):
    def m(M):
        rp, col, val = M
        return ([int(x) for x in rp], [int(x) for x in col], codec.fr_to_ints(val))
    return g.R1CS(s.n_in, s.n_aux, m(s.A), m(s.B), m(s.C))


def py_witness(s, cs):
    z = codec.fr_to_ints(s.z_head)
    base = 1 + s.n_in + s.n_free
    for i in range(s.n_cons):
        sa = sum(cs.A[2][k] * z[cs.A[1][k]] for k in range(cs.A[0][i], cs.A[0][i + 1]))
        sb = sum(cs.B[2][k] * z[cs.B[1][k]] for k in range(cs.B[0][i], cs.B[0][i + 1]))
        assert len(z) == base + i
        z.append(sa * sb % c.R)
    return z


def groth(log_m, n_in, seed):
    s = synth.SynthR1cs(log_m, n_in, seed, n_free=3)
    cs = to_py_r1cs(s)
    z = py_witness(s, cs)
    assert cs.is_satisfied(z) and cs.domain_log() == log_m
    rng = random.Random(seed ^ 0xABCDEF)
    toxic = [rng.randrange(1, c.R) for _ in range(5)]
    r, sr = rng.randrange(c.R), rng.randrange(c.R)
    pk = g.setup(cs, *toxic)
    tr = g.prove(pk, z, r, sr)
    assert g.verify(pk, z[1:1 + cs.n_in], tr.proof)
    bad = list(z[1:1 + cs.n_in]); bad[0] = (bad[0] + 1) % c.R
    assert not g.verify(pk, bad, tr.proof)
    d = dict(log_m=log_m, n_in=n_in, seed=seed, n_free=3,
             toxic=codec.fr_from_ints(toxic), r=codec.fr_from_ints([r])[0], s=codec.fr_from_ints([sr])[0],
             z=codec.fr_from_ints(z),
             alpha_g1=codec.g1_from_points([pk.alpha_g1])[0], beta_g1=codec.g1_from_points([pk.beta_g1])[0],
             delta_g1=codec.g1_from_points([pk.delta_g1])[0], beta_g2=codec.g2_from_points([pk.beta_g2])[0],
             gamma_g2=codec.g2_from_points([pk.gamma_g2])[0], delta_g2=codec.g2_from_points([pk.delta_g2])[0],
             ic=codec.g1_from_points(pk.ic), a=codec.g1_from_points(pk.a), b_g1=codec.g1_from_points(pk.b_g1),
             b_g2=codec.g2_from_points(pk.b_g2), l=codec.g1_from_points(pk.l), h=codec.g1_from_points(pk.h),
             ev_a=codec.fr_from_ints(tr.a), ev_b=codec.fr_from_ints(tr.b), ev_c=codec.fr_from_ints(tr.c),
             h_coeffs=codec.fr_from_ints(tr.h),
             msm_a=codec.g1_from_points([tr.msm_a])[0], msm_b1=codec.g1_from_points([tr.msm_b1])[0],
             msm_l=codec.g1_from_points([tr.msm_l])[0], msm_h=codec.g1_from_points([tr.msm_h])[0],
             msm_b2=codec.g2_from_points([tr.msm_b2])[0],
             A=codec.g1_from_points([tr.A])[0], B=codec.g2_from_points([tr.B])[0], C=codec.g1_from_points([tr.C])[0],
             proof=np.frombuffer(tr.proof, dtype=np.uint8))
    np.savez_compressed(os.path.join(OUT, f"groth16_m{log_m}.npz"), **d)
    print("groth16", log_m, "verified", flush=True)


if __name__ == "__main__":
    assert c.check_generators_order()
    fields()
    ntts()
    msms()
    groth(4, 2, 41)
    groth(7, 5, 42)
