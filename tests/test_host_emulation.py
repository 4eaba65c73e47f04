"""CPU check of the device arithmetic headers (csrc/fp.cuh, csrc/ec.cuh): they are
compiled for the host with the PTX carry-chain primitives emulated and compared
bit-for-bit with the golden vectors / oracle.  Catches limb-algorithm bugs without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from b2g16 import codec
import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host_emu", "emu.cpp")
CSRC = os.path.join(ROOT, "crypto3-fil-proofs_b200", "csrc")
SO = os.path.join(ROOT, "tests", "host_emu", "_build", "libemu.so")


@pytest.fixture(scope="module")
def emu():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    deps = [SRC] + [os.path.join(CSRC, f) for f in ("fp.cuh", "ec.cuh")]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-x", "c++", "-I", CSRC, SRC, "-o", SO])
    return C.CDLL(SO)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


OPS = {"add": 0, "sub": 1, "mul": 2, "inv": 3, "sqr": 4, "neg": 5, "inv_gcd": 6}


def _op(emu, field, op, a, b=None):
    a = np.ascontiguousarray(a)
    b = a if b is None else np.ascontiguousarray(b)
    out = np.empty_like(a)
    fn = emu.emu_fr_op if field == "fr" else emu.emu_fq_op
    assert fn(OPS[op], _p(a), _p(b), C.c_size_t(a.shape[0]), _p(out)) == 0
    return out


def test_montgomery_constants():
    """fp.cuh constants vs big-int derivation from the BLS parameter."""
    import re
    x = -0xd201000000010000
    r = x**4 - x**2 + 1
    p = (x - 1) ** 2 * r // 3 + x
    txt = open(os.path.join(CSRC, "fp.cuh")).read()

    def arr(struct, name):
        body = txt.split("struct " + struct)[1]
        m = re.search(name + r"\(int i\) \{\s*constexpr uint32_t t\[\d+\] = \{([^}]*)\}", body)
        limbs = [int(t.strip().rstrip("u"), 16) for t in m.group(1).split(",")]
        return sum(v << (32 * i) for i, v in enumerate(limbs))

    def inv(struct):
        body = txt.split("struct " + struct)[1]
        return int(re.search(r"INV = (0x[0-9a-f]+)u", body).group(1), 16)
    for struct, mod, n in (("FrParams", r, 8), ("FqParams", p, 12)):
        assert arr(struct, "P") == mod
        assert arr(struct, "ONE") == pow(2, 32 * n, mod)
        assert arr(struct, "R2") == pow(2, 64 * n, mod)
        assert inv(struct) == (-pow(mod, -1, 2**32)) % 2**32


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_field_golden(emu, field):
    g = util.golden("fields")
    a, b = g[field + "_a"], g[field + "_b"]
    for op in ("add", "sub", "mul"):
        assert np.array_equal(_op(emu, field, op, a, b), g[f"{field}_{op}"]), op
    for op in ("sqr", "neg", "inv"):
        assert np.array_equal(_op(emu, field, op, a), g[f"{field}_{op}"]), op


@pytest.mark.parametrize("field,size", [("fr", 32), ("fq", 48)])
def test_field_random_vs_oracle(emu, oracle, field, size):
    rng = np.random.Generator(np.random.PCG64(2024))
    n = 20000
    if field == "fr":
        a, b = codec.random_fr(n, rng), codec.random_fr(n, rng)
    else:
        # random Fq: 380-bit values are always < p
        a = rng.integers(0, 256, size=(n, 48), dtype=np.uint8); a[:, 47] &= 0x0F
        b = rng.integers(0, 256, size=(n, 48), dtype=np.uint8); b[:, 47] &= 0x0F
    # limbs with extreme values to stress carries
    a[:50, :8] = 0xFF; b[:50, 4:12] = 0xFF; a[50:100, 8:16] = 0; b[100:150, :4] = 0
    for op in ("add", "sub", "mul"):
        assert np.array_equal(_op(emu, field, op, a, b), oracle.field_op(field, op, a, b)), op
    assert np.array_equal(_op(emu, field, "sqr", a), oracle.field_op(field, "sqr", a))


def test_wide_products_vs_oracle(emu, oracle):
    """fp_wide.cuh: the split product + separate Montgomery reduction, its Karatsuba form and the lazy-reduction Fq2
    product (the one Fq2::operator* uses) against big-integer arithmetic."""
    rng = np.random.Generator(np.random.PCG64(31))
    n = 6000
    P = 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
    a = rng.integers(0, 256, size=(2 * n, 48), dtype=np.uint8); a[:, 47] &= 0x0F
    b = rng.integers(0, 256, size=(2 * n, 48), dtype=np.uint8); b[:, 47] &= 0x0F
    edge = [0, 1, P - 1, P - 2, (1 << 192) - 1, ((1 << 380) - 1) & ~((1 << 192) - 1), (1 << 192), P >> 1]
    for i, v in enumerate(edge):
        a[i] = np.frombuffer(v.to_bytes(48, "little"), dtype=np.uint8)
        b[len(edge) - 1 - i] = np.frombuffer(v.to_bytes(48, "little"), dtype=np.uint8)
        a[2 * i + 20] = a[i]; b[2 * i + 21] = b[i]
    ai = [int.from_bytes(r.tobytes(), "little") for r in a]
    bi = [int.from_bytes(r.tobytes(), "little") for r in b]

    def enc(vals):
        return np.frombuffer(b"".join(v.to_bytes(48, "little") for v in vals), dtype=np.uint8).reshape(-1, 48)

    want = enc([x * y % P for x, y in zip(ai, bi)])
    for which in (0, 1):
        out = np.zeros_like(a)
        assert emu.emu_wide_mul(which, _p(a), _p(b), C.c_size_t(2 * n), _p(out)) == 0
        assert np.array_equal(out, want), which
    out = np.zeros_like(a)                                                  # which = 4: squaring through the 6-limb halves
    assert emu.emu_wide_mul(4, _p(a), _p(b), C.c_size_t(2 * n), _p(out)) == 0
    assert np.array_equal(out, enc([x * x % P for x in ai]))
    w2 = []
    for k in range(n):
        a0, a1, b0, b1 = ai[2 * k], ai[2 * k + 1], bi[2 * k], bi[2 * k + 1]
        w2 += [(a0 * b0 - a1 * b1) % P, (a0 * b1 + a1 * b0) % P]
    want2 = enc(w2)
    for which in (2, 3):
        out = np.zeros_like(a)
        assert emu.emu_wide_mul(which, _p(a), _p(b), C.c_size_t(n), _p(out)) == 0
        assert np.array_equal(out, want2), which


def test_curve_ops_vs_oracle(emu, oracle):
    rng = np.random.Generator(np.random.PCG64(77))
    k = codec.random_fr(40, rng)
    for name, gen, size, fn in (("g1", oracle.g1_gen_mul, 96, "emu_g1_sum"), ("g2", oracle.g2_gen_mul, 192, "emu_g2_sum")):
        pts = gen(k)
        pts[5] = 0                      # identity in the middle
        pts[9] = pts[8]                 # forces the doubling branch in mode 0? (acc != P) no; see below
        tot = sum(v for i, v in enumerate(codec.fr_to_ints(k)) if i not in (5, 9)) + codec.fr_to_ints(k)[8]
        want = gen(codec.fr_from_ints([tot % codec.R_MOD]))[0]
        want8 = gen(codec.fr_from_ints([tot * 8 % codec.R_MOD]))[0]
        for mode, w in ((0, want), (1, want), (2, want8)):
            out = np.zeros(size, dtype=np.uint8)
            assert getattr(emu, fn)(mode, _p(pts), C.c_size_t(40), _p(out)) == 0
            assert np.array_equal(out, w), (name, mode)
        # P + P (doubling branch of add_mixed) and P + (-P) (cancellation branch)
        two = np.stack([pts[0], pts[0]])
        out = np.zeros(size, dtype=np.uint8)
        getattr(emu, fn)(0, _p(two), C.c_size_t(2), _p(out))
        assert np.array_equal(out, gen(codec.fr_from_ints([2 * codec.fr_to_ints(k)[0] % codec.R_MOD]))[0])
        negk = codec.fr_from_ints([(-codec.fr_to_ints(k)[0]) % codec.R_MOD])
        pm = np.stack([pts[0], gen(negk)[0]])
        getattr(emu, fn)(0, _p(pm), C.c_size_t(2), _p(out))
        assert not out.any()


# ---- NTT pass kernels and MSM pipeline logic, executed on the CPU ------------------------
def _deps_all():
    return [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]


@pytest.fixture(scope="module")
def emu2():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    deps = _deps_all()
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-x", "c++", "-I", CSRC, SRC, "-o", SO])
    return C.CDLL(SO)


@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 11, 12, 13, 14])
@pytest.mark.parametrize("inverse,coset", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("tw_tables,rounds", [(1, 1), (0, 1), (1, 0)])
def test_ntt_passes_vs_oracle(emu2, oracle, log_n, inverse, coset, tw_tables, rounds):
    """tw_tables = 1: strided passes take combined butterfly twiddles from per-pass root tables (no inter-pass twiddle);
    0: the fallback that rebuilds inter-pass twiddles as tlo * thi.  rounds = 1: three stages at a time on eight elements
    in registers (ntt_phase_round, what the kernel runs); 0: one stage at a time (ntt_phase_stage, the reference form)."""
    if not tw_tables and log_n <= 11:
        pytest.skip("single contiguous pass: no strided pass to differ")
    emu2.emu_set_ntt_rounds(rounds)
    rng = np.random.Generator(np.random.PCG64(1000 + log_n))
    x = codec.random_fr(1 << log_n, rng)
    got = x.copy()
    emu2.emu_set_tw_tables(tw_tables)
    try:
        n_pass = emu2.emu_ntt(_p(got), C.c_uint(log_n), inverse, coset)
    finally:
        emu2.emu_set_tw_tables(1)
        emu2.emu_set_ntt_rounds(1)
    assert n_pass == (1 if log_n <= 11 else 2)
    want = oracle.ntt(x, inverse=bool(inverse), coset=bool(coset))
    assert np.array_equal(got, want)


def test_ntt_three_passes_vs_oracle(emu2, oracle):
    """2^21: the smallest domain with two strided passes (9 + 9 + 3 index bits), inverse coset transform."""
    log_n = 21
    x = codec.random_fr(1 << log_n, np.random.Generator(np.random.PCG64(21)))
    got = x.copy()
    assert emu2.emu_ntt(_p(got), C.c_uint(log_n), 1, 1) == 3
    assert np.array_equal(got, oracle.ntt(x, inverse=True, coset=True))


def test_ntt_golden(emu2):
    g = util.golden("ntt")
    checked = 0
    for key in g.files:
        if not key.startswith("in_"):
            continue
        log_n = int(key.split("_")[1])
        for inv in (0, 1):
            for cos in (0, 1):
                got = g[key].copy()
                emu2.emu_ntt(_p(got), C.c_uint(log_n), inv, cos)
                assert np.array_equal(got, g[f"out_{log_n}_{inv}{cos}"]), (log_n, inv, cos)
                checked += 1
    assert checked >= 8


@pytest.mark.parametrize("log_m", [1, 4, 11, 13])
def test_h_chain_vs_oracle(emu2, oracle, log_m):
    rng = np.random.Generator(np.random.PCG64(2000 + log_m))
    m = 1 << log_m
    # a*b - c must vanish on the domain for H to be a polynomial: take c = a*b pointwise
    a, b = codec.random_fr(m, rng), codec.random_fr(m, rng)
    c = oracle.field_op("fr", "mul", a, b)
    want = oracle.h_from_abc(a, b, c)
    got = np.zeros((m, 32), dtype=np.uint8)
    assert emu2.emu_h_from_abc(_p(a), _p(b), _p(c), C.c_uint(log_m), _p(got)) == 0
    assert np.array_equal(got, want)
    assert not got[m - 1].any()


@pytest.mark.parametrize("c", [8, 11, 13, 16, 20])
def test_msm_signed_digits(emu2, c):
    rng = np.random.Generator(np.random.PCG64(c))
    vals = codec.fr_to_ints(codec.random_fr(200, rng)) + [0, 1, codec.R_MOD - 1, 2**c - 1, 2**(c - 1), 2**(c - 1) + 1, 2**254]
    for v in vals:
        out = (C.c_int * 64)()
        s = np.frombuffer(int(v).to_bytes(32, "little"), dtype=np.uint8).copy()
        W = emu2.emu_msm_digits(_p(s), c, out)
        assert W == -(-256 // c)
        ds = list(out[:W])
        assert all(-(1 << (c - 1)) <= d <= (1 << (c - 1)) for d in ds)
        assert sum(d << (c * w) for w, d in enumerate(ds)) == v


@pytest.mark.parametrize("n,c,S", [(1, 8, 4), (37, 8, 4), (300, 9, 3), (300, 11, 1000)])
def test_msm_pipeline_logic_vs_oracle(emu2, oracle, n, c, S):
    rng = np.random.Generator(np.random.PCG64(n * 31 + c))
    bases = oracle.g1_gen_mul(codec.random_fr(n, rng))
    scal = codec.random_fr(n, rng)
    scal[0] = 0
    if n > 4:
        scal[1] = 0; scal[1, 0] = 1                       # scalar 1
        scal[2] = np.frombuffer((codec.R_MOD - 1).to_bytes(32, "little"), dtype=np.uint8)
        bases[3] = 0                                       # identity base
        bases[4] = bases[5]; scal[4] = scal[5]             # duplicate -> doubling branch inside a bucket
    out = np.zeros(96, dtype=np.uint8)
    assert emu2.emu_msm_g1(_p(bases), _p(scal), C.c_size_t(n), c, C.c_uint32(S), _p(out)) == 0
    assert np.array_equal(out, oracle.msm_g1(bases, scal))


@pytest.mark.parametrize("group,n,c,K,K2,levels", [(1, 1, 8, 4, 4, 2), (1, 37, 8, 3, 2, 3), (1, 300, 8, 2, 3, 4), (1, 300, 9, 16, 32, 9),
                                                   (1, 2000, 8, 5, 7, 3), (2, 120, 8, 2, 3, 3)])
def test_msm_tree_levels_vs_oracle(emu2, oracle, group, n, c, K, K2, levels):
    """msm_tree.cuh kernel bodies (batched-affine pairwise tree, prefix-product tree of several depths, one
    inversion per level) executed thread by thread on the CPU; result against the oracle's multiexp.  The edge
    inputs put a doubling and a cancellation into single buckets, an identity base and zero / one / r-1 scalars."""
    rng = np.random.Generator(np.random.PCG64(n * 131 + c + levels))
    gen = oracle.g1_gen_mul if group == 1 else oracle.g2_gen_mul
    bases = gen(codec.random_fr(n, rng))
    scal = codec.random_fr(n, rng)
    scal[0] = 0
    if n > 12:
        scal[1] = 0; scal[1, 0] = 1
        scal[2] = np.frombuffer((codec.R_MOD - 1).to_bytes(32, "little"), dtype=np.uint8)
        bases[3] = 0                                       # identity base
        bases[4] = bases[5]; scal[4] = scal[5]             # equal points in every bucket they share -> doubling
        bases[6] = bases[7]                                # P and -P in the same buckets -> the pair sums to the identity
        scal[6] = codec.fr_from_ints([(codec.R_MOD - codec.fr_to_ints(scal[7:8])[0]) % codec.R_MOD])[0]
        bases[8] = bases[9]; bases[10] = bases[9]; scal[8] = scal[9]; scal[10] = scal[9]    # three equal points: 2P then 2P + P
    out = np.zeros(96 * group, dtype=np.uint8)
    fn = emu2.emu_msm_tree_g1 if group == 1 else emu2.emu_msm_tree_g2
    assert fn(_p(bases), _p(scal), C.c_size_t(n), c, C.c_uint32(K), C.c_uint32(K2), levels, _p(out)) == 0
    want = oracle.msm_g1(bases, scal) if group == 1 else oracle.msm_g2(bases, scal)
    assert np.array_equal(out, want)




@pytest.mark.parametrize("count,log_s", [(1, 0), (2, 3), (7, 0), (64, 3), (100, 6), (513, 3)])
def test_reduce_bits_tail_vs_oracle(emu2, oracle, count, log_s):
    """The bit-decomposition tail of the bucket reduction (msm_bits_index enumeration + Horner over the bits) equals
    sum_t a_t + 2^log_s sum_t t r_t computed by the oracle's multiexp."""
    rng = np.random.Generator(np.random.PCG64(count * 31 + log_s))
    a = oracle.g1_gen_mul(codec.random_fr(count, rng))
    r = oracle.g1_gen_mul(codec.random_fr(count, rng))
    if count > 3:
        r[2] = 0; a[1] = 0                                   # identities among the items
    # the enumeration itself: every index with bit b set, once, in increasing order
    for b in range(max(1, (count - 1).bit_length())):
        got = []
        j = 0
        while True:
            t = ((j >> b) << (b + 1)) | (1 << b) | (j & ((1 << b) - 1))
            if t >= count:
                break
            got.append(t); j += 1
        assert got == [t for t in range(count) if (t >> b) & 1]
    out = np.zeros(96, dtype=np.uint8)
    nb = emu2.emu_reduce_bits_g1(_p(a), _p(r), C.c_uint32(count), log_s, _p(out))
    assert nb == (0 if count <= 1 else (count - 1).bit_length())
    scal = codec.fr_from_ints([1] * count + [(t << log_s) % codec.R_MOD for t in range(count)])
    want = oracle.msm_g1(np.concatenate([a, r]), scal)
    assert np.array_equal(out, np.asarray(want).reshape(-1)[:96])


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_inverse_by_gcd_equals_fermat(emu, oracle, field):
    """Fp::inv_gcd (binary extended Euclid on the limbs, used for the one inversion per tree level) against the
    Fermat ladder and the oracle, with 0, 1, p - 1, 2 and a power of two among the inputs."""
    rng = np.random.Generator(np.random.PCG64(99))
    width = 32 if field == "fr" else 48
    mod = codec.R_MOD if field == "fr" else 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
    vals = [0, 1, 2, mod - 1, mod - 2, 1 << 200, (mod + 1) // 2] + [int.from_bytes(rng.bytes(width), "little") % mod for _ in range(150)]
    a = np.stack([np.frombuffer(v.to_bytes(width, "little"), dtype=np.uint8) for v in vals])
    got = _op(emu, field, "inv_gcd", a)
    assert np.array_equal(got, _op(emu, field, "inv", a))
    assert np.array_equal(got, oracle.field_op(field, "inv", a))
