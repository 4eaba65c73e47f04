"""Host-side formats and constants of the verifier (no GPU): the 192-byte proof codec, the
bellperson-layout parameter file, and the constants of csrc/host_bls.hpp against Python big-int
arithmetic.  Anchors: F/constants.hpp:93 (192-byte proofs), S/core/parameter_cache.hpp:50-94,125-153
(.params / .vk cache files), SURVEY App. D (encodings)."""
import os
import re

import numpy as np
import pytest

import util
from b2g16 import lib
from oracle.pyref import bls12_381 as ref

HDR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "crypto3-fil-proofs_b200", "csrc", "host_bls.hpp")


def _const(name):
    src = open(HDR).read()
    m = re.search(r"%s(?:\[\d+\])? = \{?([^;]*?)\}?;" % name, src, re.S)
    limbs = [int(x, 16) for x in re.findall(r"0x([0-9a-f]+)ull", m.group(1))]
    return sum(v << (64 * i) for i, v in enumerate(limbs))


def test_host_constants_match_bigint():
    p, r = ref.P, ref.R
    R = 1 << 384
    assert _const("FQ_P") == p
    assert _const("FQ_INV") == (-pow(p, -1, 1 << 64)) % (1 << 64)
    assert _const("FQ_ONE") == R % p
    assert _const("FQ_R2") == R * R % p
    assert _const("FQ_PM2") == p - 2
    assert _const("FQ_SQRT_EXP") == (p + 1) // 4
    assert _const("FQ_HALF") == (p - 1) // 2
    assert _const("FR_MOD") == r
    assert _const("BLS_X_ABS") == -ref.X_PARAM
    assert (p**2 + 1) * (p**4 - p**2 + 1) % r == 0
    assert _const("FINAL_EXP_TAIL") == (p**2 + 1) * (p**4 - p**2 + 1) // r
    assert (p**6 - 1) * _const("FINAL_EXP_TAIL") == ref.FINAL_EXP


@pytest.mark.parametrize("name", ["groth16_m4", "groth16_m7"])
def test_proof_codec_golden(name):
    g = util.golden(name)
    a, b, c = lib.proof_decode(g["proof"].tobytes())
    assert np.array_equal(a, g["A"]) and np.array_equal(b, g["B"]) and np.array_equal(c, g["C"])
    assert lib.proof_encode(g["A"], g["B"], g["C"]) == g["proof"].tobytes()


def test_proof_codec_generators_and_infinity():
    # the two KATs of SURVEY App. D: compressed generators
    from b2g16 import codec
    g1, g2 = codec.g1_from_points([ref.G1_GEN])[0], codec.g2_from_points([ref.G2_GEN])[0]
    enc = lib.proof_encode(g1, g2, np.zeros(96, np.uint8))
    assert enc[:48] == ref.g1_compress(ref.G1_GEN) and enc[48:144] == ref.g2_compress(ref.G2_GEN)
    assert enc[:4].hex() == "97f1d3a7" and enc[48:52].hex() == "93e02b60"
    assert enc[144] == 0xC0 and not any(enc[145:])
    # the encoder knows the point at infinity (0xC0, zeros), but a *proof* holding it is refused, as bellman's
    # Proof::read refuses it: decode validates proofs, not arbitrary points
    with pytest.raises(lib.B2g16Error):
        lib.proof_decode(enc)
    a, b, c = lib.proof_decode(lib.proof_encode(g1, g2, g1))
    assert np.array_equal(a, g1) and np.array_equal(b, g2) and np.array_equal(c, g1)
    # negated point flips the sort flag only
    neg = codec.g1_from_points([ref.g1_neg(ref.G1_GEN)])[0]
    enc2 = lib.proof_encode(neg, g2, g1)
    assert enc2[0] == enc[0] ^ 0x20 and enc2[1:48] == enc[1:48]
    assert np.array_equal(lib.proof_decode(enc2)[0], neg)


def test_proof_decode_rejects_malformed():
    g = util.golden("groth16_m4")
    good = bytearray(g["proof"].tobytes())
    for mutate in (lambda b: b.__setitem__(0, b[0] & 0x7F),            # compression flag cleared
                   lambda b: b.__setitem__(0, b[0] | 0x40),            # infinity flag with a non-zero x
                   lambda b: b.__setitem__(slice(1, 48), b"\xff" * 47)):  # x >= p
        bad = bytearray(good)
        mutate(bad)
        with pytest.raises(lib.B2g16Error):
            lib.proof_decode(bytes(bad))
    # an x with no point on the curve, or a curve point outside the order-r subgroup: find the first of each
    seen_no_root = seen_off_group = False
    for x in range(1, 200):
        rhs = (x**3 + 4) % ref.P
        y = ref.fq_sqrt(rhs)
        enc = bytearray(x.to_bytes(48, "big")); enc[0] |= 0x80
        bad = bytes(enc) + bytes(good[48:])
        if y is None:
            seen_no_root = True
        elif ref.pt_mul(ref.FQ, (x, y), ref.R) is not None:
            seen_off_group = True
        else:
            continue
        with pytest.raises(lib.B2g16Error):
            lib.proof_decode(bad)
        if seen_no_root and seen_off_group:
            break
    assert seen_no_root and seen_off_group
    with pytest.raises(lib.B2g16Error):
        lib.proof_encode(np.ones(96, np.uint8), g["B"], g["C"])          # not on the curve


def _golden_params(g):
    return {n: g[n] for n in ("alpha_g1", "beta_g1", "beta_g2", "gamma_g2", "delta_g1", "delta_g2", "ic", "h", "l", "a", "b_g1", "b_g2")}


def test_params_file_layout_and_round_trip(tmp_path):
    g = util.golden("groth16_m4")
    path = str(tmp_path / "v28-test.params")
    lib.params_write(path, _golden_params(g))
    raw = open(path, "rb").read()
    # bellperson layout, restated independently with the Python oracle's encoders
    from b2g16 import codec

    def g1u(b):
        p = codec.g1_to_point(b)
        return bytes([0x40]) + bytes(95) if p is None else p[0].to_bytes(48, "big") + p[1].to_bytes(48, "big")

    def g2u(b):
        p = codec.g2_to_point(b)
        if p is None:
            return bytes([0x40]) + bytes(191)
        (x0, x1), (y0, y1) = p
        return b"".join(v.to_bytes(48, "big") for v in (x1, x0, y1, y0))

    want = g1u(g["alpha_g1"]) + g1u(g["beta_g1"]) + g2u(g["beta_g2"]) + g2u(g["gamma_g2"]) + g1u(g["delta_g1"]) + g2u(g["delta_g2"])
    for n, enc in (("ic", g1u), ("h", g1u), ("l", g1u), ("a", g1u), ("b_g1", g1u), ("b_g2", g2u)):
        want += len(g[n]).to_bytes(4, "big") + b"".join(enc(row) for row in g[n])
    assert raw == want
    back = lib.params_read(path)
    for n, v in _golden_params(g).items():
        assert np.array_equal(back[n].reshape(v.shape), v), n
    assert np.array_equal(back["vk"][:96], g["alpha_g1"]) and np.array_equal(back["vk"][672:].reshape(-1, 96), g["ic"])
    # truncated file and bad flag byte are errors, not crashes
    open(path, "wb").write(raw[:-7])
    with pytest.raises(lib.B2g16Error):
        lib.params_read(path)
    open(path, "wb").write(bytes([0x80]) + raw[1:])
    with pytest.raises(lib.B2g16Error):
        lib.params_read(path)
    with pytest.raises(lib.B2g16Error):
        lib.params_read(str(tmp_path / "missing.params"))


def test_density_filtered_queries():
    """bellman-style files keep a / b_g1 / b_g2 without their points at infinity; the bitmap that puts them
    back is the column occupancy of A (resp. B).  Pinned on the golden key: a query point is the identity
    exactly where the matrix has no entry for that variable."""
    from b2g16 import synth
    g = util.golden("groth16_m4")
    s = synth.SynthR1cs(int(g["log_m"]), int(g["n_in"]), int(g["seed"]), n_free=int(g["n_free"]))
    n_vars = g["a"].shape[0]
    for mat, n_always, names in ((s.A, s.n_in + 1, (("a", 96),)), (s.B, 0, (("b_g1", 96), ("b_g2", 192)))):
        dens, cnt = lib.csr_density(mat[1], n_vars, n_always)
        assert cnt == int(dens.sum())
        for n, w in names:
            assert np.array_equal(dens, g[n].any(axis=1).astype(np.uint8)), n
            filt, dens2 = lib.query_filter(g[n], w)
            assert np.array_equal(dens2, dens) and filt.shape == (cnt, w)
            assert np.array_equal(filt, g[n][dens.astype(bool)])
            assert np.array_equal(lib.query_expand(filt, dens, w), g[n])
            with pytest.raises(lib.B2g16Error):
                lib.query_expand(filt[:-1], dens, w)                         # file and constraint system disagree
    # empty matrix / nothing filtered / everything filtered
    dens, cnt = lib.csr_density(np.zeros(0, np.uint32), 5)
    assert cnt == 0 and not dens.any()
    assert not lib.query_expand(np.zeros((0, 96), np.uint8), dens, 96).any()
    full = np.arange(3 * 192, dtype=np.uint32).astype(np.uint8).reshape(3, 192) | 1
    filt, d = lib.query_filter(full, 192)
    assert d.all() and np.array_equal(filt, full)
    with pytest.raises(lib.B2g16Error):
        lib.csr_density(np.array([7], np.uint32), 5)                         # column out of range
    import ctypes as C
    ones, out = np.ones(3, np.uint8), np.zeros(3 * 192, np.uint8)
    rc = lib.load().b2g16_query_expand(lib._p(full), C.c_uint64(3), lib._p(ones), C.c_uint64(3), C.c_size_t(100), lib._p(out))
    assert rc == -1                                                          # not a point size


def test_density_of_a_query_keeps_inputs_absent_from_A(oracle):
    """a public input that occurs only in C (pack_into_inputs-style) has no entry in any row of A, yet its
    `a` point is not the identity: the input-consistency row n_cons + j contributes L_{n_cons+j}(tau) to u_j (SURVEY
    App. B step 2), and bellman keeps every input in the a query.  Plain column occupancy of A would drop it and
    then refuse the (correct) filtered file; b2g16_query_density(A, n_always = n_in + 1) must not."""
    from b2g16 import synth
    s = synth.SynthR1cs(5, 3, 77, n_free=4)
    acol = s.A[1].copy()
    acol[acol == 2] = 0                                       # input 2 no longer occurs in A
    bcol = s.B[1].copy()
    bcol[bcol == 2] = 0
    s.A = (s.A[0], acol, s.A[2]); s.B = (s.B[0], bcol, s.B[2])
    cs = util.oracle_cs(oracle, s)
    toxic, _, _ = util.toxic_from_seed(77)
    key = oracle.groth16_setup(cs, toxic)
    assert key.a[2].any(), "input-consistency row: the a point of an input is never the identity"
    assert not key.b_g1[2].any()
    plain, cnt_plain = lib.csr_density(acol, s.n_vars)
    dens, cnt = lib.csr_density(acol, s.n_vars, s.n_in + 1)
    assert plain[2] == 0 and dens[2] == 1 and cnt == cnt_plain + 1
    filt, dens_file = lib.query_filter(key.a, 96)
    assert np.array_equal(dens_file, dens) and filt.shape[0] == cnt
    assert np.array_equal(lib.query_expand(filt, dens, 96), key.a)
    with pytest.raises(lib.B2g16Error):
        lib.query_expand(filt, plain, 96)                     # what round 1 did: refuses a correct file
    dens_b, cnt_b = lib.csr_density(bcol, s.n_vars, 0)
    filt_b, dens_b_file = lib.query_filter(key.b_g1, 96)
    assert np.array_equal(dens_b_file, dens_b) and filt_b.shape[0] == cnt_b
