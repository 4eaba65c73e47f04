"""Pins the fast C++ oracle (oracle/c/oracle.cpp) against the golden vectors made by
the slow Python oracle and against the reference's own KATs.
CPU only."""
import numpy as np
import pytest

from b2g16 import codec, synth
import util


def test_fr32_kats_reference(oracle):
    # libs/storage/test/core/fr32.cpp:51-85 and libs/filecoin/test/api/mod.cpp:35-43
    g = util.golden("fields")
    for b, ok in zip(g["fr32_kat_bytes"], g["fr32_kat_ok"]):
        assert oracle.fr_is_canonical(b.tobytes()) == bool(ok)


def test_fr_roundtrip_100(oracle):
    # libs/filecoin/test/api/mod.cpp:91-100 : Fr -> bytes -> Fr round trip x100
    rng = np.random.Generator(np.random.PCG64(7))
    v = codec.random_fr(100, rng)
    one = codec.fr_from_ints([1] * 100)
    assert np.array_equal(oracle.field_op("fr", "mul", v, one), v)


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_field_ops(oracle, field):
    g = util.golden("fields")
    a, b = g[field + "_a"], g[field + "_b"]
    for op in ("add", "sub", "mul"):
        assert np.array_equal(oracle.field_op(field, op, a, b), g[f"{field}_{op}"]), op
    for op in ("sqr", "neg", "inv"):
        assert np.array_equal(oracle.field_op(field, op, a), g[f"{field}_{op}"]), op


def test_ntt_golden(oracle):
    g = util.golden("ntt")
    for ln in range(0, 11):
        for inv in (0, 1):
            for cos in (0, 1):
                out = oracle.ntt(g[f"in_{ln}"], bool(inv), bool(cos))
                assert np.array_equal(out, g[f"out_{ln}_{inv}{cos}"]), (ln, inv, cos)


def test_ntt_roundtrip_large(oracle):
    rng = np.random.Generator(np.random.PCG64(3))
    v = codec.random_fr(1 << 14, rng)
    assert np.array_equal(oracle.ntt(oracle.ntt(v), inverse=True), v)
    assert np.array_equal(oracle.ntt(oracle.ntt(v, coset=True), inverse=True, coset=True), v)


def test_msm_golden(oracle):
    g = util.golden("msm")
    for n in (1, 2, 3, 17, 256):
        assert np.array_equal(oracle.msm_g1(g[f"g1_bases_{n}"], g[f"scalars_{n}"]), g[f"g1_out_{n}"]), n
        assert np.array_equal(oracle.msm_g2(g[f"g2_bases_{n}"], g[f"scalars_{n}"]), g[f"g2_out_{n}"]), n


def test_msm_linearity(oracle):
    # sum k_i (s_i G) == (sum k_i s_i) G  -- size-independent property
    rng = np.random.Generator(np.random.PCG64(11))
    n = 3000
    s, k = codec.random_fr(n, rng), codec.random_fr(n, rng)
    tot = sum(x * y for x, y in zip(codec.fr_to_ints(s), codec.fr_to_ints(k))) % codec.R_MOD
    t = codec.fr_from_ints([tot])
    assert np.array_equal(oracle.msm_g1(oracle.g1_gen_mul(s), k), oracle.g1_gen_mul(t)[0])
    assert np.array_equal(oracle.msm_g2(oracle.g2_gen_mul(s[:500]), k[:500]),
                          oracle.g2_gen_mul(codec.fr_from_ints(
                              [sum(x * y for x, y in zip(codec.fr_to_ints(s[:500]), codec.fr_to_ints(k[:500]))) % codec.R_MOD]))[0])


def test_msm_empty_and_identity(oracle):
    assert not oracle.msm_g1(np.zeros((0, 96), np.uint8), np.zeros((0, 32), np.uint8)).any()
    rng = np.random.Generator(np.random.PCG64(12))
    k = codec.random_fr(10, rng)
    assert not oracle.msm_g1(np.zeros((10, 96), np.uint8), k).any()
    assert not oracle.msm_g2(np.zeros((10, 192), np.uint8), k).any()


@pytest.mark.parametrize("log_m", [4, 7])
def test_groth16_golden(oracle, log_m):
    g = util.golden(f"groth16_m{log_m}")
    s = synth.SynthR1cs(int(g["log_m"]), int(g["n_in"]), int(g["seed"]), n_free=int(g["n_free"]))
    cs = util.oracle_cs(oracle, s)
    z = oracle.synth_witness(cs, s.z_head, s.n_free)
    assert np.array_equal(z, g["z"])
    key = oracle.groth16_setup(cs, g["toxic"])
    for name in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "gamma_g2", "delta_g2", "ic", "a", "b_g1", "b_g2", "l", "h"):
        assert np.array_equal(getattr(key, name), g[name]), name
    proof, tr = oracle.groth16_prove(key, z, g["r"], g["s"])
    assert len(proof) == 192          # constants.hpp:93 SINGLE_PARTITION_PROOF_LEN
    for name, gname in (("a", "ev_a"), ("b", "ev_b"), ("c", "ev_c"), ("h", "h_coeffs"), ("msm_a", "msm_a"),
                        ("msm_b1", "msm_b1"), ("msm_b2", "msm_b2"), ("msm_l", "msm_l"), ("msm_h", "msm_h"),
                        ("A", "A"), ("B", "B"), ("C", "C")):
        assert np.array_equal(tr[name], g[gname]), name
    assert proof == g["proof"].tobytes()


def test_groth16_verifies_with_pairing(oracle):
    """prove -> verify == true round trip (libs/storage/test/**/compound.cpp), on an
    instance bigger than the golden ones; the verifier is the Python pairing."""
    from oracle.pyref import bls12_381 as c, groth16 as pg
    s, cs, z = util.synth_instance(oracle, 9, 4, seed=99)
    toxic, r, sr = util.toxic_from_seed(99)
    key = oracle.groth16_setup(cs, toxic)
    proof, tr = oracle.groth16_prove(key, z, r, sr)
    pk = pg.ProvingKey()
    pk.alpha_g1, pk.beta_g2 = codec.g1_to_point(key.alpha_g1), codec.g2_to_point(key.beta_g2)
    pk.gamma_g2, pk.delta_g2 = codec.g2_to_point(key.gamma_g2), codec.g2_to_point(key.delta_g2)
    pk.ic = [codec.g1_to_point(p) for p in key.ic]
    pub = codec.fr_to_ints(z[1:1 + s.n_in])
    assert pg.verify(pk, pub, proof)
    pub[0] = (pub[0] + 1) % c.R
    assert not pg.verify(pk, pub, proof)
    # h has degree <= m-2
    assert not tr["h"][-1].any()
