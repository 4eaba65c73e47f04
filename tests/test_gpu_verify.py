"""GPU parity for the verifier, the caller-supplied-abc prover entry and the parameter-file round trip
(SURVEY §8 rows a18, f2, f3).  The oracle here is the big-int Python pairing (oracle/pyref) through
util.pairing_verify and the golden proofs; the reference pins prove -> verify == true
(libs/storage/test/porep/stacked/circuit/proof.cpp:257-273, libs/storage/test/core/components/por.cpp:353-362)."""
import numpy as np
import pytest

import util
from b2g16 import lib, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def oracle():
    from oracle import coracle
    coracle.build()
    return coracle


def _golden_vk(g):
    return np.concatenate([g["alpha_g1"], g["beta_g2"], g["gamma_g2"], g["delta_g2"], g["ic"].reshape(-1)])


@pytest.mark.parametrize("name", ["groth16_m4", "groth16_m7"])
def test_verify_golden(ctx, oracle, name):
    g = util.golden(name)
    n_in = int(g["n_in"])
    pvk = ctx.vk_prepare(_golden_vk(g))
    pub = g["z"][1:1 + n_in]
    proof = g["proof"].tobytes()
    assert ctx.verify(pvk, pub, proof)
    assert util.pairing_verify(oracle, _golden_vk(g), pub, proof)            # same decision as the Python pairing
    # a perturbed public input, a swapped A/C and a flipped y-sign are rejected (and the oracle agrees)
    bad = pub.copy(); bad[0, 0] ^= 1
    assert not ctx.verify(pvk, bad, proof) and not util.pairing_verify(oracle, _golden_vk(g), bad, proof)
    swapped = proof[144:] + proof[48:144] + proof[:48]
    assert not ctx.verify(pvk, pub, swapped)
    flipped = bytes([proof[0] ^ 0x20]) + proof[1:]
    assert not ctx.verify(pvk, pub, flipped)
    assert not ctx.verify(pvk, pub, bytes([proof[0] & 0x7F]) + proof[1:])     # malformed: reject, not raise
    with pytest.raises(lib.B2g16Error):
        ctx.verify(pvk, g["z"][1:n_in], proof)                                # wrong input count
    pvk.free()


def test_vk_prepare_rejects_bad_key(ctx):
    g = util.golden("groth16_m4")
    vk = _golden_vk(g).copy()
    vk[0] ^= 1                                                                # alpha_g1 off the curve
    with pytest.raises(lib.B2g16Error):
        ctx.vk_prepare(vk)
    vk = _golden_vk(g).copy()
    vk[672 + 5] ^= 1                                                          # an IC point off the curve
    with pytest.raises(lib.B2g16Error):
        ctx.vk_prepare(vk)
    # an IC point on the curve but outside the order-r subgroup (bellman's VerifyingKey::read checks this)
    from oracle.pyref import bls12_381 as ref
    from b2g16 import codec
    for x in range(1, 200):
        y = ref.fq_sqrt((x**3 + 4) % ref.P)
        if y is not None and ref.pt_mul(ref.FQ, (x, y), ref.R) is not None:
            break
    vk = _golden_vk(g).copy()
    vk[672 + 96:672 + 192] = codec.g1_from_points([(x, y)])[0]
    with pytest.raises(lib.B2g16Error) as e:
        ctx.vk_prepare(vk)
    assert e.value.code == -3 and "subgroup" in str(e.value)
    # a proof holding the point at infinity is a reject, not an error and not an accept
    pvk = ctx.vk_prepare(_golden_vk(g))
    inf_c = g["proof"].tobytes()[:144] + bytes([0xC0]) + bytes(47)
    assert not ctx.verify(pvk, g["z"][1:1 + int(g["n_in"])], inf_c)
    pvk.free()


def test_prove_verify_batch(ctx, oracle):
    """Several partitions under one key (multi_proof, F/api/seal.hpp:304-308): every proof verifies alone,
    the batch verifies as one product, and one bad proof or input sinks the batch."""
    s, cs, z0 = util.synth_instance(oracle, 9, 5, seed=77)
    toxic, r, sr = util.toxic_from_seed(77)
    pk, vk = ctx.pk_generate(cs, toxic, want_vk=True)
    pvk = ctx.vk_prepare(vk)
    zs, rs, ss = [], [], []
    rng = np.random.Generator(np.random.PCG64(9))
    from b2g16 import codec
    for k in range(3):
        head = s.z_head.copy()
        head[1:] = codec.random_fr(head.shape[0] - 1, rng)
        zs.append(oracle.synth_witness(cs, head, s.n_free))
        t = codec.random_fr(2, rng)
        rs.append(t[0]); ss.append(t[1])
    proofs = ctx.prove_batch(pk, np.concatenate(zs), np.stack(rs), np.stack(ss))
    assert len(proofs) == 3 * 192
    pubs = np.stack([z[1:1 + s.n_in] for z in zs])
    for k in range(3):
        assert ctx.verify(pvk, pubs[k], proofs[192 * k:192 * (k + 1)])
    want, _ = oracle.groth16_prove(oracle.groth16_setup(cs, toxic), zs[1], rs[1], ss[1], want_trace=False)
    assert proofs[192:384] == want
    assert ctx.verify_batch(pvk, pubs, proofs)
    assert ctx.verify_batch(pvk, pubs, proofs, rand=codec.random_fr(3, rng))
    assert ctx.verify_batch(pvk, pubs[:0], b"")                              # empty batch
    mixed = proofs[:192] + proofs[384:] + proofs[192:384]                     # proofs 1 and 2 swapped
    assert not ctx.verify_batch(pvk, pubs, mixed)
    bad = pubs.copy(); bad[2, 1, 0] ^= 1
    assert not ctx.verify_batch(pvk, bad, proofs)
    pvk.free(); pk.free()


def test_prove_abc_and_params_file(ctx, oracle, tmp_path):
    """The bellperson-style flow: key from a .params file, a/b/c from the caller's synthesis.  Same
    192 bytes as the CSR path and as the CPU oracle."""
    s, cs, z = util.synth_instance(oracle, 10, 3, seed=31)
    toxic, r, sr = util.toxic_from_seed(31)
    pk, vk = ctx.pk_generate(cs, toxic, want_vk=True)
    m, nv = 1 << s.log_m, cs.n_vars
    params = {"gamma_g2": vk[288:480], "ic": vk[672:].reshape(-1, 96)}
    for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", cs.n_aux), (4, "h", m - 1)):
        params[n] = pk.export(which, cnt)
    for which, n in ((5, "alpha_g1"), (6, "beta_g1"), (7, "delta_g1"), (8, "beta_g2"), (9, "delta_g2")):
        params[n] = pk.export(which, 1)[0]
    want = ctx.prove(pk, z, r, sr)
    pk.free()
    path = str(tmp_path / "v28-synth.params")
    lib.params_write(path, params)
    back = lib.params_read(path)
    assert np.array_equal(back["vk"], vk)

    class Key:
        pass
    key = Key()
    for n in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h"):
        setattr(key, n, back[n])

    class ShapeOnly:                                                          # no matrices: a, b, c come from the caller
        log_m, n_in, n_aux, n_cons, mats = s.log_m, s.n_in, s.n_aux, s.n_cons, None
    pk2 = ctx.pk_load(key, ShapeOnly)
    with pytest.raises(lib.B2g16Error):
        ctx.prove(pk2, z, r, sr)                                              # needs matrices
    okey = oracle.groth16_setup(cs, toxic)
    oproof, tr = oracle.groth16_prove(okey, z, r, sr, want_trace=True)
    got = ctx.prove_abc(pk2, tr["a"], tr["b"], tr["c"], z, r, sr)
    assert got == want == oproof
    pvk = ctx.vk_prepare(back["vk"])
    assert ctx.verify(pvk, z[1:1 + s.n_in], got)
    pvk.free(); pk2.free()


def test_streamed_key_load_from_file(ctx, oracle, tmp_path):
    """b2g16_pk_load_file: the `.params` cache entry streamed to the device (pinned ring, device-side byte order /
    Montgomery conversion, device-side expansion of density-filtered queries) gives the same resident key -- hence the same
    192 proof bytes -- as b2g16_pk_load from host arrays, for a dense file and for a bellman-style filtered one.
    (mapped_scheme_params.hpp:43-84, parameter_cache.hpp:125-129)"""
    s, cs, z = util.synth_instance(oracle, 10, 3, seed=57)
    toxic, r, sr = util.toxic_from_seed(57)
    pk, vk = ctx.pk_generate(cs, toxic, want_vk=True)
    m, nv = 1 << s.log_m, cs.n_vars
    params = {"gamma_g2": vk[288:480], "ic": vk[672:].reshape(-1, 96)}
    for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", cs.n_aux), (4, "h", m - 1)):
        params[n] = pk.export(which, cnt)
    for which, n in ((5, "alpha_g1"), (6, "beta_g1"), (7, "delta_g1"), (8, "beta_g2"), (9, "delta_g2")):
        params[n] = pk.export(which, 1)[0]
    want = ctx.prove(pk, z, r, sr)
    pk.free()
    dense_path, filt_path = str(tmp_path / "dense.params"), str(tmp_path / "filtered.params")
    lib.params_write(dense_path, params)
    filt = dict(params)
    kept = {}
    for n, w in (("a", 96), ("b_g1", 96), ("b_g2", 192)):
        filt[n], _ = lib.query_filter(params[n], w)
        kept[n] = len(filt[n])
    assert kept["a"] < nv and kept["b_g1"] == kept["b_g2"] < nv      # the synthetic instance has empty columns
    lib.params_write(filt_path, filt)
    mp = lib.params_map(filt_path)
    assert mp["ic"] == (cs.n_in + 1, 868) and mp["h"][0] == m - 1 and mp["a"][0] == kept["a"] and mp["b_g2"][0] == kept["b_g2"]
    assert mp["h"][1] == 868 + 96 * (cs.n_in + 1) + 4
    ctx.set_tuning("key_chunk_mib", 1)                               # several ring turns even at this size
    for path in (dense_path, filt_path):
        pk2, vk2 = ctx.pk_load_file(path, cs, want_vk=True)
        assert np.array_equal(vk2, vk)
        for which, n, cnt in ((0, "a", nv), (1, "b_g1", nv), (2, "b_g2", nv), (3, "l", cs.n_aux), (4, "h", m - 1)):
            assert np.array_equal(pk2.export(which, cnt), params[n]), (path, n)
        assert ctx.prove(pk2, z, r, sr) == want
        pk2.free()
    ctx.set_tuning("key_chunk_mib", 32)
    # a corrupted coordinate (not on the curve) and a truncated file are errors
    raw = bytearray(open(dense_path, "rb").read())
    raw[mp["h"][1] + 96 * 5 + 95] ^= 1
    bad_path = str(tmp_path / "bad.params")
    open(bad_path, "wb").write(raw)
    with pytest.raises(lib.B2g16Error) as e:
        ctx.pk_load_file(bad_path, cs)
    assert e.value.code in (-2, -3)
    open(bad_path, "wb").write(bytes(raw[:-50]))
    with pytest.raises(lib.B2g16Error):
        ctx.pk_load_file(bad_path, cs)
