"""The crypto3-shaped C++ shim (include/nil/crypto3/b2g16_shim.hpp) driven from a C++ program the
way compound_proof::circuit_proofs would call crypto3
(libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137):
multiexp, evaluation_domain and prove<r1cs_gg_ppzksnark<bls12<381>>> against the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from b2g16 import codec
import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "crypto3-fil-proofs_b200")
BIN = os.path.join(PKG, "lib", "shim_test")


def _build():
    subprocess.check_call(["make", "-C", PKG, "-s", "lib/shim_test"])


def test_shim_compiles_and_has_no_cpu_fallback():
    """-m "not gpu": the shim compiles against include/b2g16.h and, with no device, the first call
    throws b2g16::error(B2G16_E_CUDA) instead of computing anything on the CPU."""
    import torch
    _build()
    if torch.cuda.is_available():
        pytest.skip("a device is present; the no-device behaviour is checked on the CPU box")
    r = subprocess.run([BIN, "--no-gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "no device" in r.stdout


@pytest.mark.gpu
def test_shim_against_oracle(oracle, tmp_path):
    if not os.path.exists(BIN):
        _build()
    rng = np.random.Generator(np.random.PCG64(77))
    d = str(tmp_path)

    def put(name, arr):
        np.ascontiguousarray(arr).tofile(os.path.join(d, name + ".bin"))

    n = 700
    s = codec.random_fr(n, rng)
    b1 = oracle.g1_gen_mul(codec.random_fr(n, rng))
    b2 = oracle.g2_gen_mul(codec.random_fr(200, rng))
    put("msm_g1_bases", b1); put("msm_g2_bases", b2); put("msm_scalars", s)
    put("msm_g1_out", oracle.msm_g1(b1, s)); put("msm_g2_out", oracle.msm_g2(b2, s[:200]))
    x = codec.random_fr(1 << 12, rng)
    put("ntt_in", x); put("ntt_fwd", oracle.ntt(x)); put("ntt_icoset", oracle.ntt(x, inverse=True, coset=True))
    sy, cs, z = util.synth_instance(oracle, 9, 3, seed=4242)
    toxic, r, sr = util.toxic_from_seed(4242)
    key = oracle.groth16_setup(cs, toxic)
    proof, tr = oracle.groth16_prove(key, z, r, sr)
    put("cs_shape", np.array([cs.log_m, cs.n_cons, cs.n_in, cs.n_aux], dtype=np.uint64))
    for name, (rp, col, val) in zip("ABC", cs.mats):
        put(name + "_rp", rp); put(name + "_col", col); put(name + "_val", val)
    put("toxic", toxic); put("z", z); put("rs", np.stack([r, sr])); put("ic", key.ic)
    put("proof", np.frombuffer(proof, dtype=np.uint8)); put("proof_A", tr["A"]); put("proof_B", tr["B"]); put("proof_C", tr["C"])
    res = subprocess.run([BIN, d], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "shim test passed" in res.stdout
