"""Shared helpers for the parity tests."""
import os

import numpy as np

from b2g16 import codec, synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def oracle_cs(coracle, s):
    """SynthR1cs -> oracle R1csArrays."""
    return coracle.R1csArrays(s.log_m, s.n_in, s.n_aux, s.A, s.B, s.C)


def synth_instance(coracle, log_m, n_in, seed, n_free=8, n_cons=None):
    """(SynthR1cs, oracle cs, witness bytes) with the witness from the CPU oracle."""
    s = synth.SynthR1cs(log_m, n_in, seed, n_free=n_free, n_cons=n_cons)
    cs = oracle_cs(coracle, s)
    z = coracle.synth_witness(cs, s.z_head, s.n_free)
    return s, cs, z


def toxic_from_seed(seed):
    rng = np.random.Generator(np.random.PCG64(seed ^ 0x70C51C))
    t = codec.random_fr(7, rng)
    return t[:5], t[5], t[6]


def pairing_verify(coracle, vk, public_inputs, proof):
    """Groth16 verification e(A,B) = e(alpha,beta) e(sum x_j IC_j, gamma) e(C,delta) with the big-int
    Python oracle's pairing (oracle/pyref) -- the prove -> verify == true round trip the reference's
    compound tests pin (libs/storage/test/**/compound.cpp).  vk: bytes from b2g16_pk_generate
    (alpha_g1 | beta_g2 | gamma_g2 | delta_g2 | IC...); public_inputs: uint8[n_in, 32]; the IC
    combination is done by the C++ oracle (fast), the pairing product in Python (~3 s)."""
    from oracle.pyref import bls12_381 as c
    vk = np.asarray(vk, dtype=np.uint8)
    ic = vk[672:].reshape(-1, 96)
    xs = np.zeros((ic.shape[0], 32), dtype=np.uint8)
    xs[0, 0] = 1
    xs[1:] = np.asarray(public_inputs, dtype=np.uint8).reshape(-1, 32)
    acc = codec.g1_to_point(coracle.msm_g1(ic, xs))
    A = c.g1_decompress(proof[:48]); B = c.g2_decompress(proof[48:144]); C = c.g1_decompress(proof[144:])
    return c.pairing_product_is_one([(c.g1_neg(A), B), (codec.g1_to_point(vk[:96]), codec.g2_to_point(vk[96:288])),
                                     (acc, codec.g2_to_point(vk[288:480])), (C, codec.g2_to_point(vk[480:672]))])
