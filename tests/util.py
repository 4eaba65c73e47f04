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
