"""Synthetic R1CS instances of the SURVEY §8(d) shape, built with numpy only.

Each constraint i is (sum_{j<3} alpha_ij z[a_ij]) * (sum_{j<3} beta_ij z[b_ij]) = z[out_i]:
A and B have 3 nnz/row, C has 1 nnz/row (coefficient 1 on a fresh aux variable), so a
satisfying witness is obtained by evaluating the rows forward.  Rows are grouped in
doubling layers that read only variables of earlier layers, so a layer is one
data-parallel SpMV + pointwise product (on the GPU via b2g16, or in the CPU oracle).

Instance shapes follow the reference: PoRep public inputs = 3 + 18*challenges
(libs/storage/include/nil/filecoin/storage/proofs/porep/stacked/circuit/proof.hpp:186-269,
pinned 22 incl. ONE for one challenge at libs/storage/test/porep/stacked/circuit/proof.cpp:137-155).
"""
import numpy as np

from .codec import random_fr


def layers(n_cons, n_in, n_free):
    """[(first_row, end_row, n_visible_vars)] -- must match oracle orc_synth_witness."""
    out, base, row = [], 1 + n_in + n_free, 0
    while row < n_cons:
        visible = base + row
        size = min(max(visible // 2, 1), n_cons - row)
        out.append((row, row + size, visible))
        row += size
    return out


class SynthR1cs:
    """row_ptr/col/val numpy arrays for A, B, C plus shape numbers."""

    def __init__(self, log_m, n_in, seed, n_free=8, n_cons=None):
        m = 1 << log_m
        if n_cons is None:
            n_cons = m - n_in - 1            # domain trap: n_cons + n_in + 1 <= m exactly
        assert n_cons + n_in + 1 <= m and n_cons >= 1
        rng = np.random.Generator(np.random.PCG64(seed))
        self.log_m, self.n_in, self.n_free, self.n_cons = log_m, n_in, n_free, n_cons
        self.n_aux = n_free + n_cons
        self.n_vars = 1 + n_in + self.n_aux
        acol = np.empty((n_cons, 3), dtype=np.uint32)
        bcol = np.empty((n_cons, 3), dtype=np.uint32)
        for (r0, r1, visible) in layers(n_cons, n_in, n_free):
            acol[r0:r1] = rng.integers(0, visible, size=(r1 - r0, 3), dtype=np.uint32)
            bcol[r0:r1] = rng.integers(0, visible, size=(r1 - r0, 3), dtype=np.uint32)
        rp3 = (np.arange(n_cons + 1, dtype=np.uint64) * 3).astype(np.uint32)
        rp1 = np.arange(n_cons + 1, dtype=np.uint32)
        one = np.zeros((n_cons, 32), dtype=np.uint8)
        one[:, 0] = 1
        ccol = (1 + n_in + n_free + np.arange(n_cons, dtype=np.uint64)).astype(np.uint32)
        self.A = (rp3, acol.reshape(-1), random_fr(3 * n_cons, rng))
        self.B = (rp3.copy(), bcol.reshape(-1), random_fr(3 * n_cons, rng))
        self.C = (rp1, ccol, one)
        head = random_fr(1 + n_in + n_free, rng)
        head[0] = 0
        head[0, 0] = 1                       # z_0 = ONE
        self.z_head = head
        self.seed = seed


def porep_public_inputs(challenges):
    """n_in (excluding ONE) of a stacked-PoRep partition circuit."""
    return 3 + 18 * challenges


# ---- hashed variant: every column / coefficient is a counter-based hash, so libb2g16 can generate the
# same instance on the device (csrc/prove.cu synth_rows_kernel, b2g16_pk_generate_synth) at sizes whose
# host arrays would not fit.  This numpy mirror exists so small instances can be checked against the oracle.
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix(x):
    """splitmix64 on uint64 arrays (wrap-around arithmetic)."""
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def _stream(seed, tag):
    return _mix(np.array([(seed ^ ((tag * 0xD6E8FEB86659FD93) & 0xFFFFFFFFFFFFFFFF)) & 0xFFFFFFFFFFFFFFFF], dtype=np.uint64))[0]


def _hashed_fr(stream, count):
    with np.errstate(over="ignore"):
        idx = np.arange(4 * count, dtype=np.uint64) + stream
    w = _mix(idx).reshape(count, 4)
    w[:, 3] &= np.uint64(0x3FFFFFFFFFFFFFFF)             # < 2^254 < r
    return w.view(np.uint8).reshape(count, 32)


class HashedR1cs(SynthR1cs):
    def __init__(self, log_m, n_in, seed, n_free=8):
        m = 1 << log_m
        n_cons = m - n_in - 1
        assert n_cons >= 1
        self.log_m, self.n_in, self.n_free, self.n_cons, self.seed = log_m, n_in, n_free, n_cons, seed
        self.n_aux = n_free + n_cons
        self.n_vars = 1 + n_in + self.n_aux
        visible = np.empty(n_cons, dtype=np.uint64)
        for (r0, r1, vis) in layers(n_cons, n_in, n_free):
            visible[r0:r1] = 1 + n_in + n_free + r0
        vis3 = np.repeat(visible, 3)
        with np.errstate(over="ignore"):
            e = np.arange(3 * n_cons, dtype=np.uint64)
            acol = ((_mix(e + _stream(seed, 1)) >> np.uint64(11)) % vis3).astype(np.uint32)
            bcol = ((_mix(e + _stream(seed, 2)) >> np.uint64(11)) % vis3).astype(np.uint32)
        rp3 = (np.arange(n_cons + 1, dtype=np.uint64) * 3).astype(np.uint32)
        one = np.zeros((n_cons, 32), dtype=np.uint8)
        one[:, 0] = 1
        ccol = (1 + n_in + n_free + np.arange(n_cons, dtype=np.uint64)).astype(np.uint32)
        self.A = (rp3, acol, _hashed_fr(_stream(seed, 3), 3 * n_cons))
        self.B = (rp3.copy(), bcol, _hashed_fr(_stream(seed, 4), 3 * n_cons))
        self.C = (np.arange(n_cons + 1, dtype=np.uint32), ccol, one)
        head = _hashed_fr(_stream(seed, 5), 1 + n_in + n_free).copy()
        head[0] = 0
        head[0, 0] = 1
        self.z_head = head
