"""Groth16 (r1cs_gg_ppzksnark) setup / prove / verify + radix-2 evaluation domain,
restated from the definitions in SURVEY.md Appendix B.  TEST INFRASTRUCTURE ONLY.

Reference anchors (paths relative to /root/reference):
  * key layout  vk,h,l,a,b_g1,b_g2:
      libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:46-66
      (h[i] = tau^i t(tau)/delta :48-50, l :52-55, a :57-60, b_g1/b_g2 :62-66)
  * call site of the prover: .../core/proof/compound_proof.hpp:127-137 (stub)
  * proof = (A in G1, B in G2, C in G1), 192 bytes:
      .../core/proof/multi_proof.hpp:38-58, libs/filecoin/include/nil/filecoin/proofs/constants.hpp:93
The prover/witness-map algorithm itself lives in the un-vendored
nilfoundation/zk @ 44a19899 (r1cs_gg_ppzksnark_prover, r1cs_to_qap::witness_map) and
crypto3-math @ 16c55f5c (basic_radix2_domain); restated here from the published
Groth16 construction.  Parity unpinned at byte level (no golden vectors upstream).
"""
import random

from .bls12_381 import (R, FR_GENERATOR, fr_inv, fr_root_of_unity, G1_GEN, G2_GEN,
                        g1_add, g1_mul, g1_neg, g2_add, g2_mul, g1_gen_mul, g2_gen_mul,
                        msm_g1, msm_g2, g1_compress, g2_compress, g1_decompress,
                        g2_decompress, pairing_product_is_one)


# ------------------------------------------------------------ evaluation domain
def ntt(a, inverse=False):
    """Natural-order in/out radix-2 NTT over Fr.  X[k] = sum_n x[n] w^(nk);
    inverse uses w^-1 and scales by 1/n.  (evaluation_domain::fft / inverse_fft)"""
    n = len(a)
    log_n = n.bit_length() - 1
    assert 1 << log_n == n
    a = [v % R for v in a]
    # bit reversal
    j = 0
    for i in range(1, n):
        bit = n >> 1
        while j & bit:
            j ^= bit
            bit >>= 1
        j |= bit
        if i < j:
            a[i], a[j] = a[j], a[i]
    w_n = fr_root_of_unity(log_n)
    if inverse:
        w_n = fr_inv(w_n)
    length = 2
    while length <= n:
        w_len = pow(w_n, n // length, R)
        half = length >> 1
        for start in range(0, n, length):
            w = 1
            for k in range(start, start + half):
                u, v = a[k], a[k + half] * w % R
                a[k] = (u + v) % R
                a[k + half] = (u - v) % R
                w = w * w_len % R
        length <<= 1
    if inverse:
        n_inv = fr_inv(n)
        a = [v * n_inv % R for v in a]
    return a


def dft_naive(a, inverse=False):
    """O(n^2) definition, used to pin ntt() itself."""
    n = len(a)
    log_n = n.bit_length() - 1
    w = fr_root_of_unity(log_n)
    if inverse:
        w = fr_inv(w)
    out = []
    for k in range(n):
        wk = pow(w, k, R)
        acc, cur = 0, 1
        for v in a:
            acc = (acc + v * cur) % R
            cur = cur * wk % R
        out.append(acc)
    if inverse:
        ni = fr_inv(n)
        out = [v * ni % R for v in out]
    return out


def coset_ntt(a, g=FR_GENERATOR):
    """Evaluate the polynomial with coefficients a on the coset g*D."""
    cur, out = 1, []
    for v in a:
        out.append(v * cur % R)
        cur = cur * g % R
    return ntt(out)


def coset_intt(a, g=FR_GENERATOR):
    c = ntt(a, inverse=True)
    gi = fr_inv(g)
    cur, out = 1, []
    for v in c:
        out.append(v * cur % R)
        cur = cur * gi % R
    return out


def transform(a, inverse, coset):
    """The four modes the C ABI `b2g16_ntt` exposes."""
    if coset:
        return coset_intt(a) if inverse else coset_ntt(a)
    return ntt(a, inverse)


# ------------------------------------------------------------------------ R1CS
class R1CS:
    """Sparse R1CS in CSR form.  Variable order z = (1, x_1..x_nin, w_1..w_naux).
    Each of A,B,C is (row_ptr[n_cons+1], col[nnz], val[nnz])."""

    def __init__(self, n_in, n_aux, A, B, C):
        self.n_in, self.n_aux = n_in, n_aux
        self.A, self.B, self.C = A, B, C
        self.n_cons = len(A[0]) - 1
        assert len(B[0]) - 1 == self.n_cons and len(C[0]) - 1 == self.n_cons

    @property
    def n_vars(self):
        return 1 + self.n_in + self.n_aux

    def domain_log(self):
        need = self.n_cons + self.n_in + 1
        k = 0
        while (1 << k) < need:
            k += 1
        return k

    @staticmethod
    def spmv(M, z):
        rp, col, val = M
        return [sum(val[k] * z[col[k]] for k in range(rp[i], rp[i + 1])) % R
                for i in range(len(rp) - 1)]

    def is_satisfied(self, z):
        a, b, c = self.spmv(self.A, z), self.spmv(self.B, z), self.spmv(self.C, z)
        return all((x * y - w) % R == 0 for x, y, w in zip(a, b, c))


def synth_layout(n_cons, n_in, n_free):
    """Layer boundaries of the synthetic instance: constraint i defines aux
    variable  1+n_in+n_free+i  and may reference only variables of earlier layers;
    layers double in size so a whole layer can be evaluated in parallel.
    Returns list of (first_row, end_row, n_visible_vars)."""
    layers = []
    base = 1 + n_in + n_free
    row = 0
    while row < n_cons:
        visible = base + row
        size = min(max(visible // 2, 1), n_cons - row)
        layers.append((row, row + size, visible))
        row += size
    return layers


def synth_r1cs(n_cons, n_in, seed, n_free=8, small_coeff=False):
    """Random satisfiable R1CS of the SURVEY §8(d) shape: each constraint
    (sum_{j<3} alpha_j z_aj) * (sum_{j<3} beta_j z_bj) = z_out, A,B 3 nnz/row, C 1 nnz/row,
    witness built forward.  Returns (R1CS, z)."""
    rng = random.Random(seed)
    n_aux = n_free + n_cons
    z = [1] + [rng.randrange(R) for _ in range(n_in + n_free)]
    arp, acol, aval = [0], [], []
    brp, bcol, bval = [0], [], []
    crp, ccol, cval = [0], [], []
    base = 1 + n_in + n_free
    for (r0, r1, visible) in synth_layout(n_cons, n_in, n_free):
        new = []
        for i in range(r0, r1):
            sa = sb = 0
            for _ in range(3):
                c = rng.randrange(visible)
                v = rng.randrange(1, 4) if small_coeff else rng.randrange(1, R)
                acol.append(c); aval.append(v); sa += v * z[c]
            for _ in range(3):
                c = rng.randrange(visible)
                v = rng.randrange(1, 4) if small_coeff else rng.randrange(1, R)
                bcol.append(c); bval.append(v); sb += v * z[c]
            arp.append(len(acol)); brp.append(len(bcol))
            ccol.append(base + i); cval.append(1); crp.append(len(ccol))
            new.append(sa * sb % R)
        z.extend(new)
    cs = R1CS(n_in, n_aux, (arp, acol, aval), (brp, bcol, bval), (crp, ccol, cval))
    return cs, z


# ----------------------------------------------------------------------- keys
class ProvingKey:
    """crypto3/libsnark-style *dense* key: one base per variable (identity where
    the QAP polynomial is zero), plus the constraint system."""

    def __init__(self):
        self.alpha_g1 = self.beta_g1 = self.delta_g1 = None
        self.beta_g2 = self.gamma_g2 = self.delta_g2 = None
        self.ic = []      # n_in+1 G1
        self.a = []       # n_vars G1
        self.b_g1 = []    # n_vars G1
        self.b_g2 = []    # n_vars G2
        self.l = []       # n_aux G1
        self.h = []       # m-1 G1
        self.log_m = 0
        self.cs = None


def lagrange_at(tau, log_m):
    """L_i(tau) for the size-2^log_m domain: (tau^m - 1)/m * w^i / (tau - w^i)."""
    m = 1 << log_m
    w = fr_root_of_unity(log_m)
    zt = (pow(tau, m, R) - 1) % R
    pref = zt * fr_inv(m) % R
    out, wi = [], 1
    for _ in range(m):
        out.append(pref * wi % R * fr_inv((tau - wi) % R) % R)
        wi = wi * w % R
    return out, zt


def qap_at_tau(cs, tau):
    """u_j(tau), v_j(tau), w_j(tau) for every variable j, with the n_in+1
    input-consistency rows appended to A."""
    log_m = cs.domain_log()
    L, zt = lagrange_at(tau, log_m)
    u = [0] * cs.n_vars
    v = [0] * cs.n_vars
    w = [0] * cs.n_vars
    for M, acc in ((cs.A, u), (cs.B, v), (cs.C, w)):
        rp, col, val = M
        for i in range(cs.n_cons):
            for k in range(rp[i], rp[i + 1]):
                acc[col[k]] = (acc[col[k]] + val[k] * L[i]) % R
    for j in range(cs.n_in + 1):
        u[j] = (u[j] + L[cs.n_cons + j]) % R
    return u, v, w, zt, log_m


def setup(cs, tau, alpha, beta, gamma, delta):
    u, v, w, zt, log_m = qap_at_tau(cs, tau)
    m = 1 << log_m
    pk = ProvingKey()
    pk.cs, pk.log_m = cs, log_m
    pk.alpha_g1, pk.beta_g1, pk.delta_g1 = g1_gen_mul(alpha), g1_gen_mul(beta), g1_gen_mul(delta)
    pk.beta_g2, pk.gamma_g2, pk.delta_g2 = g2_gen_mul(beta), g2_gen_mul(gamma), g2_gen_mul(delta)
    gi, di = fr_inv(gamma), fr_inv(delta)
    pk.a = [g1_gen_mul(x) for x in u]
    pk.b_g1 = [g1_gen_mul(x) for x in v]
    pk.b_g2 = [g2_gen_mul(x) for x in v]
    k = [(beta * u[j] + alpha * v[j] + w[j]) % R for j in range(cs.n_vars)]
    pk.ic = [g1_gen_mul(k[j] * gi % R) for j in range(cs.n_in + 1)]
    pk.l = [g1_gen_mul(k[j] * di % R) for j in range(cs.n_in + 1, cs.n_vars)]
    zd = zt * di % R
    pk.h, cur = [], 1
    for _ in range(m - 1):
        pk.h.append(g1_gen_mul(cur * zd % R))
        cur = cur * tau % R
    return pk


# ---------------------------------------------------------------------- prove
def abc_evals(cs, z, log_m):
    m = 1 << log_m
    a = R1CS.spmv(cs.A, z) + [0] * (m - cs.n_cons)
    b = R1CS.spmv(cs.B, z) + [0] * (m - cs.n_cons)
    c = R1CS.spmv(cs.C, z) + [0] * (m - cs.n_cons)
    for j in range(cs.n_in + 1):
        a[cs.n_cons + j] = z[j] % R
    return a, b, c


def h_from_abc(a, b, c):
    """H = (A*B - C)/Z as m coefficients (h[m-1] == 0).  witness_map steps 3."""
    m = len(a)
    ca, cb, cc = (coset_ntt(ntt(v, inverse=True)) for v in (a, b, c))
    zinv = fr_inv((pow(FR_GENERATOR, m, R) - 1) % R)
    hc = [(x * y - w) % R * zinv % R for x, y, w in zip(ca, cb, cc)]
    return coset_intt(hc)


class Trace:
    """Every intermediate the parity tests compare."""
    pass


def prove(pk, z, r, s):
    cs = pk.cs
    tr = Trace()
    a, b, c = abc_evals(cs, z, pk.log_m)
    tr.a, tr.b, tr.c = a, b, c
    h = h_from_abc(a, b, c)
    assert h[-1] == 0
    tr.h = h
    tr.msm_a = msm_g1(pk.a, z)
    tr.msm_b1 = msm_g1(pk.b_g1, z)
    tr.msm_b2 = msm_g2(pk.b_g2, z)
    tr.msm_l = msm_g1(pk.l, z[cs.n_in + 1:])
    tr.msm_h = msm_g1(pk.h, h[:-1])
    A = g1_add(g1_add(pk.alpha_g1, tr.msm_a), g1_mul(pk.delta_g1, r))
    B2 = g2_add(g2_add(pk.beta_g2, tr.msm_b2), g2_mul(pk.delta_g2, s))
    B1 = g1_add(g1_add(pk.beta_g1, tr.msm_b1), g1_mul(pk.delta_g1, s))
    C = g1_add(tr.msm_h, tr.msm_l)
    C = g1_add(C, g1_mul(A, s))
    C = g1_add(C, g1_mul(B1, r))
    C = g1_add(C, g1_neg(g1_mul(pk.delta_g1, r * s % R)))
    tr.A, tr.B, tr.C = A, B2, C
    tr.proof = g1_compress(A) + g2_compress(B2) + g1_compress(C)
    assert len(tr.proof) == 192
    return tr


def verify(pk, public_inputs, proof_bytes):
    """e(A,B) = e(alpha,beta) e(sum x_j IC_j, gamma) e(C,delta); x_0 = 1."""
    assert len(proof_bytes) == 192
    A = g1_decompress(proof_bytes[:48])
    B = g2_decompress(proof_bytes[48:144])
    C = g1_decompress(proof_bytes[144:])
    xs = [1] + [v % R for v in public_inputs]
    assert len(xs) == len(pk.ic)
    acc = msm_g1(pk.ic, xs)
    return pairing_product_is_one([
        (g1_neg(A), B), (pk.alpha_g1, pk.beta_g2), (acc, pk.gamma_g2), (C, pk.delta_g2)])
