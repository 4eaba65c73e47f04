"""ctypes binding of oracle/_build/liboracle.so (the fast CPU restatement).

TEST INFRASTRUCTURE ONLY -- see oracle/c/oracle.cpp header.  Byte formats are the
ones of include/b2g16.h: Fr = 32 B LE canonical, G1 affine = 96 B, G2 affine = 192 B.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")


def build(force=False):
    src = os.path.join(_HERE, "c", "oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


class Csr(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("nnz", C.c_uint64), ("row_ptr", C.c_void_p),
                ("col", C.c_void_p), ("val", C.c_void_p)]


class R1cs(C.Structure):
    _fields_ = [("log_m", C.c_uint32), ("n_cons", C.c_uint64), ("n_in", C.c_uint64),
                ("n_aux", C.c_uint64), ("A", Csr), ("B", Csr), ("C", Csr)]


class Pk(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h")]


class PkOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "gamma_g2", "delta_g2", "ic",
                 "a", "b_g1", "b_g2", "l", "h")]


class Trace(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("a", "b", "c", "h", "msm_a", "msm_b1", "msm_l", "msm_h", "msm_b2", "A", "C", "B")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


def threads():
    return lib().orc_threads()


def set_threads(t):
    lib().orc_set_threads(int(t))


def fr_is_canonical(b32):
    return bool(lib().orc_fr_is_canonical(_p(_u8(np.frombuffer(bytes(b32), dtype=np.uint8)))))


def field_op(field, op, a, b=None):
    """field 'fr'|'fq'; op add sub mul inv sqr neg; a,b uint8 arrays (n, 32|48)."""
    code = {"add": 0, "sub": 1, "mul": 2, "inv": 3, "sqr": 4, "neg": 5}[op]
    a = _u8(a)
    n = a.shape[0]
    out = np.empty_like(a)
    b = _u8(b) if b is not None else a
    fn = lib().orc_fr_op if field == "fr" else lib().orc_fq_op
    rc = fn(code, _p(a), _p(b), C.c_size_t(n), _p(out))
    assert rc == 0, rc
    return out


def ntt(data, inverse=False, coset=False):
    d = _u8(data).copy()
    n = d.shape[0]
    log_n = n.bit_length() - 1
    assert 1 << log_n == n
    rc = lib().orc_ntt(_p(d), C.c_uint(log_n), int(inverse), int(coset))
    assert rc == 0, rc
    return d


def msm_g1(bases, scalars):
    bases, scalars = _u8(bases), _u8(scalars)
    out = np.zeros(96, dtype=np.uint8)
    rc = lib().orc_msm_g1(_p(bases), _p(scalars), C.c_size_t(scalars.shape[0]), _p(out))
    assert rc == 0, rc
    return out


def msm_g2(bases, scalars):
    bases, scalars = _u8(bases), _u8(scalars)
    out = np.zeros(192, dtype=np.uint8)
    rc = lib().orc_msm_g2(_p(bases), _p(scalars), C.c_size_t(scalars.shape[0]), _p(out))
    assert rc == 0, rc
    return out


def g1_gen_mul(k):
    k = _u8(k)
    out = np.zeros((k.shape[0], 96), dtype=np.uint8)
    assert lib().orc_g1_gen_mul(_p(k), C.c_size_t(k.shape[0]), _p(out)) == 0
    return out


def g2_gen_mul(k):
    k = _u8(k)
    out = np.zeros((k.shape[0], 192), dtype=np.uint8)
    assert lib().orc_g2_gen_mul(_p(k), C.c_size_t(k.shape[0]), _p(out)) == 0
    return out


def _pt2(fn, a, b, size):
    out = np.zeros(size, dtype=np.uint8)
    assert fn(_p(_u8(a)), _p(_u8(b)), _p(out)) == 0
    return out


def g1_add(a, b): return _pt2(lib().orc_g1_add, a, b, 96)
def g2_add(a, b): return _pt2(lib().orc_g2_add, a, b, 192)
def g1_mul(a, k): return _pt2(lib().orc_g1_mul, a, k, 96)
def g2_mul(a, k): return _pt2(lib().orc_g2_mul, a, k, 192)
def g1_on_curve(a): return bool(lib().orc_g1_on_curve(_p(_u8(a))))
def g2_on_curve(a): return bool(lib().orc_g2_on_curve(_p(_u8(a))))


def g1_compress(a):
    out = np.zeros(48, dtype=np.uint8)
    assert lib().orc_g1_compress(_p(_u8(a)), _p(out)) == 0
    return out.tobytes()


def g2_compress(a):
    out = np.zeros(96, dtype=np.uint8)
    assert lib().orc_g2_compress(_p(_u8(a)), _p(out)) == 0
    return out.tobytes()


class R1csArrays:
    """numpy-backed R1CS: keeps arrays alive and exposes the C struct.
    A/B/C = (row_ptr uint32[n_cons+1], col uint32[nnz], val uint8[nnz,32])."""

    def __init__(self, log_m, n_in, n_aux, A, B, Cm):
        self.log_m, self.n_in, self.n_aux = log_m, n_in, n_aux
        self.mats = []
        for (rp, col, val) in (A, B, Cm):
            self.mats.append((np.ascontiguousarray(rp, dtype=np.uint32),
                              np.ascontiguousarray(col, dtype=np.uint32),
                              np.ascontiguousarray(val, dtype=np.uint8).reshape(-1, 32)))
        self.n_cons = self.mats[0][0].shape[0] - 1
        self.n_vars = 1 + n_in + n_aux

    def struct(self, cls_r1cs=R1cs, cls_csr=Csr):
        s = cls_r1cs()
        s.log_m, s.n_cons, s.n_in, s.n_aux = self.log_m, self.n_cons, self.n_in, self.n_aux
        for name, (rp, col, val) in zip("ABC", self.mats):
            c = cls_csr()
            c.n_rows, c.nnz = self.n_cons, col.shape[0]
            c.row_ptr, c.col, c.val = _p(rp), _p(col), _p(val)
            setattr(s, name, c)
        return s


def spmv(mat, z):
    rp, col, val = mat
    c = Csr()
    c.n_rows, c.nnz = rp.shape[0] - 1, col.shape[0]
    c.row_ptr, c.col, c.val = _p(rp), _p(col), _p(val)
    z = _u8(z)
    out = np.zeros((rp.shape[0] - 1, 32), dtype=np.uint8)
    assert lib().orc_spmv(C.byref(c), _p(z), C.c_size_t(z.shape[0]), _p(out)) == 0
    return out


def h_from_abc(a, b, c):
    a, b, c = _u8(a), _u8(b), _u8(c)
    m = a.shape[0]
    out = np.zeros((m, 32), dtype=np.uint8)
    assert lib().orc_h_from_abc(_p(a), _p(b), _p(c), C.c_uint(m.bit_length() - 1), _p(out)) == 0
    return out


def abc_evals(cs, z):
    m = 1 << cs.log_m
    a, b, c = (np.zeros((m, 32), dtype=np.uint8) for _ in range(3))
    s = cs.struct()
    assert lib().orc_abc_evals(C.byref(s), _p(_u8(z)), _p(a), _p(b), _p(c)) == 0
    return a, b, c


def synth_witness(cs, z_head, n_free):
    z = np.zeros((cs.n_vars, 32), dtype=np.uint8)
    z[: z_head.shape[0]] = z_head
    s = cs.struct()
    assert lib().orc_synth_witness(C.byref(s), _p(z), C.c_uint64(n_free)) == 0
    return z


class Key:
    """Dense Groth16 key as numpy byte arrays (ABI encodings)."""
    FIELDS = (("alpha_g1", 96, None), ("beta_g1", 96, None), ("delta_g1", 96, None),
              ("beta_g2", 192, None), ("gamma_g2", 192, None), ("delta_g2", 192, None))

    def __init__(self, cs):
        m = 1 << cs.log_m
        self.cs = cs
        for name, sz, _ in self.FIELDS:
            setattr(self, name, np.zeros(sz, dtype=np.uint8))
        self.ic = np.zeros((cs.n_in + 1, 96), dtype=np.uint8)
        self.a = np.zeros((cs.n_vars, 96), dtype=np.uint8)
        self.b_g1 = np.zeros((cs.n_vars, 96), dtype=np.uint8)
        self.b_g2 = np.zeros((cs.n_vars, 192), dtype=np.uint8)
        self.l = np.zeros((cs.n_aux, 96), dtype=np.uint8)
        self.h = np.zeros((m - 1, 96), dtype=np.uint8)

    def pk_struct(self):
        s = Pk()
        for n in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h"):
            setattr(s, n, _p(getattr(self, n)))
        return s


def groth16_setup(cs, toxic):
    """toxic: uint8[5,32] = tau, alpha, beta, gamma, delta."""
    key = Key(cs)
    o = PkOut()
    for n, _ in PkOut._fields_:
        setattr(o, n, _p(getattr(key, n)))
    s = cs.struct()
    rc = lib().orc_groth16_setup(C.byref(s), _p(_u8(toxic)), C.byref(o))
    assert rc == 0, rc
    return key


def groth16_prove(key, z, r, s, abc=None, want_trace=True):
    cs = key.cs
    m = 1 << cs.log_m
    proof = np.zeros(192, dtype=np.uint8)
    tr = Trace()
    bufs = {}
    if want_trace:
        sizes = dict(a=(m, 32), b=(m, 32), c=(m, 32), h=(m, 32), msm_a=(96,), msm_b1=(96,), msm_l=(96,),
                     msm_h=(96,), msm_b2=(192,), A=(96,), C=(96,), B=(192,))
        for n, shp in sizes.items():
            bufs[n] = np.zeros(shp, dtype=np.uint8)
            setattr(tr, n, _p(bufs[n]))
    pk = key.pk_struct()
    st = cs.struct()
    rc = lib().orc_groth16_prove(C.byref(pk), C.byref(st), _p(_u8(abc)) if abc is not None else None,
                                 _p(_u8(z)), _p(_u8(r)), _p(_u8(s)), _p(proof), C.byref(tr))
    assert rc == 0, rc
    return proof.tobytes(), bufs
