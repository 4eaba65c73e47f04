"""ctypes binding of lib/libb2g16.so (the C ABI of include/b2g16.h).

This is the product path: every call goes to the CUDA library.  There is no CPU
fallback -- if the shared library is missing or no CUDA device is present the import /
Context() raises.  Host arrays are numpy uint8 in the ABI encodings (see codec.py);
device buffers are passed as raw pointers (e.g. torch uint8 CUDA tensors' data_ptr()).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_ROOT = os.path.dirname(os.path.dirname(_HERE))
SO_PATH = os.path.join(PKG_ROOT, "lib", "libb2g16.so")

ERRORS = {-1: "INVALID_ARG", -2: "NOT_CANONICAL", -3: "NOT_ON_CURVE", -4: "CUDA", -5: "NCCL", -6: "OOM", -7: "INTERNAL"}


class B2g16Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b2g16 error %s (%d): %s" % (ERRORS.get(code, "?"), code, msg))
        self.code = code


class Csr(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("nnz", C.c_uint64), ("row_ptr", C.c_void_p), ("col", C.c_void_p), ("val", C.c_void_p)]


class R1csView(C.Structure):
    _fields_ = [("log_m", C.c_uint32), ("n_cons", C.c_uint64), ("n_in", C.c_uint64), ("n_aux", C.c_uint64),
                ("A", Csr), ("B", Csr), ("C", Csr)]


class PkView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h")] + \
               [("cs", C.POINTER(R1csView))]


class Params(C.Structure):
    """b2g16_params of include/b2g16.h: a bellperson-layout parameter file in ABI encodings."""
    _fields_ = [("alpha_g1", C.c_uint8 * 96), ("beta_g1", C.c_uint8 * 96), ("beta_g2", C.c_uint8 * 192), ("gamma_g2", C.c_uint8 * 192),
                ("delta_g1", C.c_uint8 * 96), ("delta_g2", C.c_uint8 * 192)] + \
               [(n, C.c_uint64) for n in ("n_ic", "n_h", "n_l", "n_a", "n_b_g1", "n_b_g2")] + \
               [(n, C.c_void_p) for n in ("ic", "h", "l", "a", "b_g1", "b_g2")]


PARAM_ARRAYS = (("ic", 96), ("h", 96), ("l", 96), ("a", 96), ("b_g1", 96), ("b_g2", 192))
PARAM_POINTS = (("alpha_g1", 96), ("beta_g1", 96), ("beta_g2", 192), ("gamma_g2", 192), ("delta_g1", 96), ("delta_g2", 192))


def params_write(path, params):
    """params: dict of numpy uint8 arrays in ABI encodings, keys alpha_g1 .. delta_g2, ic, h, l, a, b_g1, b_g2."""
    st, keep = Params(), []
    for n, w in PARAM_POINTS:
        C.memmove(getattr(st, n), _p(_u8(params[n])), w)
    for n, w in PARAM_ARRAYS:
        arr = _u8(params[n], w)
        keep.append(arr)
        setattr(st, "n_" + n, arr.shape[0])
        setattr(st, n, _p(arr))
    rc = load().b2g16_params_write(os.fsencode(path), C.byref(st))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_params_write(%s)" % path)


def params_read(path):
    """-> dict of numpy arrays (copies; the library's buffers are freed before returning)."""
    h = C.POINTER(Params)()
    rc = load().b2g16_params_read(os.fsencode(path), C.byref(h))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_params_read(%s)" % path)
    st, out = h.contents, {}
    for n, w in PARAM_POINTS:
        out[n] = np.frombuffer(bytes(getattr(st, n)), dtype=np.uint8).copy()
    for n, w in PARAM_ARRAYS:
        cnt = int(getattr(st, "n_" + n))
        out[n] = np.ctypeslib.as_array(C.cast(getattr(st, n), C.POINTER(C.c_uint8)), shape=(cnt * w,)).reshape(cnt, w).copy() if cnt else np.zeros((0, w), np.uint8)
    vk = np.zeros(672 + 96 * int(st.n_ic), dtype=np.uint8)
    rc = load().b2g16_params_vk_bytes(h, _p(vk), C.c_size_t(vk.nbytes))
    load().b2g16_params_free(h)
    if rc != 0:
        raise B2g16Error(rc, "b2g16_params_vk_bytes")
    out["vk"] = vk
    return out


# ---- parameter-cache entries (S/core/parameter_cache.hpp:50-94,108-145,166-173) ----------------
def _cstr(fn, *args):
    buf = C.create_string_buffer(4096)
    rc = fn(*args, buf, C.c_size_t(len(buf)))
    if rc != 0:
        raise B2g16Error(rc, fn.__name__)
    return os.fsdecode(buf.value)


def sha256(data):
    out = (C.c_uint8 * 32)()
    rc = load().b2g16_sha256(bytes(data), C.c_size_t(len(data)), out)
    if rc != 0:
        raise B2g16Error(rc, "b2g16_sha256")
    return bytes(out)


def cache_dir():
    return _cstr(load().b2g16_cache_dir)


def cache_identifier(cache_prefix, parameter_set_identifier):
    return _cstr(load().b2g16_cache_identifier, os.fsencode(cache_prefix), os.fsencode(parameter_set_identifier))


CACHE_PARAMS, CACHE_VK, CACHE_META = 0, 1, 2


def cache_path(identifier, kind):
    return _cstr(load().b2g16_cache_path, os.fsencode(identifier), C.c_int(kind))


def meta_write(path, sector_size):
    rc = load().b2g16_meta_write(os.fsencode(path), C.c_uint64(sector_size))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_meta_write(%s)" % path)


def meta_read(path):
    v = C.c_uint64(0)
    rc = load().b2g16_meta_read(os.fsencode(path), C.byref(v))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_meta_read(%s)" % path)
    return int(v.value)


def vk_file_write(path, params):
    """params: dict with the six single points and ic (ABI encodings), as for params_write."""
    st = Params()
    for n, w in PARAM_POINTS:
        C.memmove(getattr(st, n), _p(_u8(params[n])), w)
    ic = _u8(params["ic"], 96)
    st.n_ic, st.ic = ic.shape[0], _p(ic)
    rc = load().b2g16_vk_file_write(os.fsencode(path), C.byref(st))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_vk_file_write(%s)" % path)


def vk_file_read(path):
    h = C.POINTER(Params)()
    rc = load().b2g16_vk_file_read(os.fsencode(path), C.byref(h))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_vk_file_read(%s)" % path)
    st, out = h.contents, {}
    for n, w in PARAM_POINTS:
        out[n] = np.frombuffer(bytes(getattr(st, n)), dtype=np.uint8).copy()
    cnt = int(st.n_ic)
    out["ic"] = np.ctypeslib.as_array(C.cast(st.ic, C.POINTER(C.c_uint8)), shape=(cnt * 96,)).reshape(cnt, 96).copy() if cnt else np.zeros((0, 96), np.uint8)
    load().b2g16_params_free(h)
    return out


def params_map(path):
    """b2g16_params_map -> {name: (count, byte offset)} for ic, h, l, a, b_g1, b_g2."""
    cnt, off = (C.c_uint64 * 6)(), (C.c_uint64 * 6)()
    rc = load().b2g16_params_map(os.fsencode(path), cnt, off)
    if rc != 0:
        raise B2g16Error(rc, "b2g16_params_map(%s)" % path)
    return {n: (int(cnt[i]), int(off[i])) for i, n in enumerate(("ic", "h", "l", "a", "b_g1", "b_g2"))}


def csr_density(col, n_vars, n_always=0):
    """b2g16_query_density: bitmap (uint8, n_vars) of the variables whose query point is present, given the matrix's CSR
    column indices; n_always = n_in + 1 for the a query (input-consistency rows), 0 for b_g1 / b_g2."""
    col = np.ascontiguousarray(col, dtype=np.uint32)
    m = Csr(0, col.shape[0], None, _p(col), None)
    out, cnt = np.zeros(n_vars, np.uint8), C.c_uint64()
    rc = load().b2g16_query_density(C.byref(m), C.c_uint64(n_vars), C.c_uint64(n_always), _p(out), C.byref(cnt))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_query_density: column index out of range")
    return out, int(cnt.value)


def query_expand(filtered, density, point_bytes):
    """b2g16_query_expand: bellman-style density-filtered query -> dense (n_vars, point_bytes), identity where density is 0."""
    filtered, density = _u8(filtered, point_bytes), _u8(density)
    out = np.empty((density.shape[0], point_bytes), np.uint8)
    rc = load().b2g16_query_expand(_p(filtered), C.c_uint64(filtered.shape[0]), _p(density), C.c_uint64(density.shape[0]),
                                   C.c_size_t(point_bytes), _p(out))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_query_expand: %d points for a bitmap with %d set" % (filtered.shape[0], int(np.count_nonzero(density))))
    return out


def query_filter(dense, point_bytes):
    """b2g16_query_filter: dense query -> (filtered points, density bitmap)."""
    dense = _u8(dense, point_bytes)
    n, cnt = dense.shape[0], C.c_uint64()
    dens, filt = np.zeros(n, np.uint8), np.empty((n, point_bytes), np.uint8)
    rc = load().b2g16_query_filter(_p(dense), C.c_uint64(n), C.c_size_t(point_bytes), _p(filt), _p(dens), C.byref(cnt))
    if rc != 0:
        raise B2g16Error(rc, "b2g16_query_filter")
    return filt[: int(cnt.value)].copy(), dens


def proof_decode(proof):
    a, b, c = np.zeros(96, np.uint8), np.zeros(192, np.uint8), np.zeros(96, np.uint8)
    rc = load().b2g16_proof_decode(_p(_u8(np.frombuffer(bytes(proof), dtype=np.uint8))), _p(a), _p(b), _p(c))
    if rc != 0:
        raise B2g16Error(rc, "malformed proof")
    return a, b, c


def proof_encode(a, b, c):
    out = np.zeros(192, np.uint8)
    rc = load().b2g16_proof_encode(_p(_u8(a)), _p(_u8(b)), _p(_u8(c)), _p(out))
    if rc != 0:
        raise B2g16Error(rc, "proof_encode: point not canonical / not on curve")
    return out.tobytes()


class Trace(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("a", "b", "c", "h", "msm_a", "msm_b1", "msm_l", "msm_h", "msm_b2", "A", "C", "B")] + \
               [("stage_ms", C.c_float * 8)]


_lib = None


def load():
    """dlopen the CUDA library (raises OSError if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise OSError("libb2g16.so not built: run `make -C %s` (or __graft_entry__.build())" % PKG_ROOT)
        _lib = C.CDLL(SO_PATH)
        _lib.b2g16_last_error.restype = C.c_char_p
        _lib.b2g16_launch_count.restype = C.c_uint64
        _lib.b2g16_bases_count.restype = C.c_size_t
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u8(a, width=None):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if width is not None:
        a = a.reshape(-1, width)
    return a


class Bases:
    """Device-resident affine points (proving-key query or synthetic set)."""

    def __init__(self, ctx, handle, group):
        self.ctx, self.h, self.group = ctx, handle, group

    def __len__(self):
        return load().b2g16_bases_count(self.h)

    def download(self, first=0, count=None):
        count = len(self) - first if count is None else count
        out = np.zeros((count, 96 if self.group == 1 else 192), dtype=np.uint8)
        self.ctx._chk(load().b2g16_bases_download(self.ctx.h, self.h, C.c_size_t(first), C.c_size_t(count), _p(out)))
        return out

    def free(self):
        if self.h:
            load().b2g16_bases_free(self.ctx.h, self.h)
            self.h = None


class ProvingKey:
    def __init__(self, ctx, handle, keepalive=None):
        self.ctx, self.h, self._keep = ctx, handle, keepalive

    def export(self, which, count):
        sizes = {0: 96, 1: 96, 2: 192, 3: 96, 4: 96, 5: 96, 6: 96, 7: 96, 8: 192, 9: 192}
        out = np.zeros((count, sizes[which]), dtype=np.uint8)
        self.ctx._chk(load().b2g16_pk_export(self.ctx.h, self.h, which, _p(out), C.c_size_t(out.nbytes)))
        return out

    def shape(self):
        out = (C.c_uint64 * 5)()
        self.ctx._chk(load().b2g16_pk_shape(self.h, out))
        return dict(zip(("log_m", "n_cons", "n_in", "n_aux", "n_vars"), [int(x) for x in out]))

    def free(self):
        if self.h:
            load().b2g16_pk_free(self.ctx.h, self.h)
            self.h = None


class PreparedVk:
    def __init__(self, ctx, handle, n_ic):
        self.ctx, self.h, self.n_ic = ctx, handle, n_ic

    def free(self):
        if self.h:
            load().b2g16_pvk_free(self.ctx.h, self.h)
            self.h = None


class Context:
    def __init__(self, device=0, stream=None):
        self.h = C.c_void_p()
        rc = load().b2g16_init(int(device), C.c_void_p(stream) if stream else None, C.byref(self.h))
        if rc != 0:
            raise B2g16Error(rc, "b2g16_init failed (no CUDA device / bad ordinal)")

    def close(self):
        if self.h:
            load().b2g16_shutdown(self.h)
            self.h = None

    def _chk(self, rc):
        if rc != 0:
            raise B2g16Error(rc, load().b2g16_last_error(self.h).decode())

    def synchronize(self):
        self._chk(load().b2g16_synchronize(self.h))

    @property
    def launches(self):
        return int(load().b2g16_launch_count(self.h))

    # ---- multiexp
    def msm_g1(self, bases, scalars):
        bases, scalars = _u8(bases, 96), _u8(scalars, 32)
        out = np.zeros(96, dtype=np.uint8)
        self._chk(load().b2g16_msm_g1(self.h, _p(bases), _p(scalars), C.c_size_t(scalars.shape[0]), _p(out)))
        return out

    def msm_g2(self, bases, scalars):
        bases, scalars = _u8(bases, 192), _u8(scalars, 32)
        out = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_msm_g2(self.h, _p(bases), _p(scalars), C.c_size_t(scalars.shape[0]), _p(out)))
        return out

    def bases_upload(self, group, pts):
        pts = _u8(pts, 96 if group == 1 else 192)
        h = C.c_void_p()
        fn = load().b2g16_bases_upload_g1 if group == 1 else load().b2g16_bases_upload_g2
        self._chk(fn(self.h, _p(pts), C.c_size_t(pts.shape[0]), C.byref(h)))
        return Bases(self, h, group)

    def bases_generate(self, group, d_scalars_ptr, n):
        h = C.c_void_p()
        self._chk(load().b2g16_bases_generate(self.h, int(group), C.c_void_p(d_scalars_ptr), C.c_size_t(n), C.byref(h)))
        return Bases(self, h, group)

    def msm_resident(self, bases, d_scalars_ptr, n, first=0):
        out = np.zeros(96 if bases.group == 1 else 192, dtype=np.uint8)
        self._chk(load().b2g16_msm_resident(self.h, bases.h, C.c_size_t(first), C.c_void_p(d_scalars_ptr), C.c_size_t(n), _p(out)))
        return out

    def msm_resident_async(self, bases, d_scalars_ptr, n, d_out_ptr, first=0):
        self._chk(load().b2g16_msm_resident_async(self.h, bases.h, C.c_size_t(first), C.c_void_p(d_scalars_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr)))

    def g1_sum(self, pts):
        pts = _u8(pts, 96)
        out = np.zeros(96, dtype=np.uint8)
        self._chk(load().b2g16_g1_sum(self.h, _p(pts), C.c_size_t(pts.shape[0]), _p(out)))
        return out

    def g2_sum(self, pts):
        pts = _u8(pts, 192)
        out = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_g2_sum(self.h, _p(pts), C.c_size_t(pts.shape[0]), _p(out)))
        return out

    # ---- evaluation domain
    def ntt(self, data, inverse=False, coset=False):
        d = _u8(data, 32).copy()
        n = d.shape[0]
        log_n = n.bit_length() - 1
        if n == 0 or (1 << log_n) != n:
            raise B2g16Error(-1, "domain size must be a power of two")
        self._chk(load().b2g16_ntt(self.h, _p(d), C.c_uint(log_n), int(inverse), int(coset)))
        return d

    def ntt_device(self, d_ptr, log_n, inverse=False, coset=False):
        self._chk(load().b2g16_ntt_device(self.h, C.c_void_p(d_ptr), C.c_uint(log_n), int(inverse), int(coset)))

    # per-rank steps of the transform sharded over several GPUs (dist.ntt_sharded)
    def ntt_cross_device(self, d_in, d_out, log_n, log_g, chunk, inverse=False, coset=False):
        self._chk(load().b2g16_ntt_cross_device(self.h, C.c_void_p(d_in), C.c_void_p(d_out), C.c_uint(log_n), C.c_uint(log_g), C.c_uint(chunk),
                                                int(inverse), int(coset)))

    def fr_coset_scale_device(self, d_ptr, count, log_n, offset, inverse=False):
        self._chk(load().b2g16_fr_coset_scale_device(self.h, C.c_void_p(d_ptr), C.c_size_t(count), C.c_uint(log_n), C.c_uint64(offset), int(inverse)))

    def fr_transpose_device(self, d_in, d_out, rows, cols):
        self._chk(load().b2g16_fr_transpose_device(self.h, C.c_void_p(d_in), C.c_void_p(d_out), C.c_size_t(rows), C.c_size_t(cols)))

    # peer memory (CUDA IPC) and the cross step fused with its exchanges
    def peer_alloc(self, nbytes):
        """-> (device pointer, 64-byte IPC handle) of a fresh whole allocation"""
        ptr, handle = C.c_void_p(), np.zeros(64, dtype=np.uint8)
        self._chk(load().b2g16_peer_alloc(self.h, C.c_size_t(nbytes), C.byref(ptr), _p(handle)))
        return int(ptr.value), handle.tobytes()

    def peer_open(self, handle):
        ptr = C.c_void_p()
        self._chk(load().b2g16_peer_open(self.h, _p(np.frombuffer(handle, dtype=np.uint8).copy()), C.byref(ptr)))
        return int(ptr.value)

    def peer_close(self, ptr):
        self._chk(load().b2g16_peer_close(self.h, C.c_void_p(ptr)))

    def peer_free(self, ptr):
        self._chk(load().b2g16_peer_free(self.h, C.c_void_p(ptr)))

    def ntt_cross_p2p_device(self, in_ptrs, out_ptrs, log_n, log_g, rank, inverse=False, coset=False):
        n = len(in_ptrs)
        a, b = (C.c_void_p * n)(*in_ptrs), (C.c_void_p * n)(*out_ptrs)
        self._chk(load().b2g16_ntt_cross_p2p_device(self.h, a, b, C.c_uint(log_n), C.c_uint(log_g), C.c_uint(rank), int(inverse), int(coset)))

    def h_from_abc(self, a, b, c):
        a, b, c = _u8(a, 32), _u8(b, 32), _u8(c, 32)
        m = a.shape[0]
        log_m = m.bit_length() - 1
        if m == 0 or (1 << log_m) != m or b.shape[0] != m or c.shape[0] != m:
            raise B2g16Error(-1, "a, b, c must have the same power-of-two length")
        out = np.zeros((m, 32), dtype=np.uint8)
        self._chk(load().b2g16_h_from_abc(self.h, _p(a), _p(b), _p(c), C.c_uint(log_m), _p(out)))
        return out

    # ---- parity / measurement aids
    def field_op(self, field, op, a, b=None):
        code = {"add": 0, "sub": 1, "mul": 2, "inv": 3, "sqr": 4, "neg": 5}[op]
        a = _u8(a, 32 if field == "fr" else 48)
        b = a if b is None else _u8(b, a.shape[1])
        out = np.empty_like(a)
        self._chk(load().b2g16_field_op(self.h, 0 if field == "fr" else 1, code, _p(a), _p(b), C.c_size_t(a.shape[0]), _p(out)))
        return out

    def fr_dot_device(self, d_a_ptr, d_b_ptr, n):
        out = np.zeros(32, dtype=np.uint8)
        self._chk(load().b2g16_fr_dot_device(self.h, C.c_void_p(d_a_ptr), C.c_void_p(d_b_ptr), C.c_size_t(n), _p(out)))
        return out

    def fr_dot_device_noalloc(self, d_a_ptr, d_b_ptr, n, d_out_ptr, d_status_ptr=None):
        """b2g16_fr_dot_device_noalloc: result (32 B) and status (int) are written to DEVICE memory; asynchronous."""
        self._chk(load().b2g16_fr_dot_device_noalloc(self.h, C.c_void_p(d_a_ptr), C.c_void_p(d_b_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr),
                                                     C.c_void_p(d_status_ptr) if d_status_ptr else None))

    def mad32_peak(self):
        v, ms = C.c_double(), C.c_double()
        self._chk(load().b2g16_measure_mad32_peak(self.h, C.byref(v), C.byref(ms)))
        return v.value, ms.value

    def set_msm_timing(self, on):
        self._chk(load().b2g16_set_msm_timing(self.h, int(on)))

    def set_msm_window(self, c):
        self._chk(load().b2g16_set_msm_window(self.h, int(c)))

    def set_tuning(self, key, value):
        self._chk(load().b2g16_set_tuning(self.h, key.encode(), int(value)))

    def kernel_log(self, max_entries=64):
        """[dict(group, c, n, ms, entries)] of the bucket-accumulation launches since set_tuning("kernel_log", 1)"""
        out = (C.c_double * (5 * max_entries))()
        n = load().b2g16_kernel_log(self.h, out, int(max_entries))
        if n < 0:
            self._chk(n)
        return [dict(group=int(out[5 * i]), c=int(out[5 * i + 1]), n=int(out[5 * i + 2]), ms=float(out[5 * i + 3]), entries=int(out[5 * i + 4]))
                for i in range(n)]

    def last_msm_timing(self):
        out = (C.c_float * 5)()
        self._chk(load().b2g16_last_msm_timing(self.h, out))
        return dict(zip(("digits_hist", "scan", "scatter", "accumulate", "reduce"), [float(x) for x in out]))

    # ---- prover
    def pk_load(self, key, cs):
        """key: object with numpy fields alpha_g1.. h (ABI encodings); cs: object with
        log_m, n_in, n_aux, n_cons and mats = [(row_ptr, col, val)] * 3."""
        view, keep = _pk_view(key, cs)
        h = C.c_void_p()
        self._chk(load().b2g16_pk_load(self.h, C.byref(view), C.byref(h)))
        return ProvingKey(self, h, keep)

    def pk_load_file(self, path, cs, want_vk=False):
        """b2g16_pk_load_file: `.params` file streamed to the device (pinned ring, device-side conversion / expansion)."""
        r1, keep = _r1cs_view(cs)
        h = C.c_void_p()
        vk = np.zeros(672 + 96 * (cs.n_in + 1), dtype=np.uint8) if want_vk else None
        self._chk(load().b2g16_pk_load_file(self.h, os.fsencode(path), C.byref(r1), C.byref(h), _p(vk), C.c_size_t(vk.nbytes if want_vk else 0)))
        pk = ProvingKey(self, h, keep)
        return (pk, vk) if want_vk else pk

    def pk_generate(self, cs, toxic, want_vk=False):
        r1, keep = _r1cs_view(cs)
        h = C.c_void_p()
        vk = np.zeros(96 + 3 * 192 + 96 * (cs.n_in + 1), dtype=np.uint8) if want_vk else None
        self._chk(load().b2g16_pk_generate(self.h, C.byref(r1), _p(_u8(toxic)), C.byref(h), _p(vk)))
        pk = ProvingKey(self, h, keep)
        return (pk, vk) if want_vk else pk

    def pk_generate_synth(self, log_m, n_in, n_free, seed, toxic, want_vk=False):
        h = C.c_void_p()
        vk = np.zeros(96 + 3 * 192 + 96 * (n_in + 1), dtype=np.uint8) if want_vk else None
        self._chk(load().b2g16_pk_generate_synth(self.h, C.c_uint32(log_m), C.c_uint64(n_in), C.c_uint64(n_free), C.c_uint64(seed),
                                                 _p(_u8(toxic)), C.byref(h), _p(vk)))
        pk = ProvingKey(self, h)
        return (pk, vk) if want_vk else pk

    def prove(self, pk, assignment, r, s, priority=False, trace_m=None):
        z = _u8(assignment, 32)
        proof = np.zeros(192, dtype=np.uint8)
        tr, bufs = None, {}
        if trace_m is not None:
            tr = Trace()
            sizes = dict(a=(trace_m, 32), b=(trace_m, 32), c=(trace_m, 32), h=(trace_m, 32), msm_a=(96,), msm_b1=(96,), msm_l=(96,),
                         msm_h=(96,), msm_b2=(192,), A=(96,), C=(96,), B=(192,))
            for n, shp in sizes.items():
                bufs[n] = np.zeros(shp, dtype=np.uint8)
                setattr(tr, n, _p(bufs[n]))
        self._chk(load().b2g16_prove(self.h, pk.h, _p(z), _p(_u8(r)), _p(_u8(s)), int(priority), _p(proof), C.byref(tr) if tr else None))
        if tr is not None:
            bufs["stage_ms"] = [float(x) for x in tr.stage_ms]
            return proof.tobytes(), bufs
        return proof.tobytes()

    def prove_abc(self, pk, a, b, c, assignment, r, s, priority=False):
        """prove from caller-supplied a, b, c evaluations (2^log_m canonical scalars each)"""
        proof = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_prove_abc(self.h, pk.h, _p(_u8(a, 32)), _p(_u8(b, 32)), _p(_u8(c, 32)), _p(_u8(assignment, 32)),
                                         _p(_u8(r)), _p(_u8(s)), int(priority), _p(proof), None))
        return proof.tobytes()

    # ---- verifier
    def vk_prepare(self, vk):
        vk = _u8(vk)
        n_ic = (vk.size - 672) // 96
        if vk.size < 768 or 672 + 96 * n_ic != vk.size:
            raise B2g16Error(-1, "vk must be 672 + 96 * n_ic bytes")
        h = C.c_void_p()
        self._chk(load().b2g16_vk_prepare(self.h, _p(vk), C.c_size_t(n_ic), C.byref(h)))
        return PreparedVk(self, h, n_ic)

    def verify(self, pvk, inputs, proof):
        x = _u8(inputs, 32)
        ok = C.c_int(0)
        self._chk(load().b2g16_verify(self.h, pvk.h, _p(x), C.c_size_t(x.shape[0]), _p(_u8(np.frombuffer(bytes(proof), dtype=np.uint8))), C.byref(ok)))
        return bool(ok.value)

    def verify_batch(self, pvk, inputs, proofs, rand=None):
        """inputs: (n_proofs, n_in, 32); proofs: 192 * n_proofs bytes; rand: (n_proofs, 32) or None"""
        pr = _u8(np.frombuffer(bytes(proofs), dtype=np.uint8))
        n = pr.size // 192
        x = _u8(inputs)
        n_in = pvk.n_ic - 1
        if x.size != n * n_in * 32:
            raise B2g16Error(-1, "inputs must hold n_proofs * n_in scalars")
        ok = C.c_int(0)
        self._chk(load().b2g16_verify_batch(self.h, pvk.h, _p(x), C.c_size_t(n_in), _p(pr), C.c_size_t(n), _p(_u8(rand)) if rand is not None else None, C.byref(ok)))
        return bool(ok.value)

    def prove_device(self, pk, d_assignment_ptr, r, s, priority=False):
        proof = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_prove_device(self.h, pk.h, C.c_void_p(d_assignment_ptr), _p(_u8(r)), _p(_u8(s)), int(priority), _p(proof), None))
        return proof.tobytes()

    def prove_partial(self, pk, d_assignment_ptr, part, parts):
        out = np.zeros(576, dtype=np.uint8)
        self._chk(load().b2g16_prove_partial(self.h, pk.h, C.c_void_p(d_assignment_ptr), C.c_uint(part), C.c_uint(parts), _p(out)))
        return out

    def abc_evals_strided_device(self, pk, d_assignment_ptr, row0, stride, count, d_a, d_b, d_c):
        self._chk(load().b2g16_abc_evals_strided_device(self.h, pk.h, C.c_void_p(d_assignment_ptr), C.c_uint64(row0), C.c_uint64(stride), C.c_uint64(count),
                                                        C.c_void_p(d_a), C.c_void_p(d_b), C.c_void_p(d_c)))

    def h_pointwise_device(self, d_a, d_b, d_c, count, log_m, d_out):
        self._chk(load().b2g16_h_pointwise_device(self.h, C.c_void_p(d_a), C.c_void_p(d_b), C.c_void_p(d_c), C.c_uint64(count), C.c_uint(log_m), C.c_void_p(d_out)))

    def prove_partial_h(self, pk, d_assignment_ptr, part, parts, d_h_ptr, h_first, h_count):
        out = np.zeros(576, dtype=np.uint8)
        self._chk(load().b2g16_prove_partial_h(self.h, pk.h, C.c_void_p(d_assignment_ptr), C.c_uint(part), C.c_uint(parts), C.c_void_p(d_h_ptr),
                                               C.c_uint64(h_first), C.c_uint64(h_count), _p(out)))
        return out

    def prove_assemble(self, pk, partials, r, s):
        p = _u8(partials, 576)
        proof = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_prove_assemble(self.h, pk.h, _p(p), C.c_uint(p.shape[0]), _p(_u8(r)), _p(_u8(s)), _p(proof)))
        return proof.tobytes()

    def prove_batch(self, pk, assignments, rs, ss, priority=False):
        z = _u8(assignments)
        n = _u8(rs, 32).shape[0]
        proofs = np.zeros(192 * n, dtype=np.uint8)
        self._chk(load().b2g16_prove_batch(self.h, pk.h, _p(z), C.c_size_t(n), _p(_u8(rs)), _p(_u8(ss)), int(priority), _p(proofs)))
        return proofs.tobytes()

    def synth_witness(self, pk, d_assignment_ptr, n_free):
        self._chk(load().b2g16_synth_witness(self.h, pk.h, C.c_void_p(d_assignment_ptr), C.c_uint64(n_free)))


class MultiContext:
    """b2g16_init_multi: several devices of one box driven from this process (one context + host thread per device)."""

    def __init__(self, devices):
        devs = (C.c_int * len(devices))(*devices)
        self.h = C.c_void_p()
        rc = load().b2g16_init_multi(devs, len(devices), C.byref(self.h))
        if rc != 0:
            raise B2g16Error(rc, "b2g16_init_multi(%s)" % list(devices))
        self.n = len(devices)

    def _chk(self, rc):
        if rc != 0:
            load().b2g16_multi_last_error.restype = C.c_char_p
            raise B2g16Error(rc, (load().b2g16_multi_last_error(self.h) or b"").decode())

    def close(self):
        if self.h:
            load().b2g16_shutdown_multi(self.h)
            self.h = None

    def key_generate_synth(self, log_m, n_in, n_free, seed, toxic, want_vk=False):
        k = C.c_void_p()
        vk = np.zeros(672 + 96 * (n_in + 1), dtype=np.uint8) if want_vk else None
        self._chk(load().b2g16_mpk_generate_synth(self.h, C.c_uint32(log_m), C.c_uint64(n_in), C.c_uint64(n_free), C.c_uint64(seed), _p(_u8(toxic)),
                                                 C.byref(k), _p(vk)))
        return (k, vk) if want_vk else k

    def key_load(self, key, cs):
        v, keep = _pk_view(key, cs)
        k = C.c_void_p()
        self._chk(load().b2g16_mpk_load(self.h, C.byref(v), C.byref(k)))
        return k

    def key_free(self, k):
        load().b2g16_mpk_free(self.h, k)

    def prove(self, k, assignment, r, s):
        out = np.zeros(192, dtype=np.uint8)
        self._chk(load().b2g16_prove_multi(self.h, k, _p(_u8(assignment)), _p(_u8(r)), _p(_u8(s)), 0, _p(out)))
        return bytes(out)

    def prove_batch(self, k, assignments, rs, ss):
        a = _u8(assignments)
        n = len(rs)
        out = np.zeros(192 * n, dtype=np.uint8)
        self._chk(load().b2g16_prove_batch_multi(self.h, k, _p(a), C.c_size_t(n), _p(_u8(rs)), _p(_u8(ss)), 0, _p(out)))
        return [bytes(out[192 * i:192 * (i + 1)]) for i in range(n)]


def _r1cs_view(cs):
    keep = []
    v = R1csView()
    v.log_m, v.n_cons, v.n_in, v.n_aux = cs.log_m, cs.n_cons, cs.n_in, cs.n_aux
    if getattr(cs, "mats", None) is None:        # shape only: the caller proves with prove_abc
        return v, keep
    for name, (rp, col, val) in zip("ABC", cs.mats):
        rp = np.ascontiguousarray(rp, dtype=np.uint32)
        col = np.ascontiguousarray(col, dtype=np.uint32)
        val = _u8(val, 32)
        keep += [rp, col, val]
        m = Csr()
        m.n_rows, m.nnz = cs.n_cons, col.shape[0]
        m.row_ptr, m.col, m.val = _p(rp), _p(col), _p(val)
        setattr(v, name, m)
    return v, keep


def _pk_view(key, cs):
    r1, keep = _r1cs_view(cs)
    keep.append(r1)
    v = PkView()
    for n in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "a", "b_g1", "b_g2", "l", "h"):
        arr = _u8(getattr(key, n))
        keep.append(arr)
        setattr(v, n, _p(arr))
    v.cs = C.pointer(r1)
    return v, keep
