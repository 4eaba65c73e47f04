"""Byte codecs of the C ABI (include/b2g16.h) on numpy arrays.

Fr  : 32 B little-endian canonical  (fr32 contract of the reference:
      libs/storage/include/nil/filecoin/storage/proofs/core/fr32.hpp:36-52)
G1  : x || y, 48 B LE each; identity = 96 zero bytes
G2  : x.c0 || x.c1 || y.c0 || y.c1, 48 B LE each; identity = 192 zero bytes
"""
import numpy as np

R_MOD = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
_R_LIMBS = np.array([(R_MOD >> (64 * i)) & (2**64 - 1) for i in range(4)], dtype=np.uint64)


def fr_from_ints(vals):
    out = np.zeros((len(vals), 32), dtype=np.uint8)
    for i, v in enumerate(vals):
        out[i] = np.frombuffer(int(v).to_bytes(32, "little"), dtype=np.uint8)
    return out


def fr_to_ints(arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint8).reshape(-1, 32)
    return [int.from_bytes(arr[i].tobytes(), "little") for i in range(arr.shape[0])]


def g1_from_points(pts):
    out = np.zeros((len(pts), 96), dtype=np.uint8)
    for i, p in enumerate(pts):
        if p is not None:
            out[i] = np.frombuffer(p[0].to_bytes(48, "little") + p[1].to_bytes(48, "little"), dtype=np.uint8)
    return out


def g1_to_point(b):
    b = bytes(np.asarray(b, dtype=np.uint8).tobytes())
    x, y = int.from_bytes(b[:48], "little"), int.from_bytes(b[48:96], "little")
    return None if x == 0 and y == 0 else (x, y)


def g2_from_points(pts):
    out = np.zeros((len(pts), 192), dtype=np.uint8)
    for i, p in enumerate(pts):
        if p is not None:
            (x0, x1), (y0, y1) = p
            out[i] = np.frombuffer(b"".join(v.to_bytes(48, "little") for v in (x0, x1, y0, y1)), dtype=np.uint8)
    return out


def g2_to_point(b):
    b = bytes(np.asarray(b, dtype=np.uint8).tobytes())
    v = [int.from_bytes(b[i * 48:(i + 1) * 48], "little") for i in range(4)]
    return None if not any(v) else ((v[0], v[1]), (v[2], v[3]))


def random_fr(n, rng):
    """n uniform canonical Fr as uint8[n,32]; rng = numpy Generator.  Drawn as
    4 x u64 masked to 255 bits with rejection of values >= r."""
    out = np.empty((n, 4), dtype=np.uint64)
    todo = np.arange(n)
    while todo.size:
        draw = rng.integers(0, 2**64, size=(todo.size, 4), dtype=np.uint64)
        draw[:, 3] &= np.uint64(0x7FFFFFFFFFFFFFFF)
        # lexicographic compare with r from the top limb down
        lt = np.zeros(todo.size, dtype=bool)
        eq = np.ones(todo.size, dtype=bool)
        for k in (3, 2, 1, 0):
            lt |= eq & (draw[:, k] < _R_LIMBS[k])
            eq &= draw[:, k] == _R_LIMBS[k]
        out[todo[lt]] = draw[lt]
        todo = todo[~lt]
    return out.view(np.uint8).reshape(n, 32)
