"""Slow, obviously-correct BLS12-381 arithmetic in Python big ints.

TEST INFRASTRUCTURE ONLY (oracle).  Nothing in the product path
(`crypto3-fil-proofs_b200/`) may import this module; only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may.

Parity status: "parity unpinned" at byte level -- the reference tree holds no
MSM / FFT / proof golden vectors.  What the reference *does*
pin near this path and what this module is checked against:
  * Fr canonical 32-byte LE encoding KATs
    (libs/storage/test/core/fr32.cpp:51-85, libs/filecoin/test/api/mod.cpp:35-43)
  * proof length 192 bytes (libs/filecoin/include/nil/filecoin/proofs/constants.hpp:93)
  * prove -> verify == true round trips (libs/storage/test/**/compound.cpp)
The arithmetic the reference delegates to (un-vendored, absent from
/root/reference): crypto3-algebra @ b2e1a199 (fields, curves, multiexp),
crypto3-math @ 16c55f5c (evaluation_domain), zk @ 44a19899 (r1cs_gg_ppzksnark).
Their published algorithms (Groth16, radix-2 NTT, short-Weierstrass group law,
optimal-ate pairing) are restated here from the mathematical definitions.

Constants are *derived* from the BLS parameter x and asserted, not trusted.
"""

# ----------------------------------------------------------------- constants
X_PARAM = -0xd201000000010000
R = X_PARAM**4 - X_PARAM**2 + 1                      # scalar field modulus
P = (X_PARAM - 1) ** 2 * R // 3 + X_PARAM            # base field modulus
assert R == 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
assert P == 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
FR_GENERATOR = 7            # multiplicative generator of Fr
FR_TWO_ADICITY = 32
assert (R - 1) % (1 << FR_TWO_ADICITY) == 0 and ((R - 1) >> FR_TWO_ADICITY) & 1 == 1
assert pow(FR_GENERATOR, (R - 1) // 2, R) == R - 1   # 7 is a non-residue

G1_GEN = (
    0x17f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb,
    0x08b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1,
)
G2_GEN = (
    (0x024aa2b2f08f0a91260805272dc51051c6e47ad4fa403b02b4510b647ae3d1770bac0326a805bbefd48056c8c121bdb8,
     0x13e02b6052719f607dacd3a088274f65596bd0d09920b61ab5da61bbdc7f5049334cf11213945d57e5ac7d055d042b7e),
    (0x0ce5d527727d6e118cc9cdc6da2e351aadfd9baa8cbdd3a76d429a695160d12c923ac9cc3baca289e193548608b82801,
     0x0606c4a02ea734cc32acd2b02bc28b99cb3e287e85a763af267492ab572e99ab3f370d275cec1da1aaa9075ff05f79be),
)


# --------------------------------------------------------------------- Fr
def fr_inv(a):
    return pow(a, R - 2, R)


def fr_root_of_unity(log_n):
    """Primitive 2^log_n-th root of unity: 7^((r-1)/2^log_n)."""
    assert 0 <= log_n <= FR_TWO_ADICITY
    return pow(FR_GENERATOR, (R - 1) >> log_n, R)


def fr_from_bytes_le(b):
    """fr32 contract: 32 LE bytes, canonical (< r) else None.
    (libs/storage/include/nil/filecoin/storage/proofs/core/fr32.hpp:36-52)"""
    assert len(b) == 32
    v = int.from_bytes(b, "little")
    return v if v < R else None


def fr_to_bytes_le(v):
    return int(v % R).to_bytes(32, "little")


# --------------------------------------------------------------------- Fq2
# element = (c0, c1) meaning c0 + c1*u, u^2 = -1
def fq2_add(a, b):
    return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)


def fq2_sub(a, b):
    return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)


def fq2_neg(a):
    return ((-a[0]) % P, (-a[1]) % P)


def fq2_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def fq2_inv(a):
    n = pow(a[0] * a[0] + a[1] * a[1], P - 2, P)
    return (a[0] * n % P, (-a[1]) * n % P)


def fq2_muls(a, k):
    return (a[0] * k % P, a[1] * k % P)


FQ2_ZERO = (0, 0)
FQ2_ONE = (1, 0)


# ------------------------------------------------- generic affine curve ops
class Field:
    """Tiny vtable so G1 (Fq) and G2 (Fq2) share one group law."""

    def __init__(self, add, sub, mul, inv, neg, zero, one, muls):
        self.add, self.sub, self.mul, self.inv, self.neg = add, sub, mul, inv, neg
        self.zero, self.one, self.muls = zero, one, muls


FQ = Field(lambda a, b: (a + b) % P, lambda a, b: (a - b) % P, lambda a, b: a * b % P,
           lambda a: pow(a, P - 2, P), lambda a: (-a) % P, 0, 1, lambda a, k: a * k % P)
FQ2 = Field(fq2_add, fq2_sub, fq2_mul, fq2_inv, fq2_neg, FQ2_ZERO, FQ2_ONE, fq2_muls)
B1 = 4
B2 = (4, 4)     # 4*(1+u)


def on_curve(F, pt, b):
    if pt is None:
        return True
    x, y = pt
    return F.mul(y, y) == F.add(F.mul(F.mul(x, x), x), b)


def pt_neg(F, p):
    return None if p is None else (p[0], F.neg(p[1]))


def pt_add(F, p, q):
    """Affine chord-and-tangent; None is the identity."""
    if p is None:
        return q
    if q is None:
        return p
    x1, y1 = p
    x2, y2 = q
    if x1 == x2:
        if y1 == y2 and y1 != F.zero:
            lam = F.mul(F.muls(F.mul(x1, x1), 3), F.inv(F.muls(y1, 2)))
        else:
            return None
    else:
        lam = F.mul(F.sub(y2, y1), F.inv(F.sub(x2, x1)))
    x3 = F.sub(F.sub(F.mul(lam, lam), x1), x2)
    y3 = F.sub(F.mul(lam, F.sub(x1, x3)), y1)
    return (x3, y3)


def pt_mul(F, p, k):
    k = int(k)
    if k < 0:
        return pt_mul(F, pt_neg(F, p), -k)
    acc = None
    while k:
        if k & 1:
            acc = pt_add(F, acc, p)
        p = pt_add(F, p, p)
        k >>= 1
    return acc


def g1_add(p, q): return pt_add(FQ, p, q)
def g1_mul(p, k): return pt_mul(FQ, p, k % R)
def g1_neg(p): return pt_neg(FQ, p)
def g2_add(p, q): return pt_add(FQ2, p, q)
def g2_mul(p, k): return pt_mul(FQ2, p, k % R)
def g2_neg(p): return pt_neg(FQ2, p)


assert on_curve(FQ, G1_GEN, B1) and on_curve(FQ2, G2_GEN, B2)


def check_generators_order():
    """Order-r check (slow-ish, called from tests, not at import)."""
    return pt_mul(FQ, G1_GEN, R) is None and pt_mul(FQ2, G2_GEN, R) is None


# Jacobian-free fast path for building many multiples k*G: use projective adds
# via Python ints would still be slow; tests only need <= 2^12 points, so a
# simple windowed fixed-base table is enough.
class FixedBase:
    def __init__(self, F, base, window=8, bits=255):
        self.F, self.w = F, window
        self.tables = []
        cur = base
        for _ in range((bits + window - 1) // window):
            row = [None]
            acc = None
            for _ in range((1 << window) - 1):
                acc = pt_add(F, acc, cur)
                row.append(acc)
            self.tables.append(row)
            cur = pt_add(F, acc, cur)      # 2^w * cur_prev

    def mul(self, k):
        k %= R
        acc = None
        i = 0
        mask = (1 << self.w) - 1
        while k:
            d = k & mask
            if d:
                acc = pt_add(self.F, acc, self.tables[i][d])
            k >>= self.w
            i += 1
        return acc


_fb = {}


def g1_gen_mul(k):
    if "g1" not in _fb:
        _fb["g1"] = FixedBase(FQ, G1_GEN, 6)
    return _fb["g1"].mul(k)


def g2_gen_mul(k):
    if "g2" not in _fb:
        _fb["g2"] = FixedBase(FQ2, G2_GEN, 6)
    return _fb["g2"].mul(k)


# ----------------------------------------------------------- naive MSMs
def msm(F, bases, scalars):
    acc = None
    for b, s in zip(bases, scalars):
        s %= R
        if s and b is not None:
            acc = pt_add(F, acc, pt_mul(F, b, s))
    return acc


def msm_g1(bases, scalars): return msm(FQ, bases, scalars)
def msm_g2(bases, scalars): return msm(FQ2, bases, scalars)


# ------------------------------------------------------------ ABI encodings
# (include/b2g16.h): G1 affine = x||y 48-byte LE canonical, infinity all-zero;
# G2 affine = x.c0||x.c1||y.c0||y.c1.
def g1_to_abi(p):
    if p is None:
        return bytes(96)
    return p[0].to_bytes(48, "little") + p[1].to_bytes(48, "little")


def g1_from_abi(b):
    assert len(b) == 96
    x = int.from_bytes(b[:48], "little")
    y = int.from_bytes(b[48:], "little")
    return None if (x == 0 and y == 0) else (x, y)


def g2_to_abi(p):
    if p is None:
        return bytes(192)
    (x0, x1), (y0, y1) = p
    return b"".join(v.to_bytes(48, "little") for v in (x0, x1, y0, y1))


def g2_from_abi(b):
    assert len(b) == 192
    v = [int.from_bytes(b[i * 48:(i + 1) * 48], "little") for i in range(4)]
    return None if not any(v) else ((v[0], v[1]), (v[2], v[3]))


# zcash/bellman compressed encodings used for the 192-byte proof
# (constants.hpp:93 pins 48+96+48).
def g1_compress(p):
    if p is None:
        return bytes([0xC0]) + bytes(47)
    x, y = p
    b = bytearray(x.to_bytes(48, "big"))
    b[0] |= 0x80
    if y > (P - 1) // 2:
        b[0] |= 0x20
    return bytes(b)


def _fq2_lex_larger(y):
    ny = fq2_neg(y)
    return (y[1], y[0]) > (ny[1], ny[0])


def g2_compress(p):
    if p is None:
        return bytes([0xC0]) + bytes(95)
    (x0, x1), y = p
    b = bytearray(x1.to_bytes(48, "big") + x0.to_bytes(48, "big"))
    b[0] |= 0x80
    if _fq2_lex_larger(y):
        b[0] |= 0x20
    return bytes(b)


def fq_sqrt(a):
    # p = 3 mod 4
    s = pow(a, (P + 1) // 4, P)
    return s if s * s % P == a % P else None


def fq2_sqrt(a):
    """Square root in Fq2 = Fq[u]/(u^2+1), p = 3 mod 4 (complex method)."""
    if a == FQ2_ZERO:
        return FQ2_ZERO
    a0, a1 = a
    if a1 == 0:
        s = fq_sqrt(a0)
        if s is not None:
            return (s, 0)
        s = fq_sqrt((-a0) % P)
        return (0, s)
    n = fq_sqrt((a0 * a0 + a1 * a1) % P)
    if n is None:
        return None
    inv2 = pow(2, P - 2, P)
    for nn in (n, (-n) % P):
        t = (a0 + nn) * inv2 % P
        x0 = fq_sqrt(t)
        if x0 is None or x0 == 0:
            continue
        x1 = a1 * pow(2 * x0, P - 2, P) % P
        if fq2_mul((x0, x1), (x0, x1)) == (a0 % P, a1 % P):
            return (x0, x1)
    return None


def g1_decompress(b):
    assert len(b) == 48 and b[0] & 0x80
    if b[0] & 0x40:
        return None
    sign = bool(b[0] & 0x20)
    x = int.from_bytes(bytes([b[0] & 0x1F]) + b[1:], "big")
    y = fq_sqrt((x * x * x + 4) % P)
    assert y is not None
    if (y > (P - 1) // 2) != sign:
        y = P - y
    return (x, y)


def g2_decompress(b):
    assert len(b) == 96 and b[0] & 0x80
    if b[0] & 0x40:
        return None
    sign = bool(b[0] & 0x20)
    x1 = int.from_bytes(bytes([b[0] & 0x1F]) + b[1:48], "big")
    x0 = int.from_bytes(b[48:], "big")
    x = (x0, x1)
    y = fq2_sqrt(fq2_add(fq2_mul(fq2_mul(x, x), x), B2))
    assert y is not None
    if _fq2_lex_larger(y) != sign:
        y = fq2_neg(y)
    return (x, y)


# ------------------------------------------------------------------ pairing
# Fq12 = Fq[w]/(w^12 - 2 w^6 + 2); Fq2 embeds via u = w^6 - 1 (then u^2 = -1).
# The sextic twist E'(Fq2): y^2 = x^3 + 4(1+u) maps to E(Fq12) by
# (x, y) -> (x / w^2, y / w^3) because w^6 = 1 + u.
class Fq12:
    __slots__ = ("c",)
    DEG = 12

    def __init__(self, c):
        self.c = [v % P for v in c]

    @staticmethod
    def one():
        return Fq12([1] + [0] * 11)

    @staticmethod
    def from_fq(a):
        return Fq12([a] + [0] * 11)

    @staticmethod
    def from_fq2(a):
        # c0 + c1*u = (c0 - c1) + c1*w^6
        c = [0] * 12
        c[0] = a[0] - a[1]
        c[6] = a[1]
        return Fq12(c)

    def __add__(self, o): return Fq12([a + b for a, b in zip(self.c, o.c)])
    def __sub__(self, o): return Fq12([a - b for a, b in zip(self.c, o.c)])
    def __neg__(self): return Fq12([-a for a in self.c])
    def __eq__(self, o): return self.c == o.c

    def __mul__(self, o):
        if isinstance(o, int):
            return Fq12([a * o for a in self.c])
        t = [0] * 23
        for i, a in enumerate(self.c):
            if a:
                for j, b in enumerate(o.c):
                    t[i + j] += a * b
        # reduce w^12 = 2 w^6 - 2
        for k in range(22, 11, -1):
            v = t[k]
            if v:
                t[k - 6] += 2 * v
                t[k - 12] -= 2 * v
        return Fq12(t[:12])

    def pow(self, e):
        res, base = Fq12.one(), self
        while e:
            if e & 1:
                res = res * base
            base = base * base
            e >>= 1
        return res

    def inv(self):
        # Fermat in Fq12 would be huge; use extended Euclid on polynomials.
        lm, hm = [1] + [0] * 12, [0] * 13
        low = self.c + [0]
        high = [2, 0, 0, 0, 0, 0, P - 2, 0, 0, 0, 0, 0, 1]

        def deg(p):
            d = len(p) - 1
            while d and p[d] == 0:
                d -= 1
            return d

        def poly_div(a, b):
            da, db = deg(a), deg(b)
            tmp = list(a)
            o = [0] * len(a)
            ib = pow(b[db], P - 2, P)
            for i in range(da - db, -1, -1):
                o[i] = (o[i] + tmp[db + i] * ib) % P
                for c in range(db + 1):
                    tmp[c + i] = (tmp[c + i] - o[i] * b[c]) % P
            return o[: deg(o) + 1]

        while deg(low):
            r = poly_div(high, low)
            r += [0] * (13 - len(r))
            nm, new = list(hm), list(high)
            for i in range(13):
                for j in range(13 - i):
                    nm[i + j] -= lm[i] * r[j]
                    new[i + j] -= low[i] * r[j]
            nm = [x % P for x in nm]
            new = [x % P for x in new]
            lm, low, hm, high = nm, new, lm, low
        il = pow(low[0], P - 2, P)
        return Fq12([x * il for x in lm[:12]])


_W = Fq12([0, 1] + [0] * 10)
_W2_INV = (_W * _W).inv()
_W3_INV = (_W * _W * _W).inv()


def _twist(q):
    x, y = q
    return (Fq12.from_fq2(x) * _W2_INV, Fq12.from_fq2(y) * _W3_INV)


def _line(p1, p2, t):
    """Evaluate the line through p1,p2 (Fq12-affine) at t."""
    x1, y1 = p1
    x2, y2 = p2
    xt, yt = t
    if x1 != x2:
        m = (y2 - y1) * (x2 - x1).inv()
        return m * (xt - x1) - (yt - y1)
    if y1 == y2:
        m = (x1 * x1 * 3) * (y1 * 2).inv()
        return m * (xt - x1) - (yt - y1)
    return xt - x1


def _add12(p, q):
    x1, y1 = p
    x2, y2 = q
    if x1 == x2 and y1 == y2:
        m = (x1 * x1 * 3) * (y1 * 2).inv()
    else:
        m = (y2 - y1) * (x2 - x1).inv()
    x3 = m * m - x1 - x2
    return (x3, m * (x1 - x3) - y1)


def miller_loop(p_g1, q_g2):
    """f_{|x|,Q}(P), Q twisted into E(Fq12).  Returns Fq12 (not yet final-exp'd)."""
    if p_g1 is None or q_g2 is None:
        return Fq12.one()
    q = _twist(q_g2)
    pt = (Fq12.from_fq(p_g1[0]), Fq12.from_fq(p_g1[1]))
    t = q
    f = Fq12.one()
    n = -X_PARAM
    for i in range(n.bit_length() - 2, -1, -1):
        f = f * f * _line(t, t, pt)
        t = _add12(t, t)
        if (n >> i) & 1:
            f = f * _line(t, q, pt)
            t = _add12(t, q)
    return f


FINAL_EXP = (P**12 - 1) // R


def final_exponentiation(f):
    return f.pow(FINAL_EXP)


def pairing(p_g1, q_g2):
    """A bilinear non-degenerate map G1 x G2 -> GT (the ate pairing for |x|;
    it is the inverse of the x<0 optimal-ate value, which is irrelevant for
    product-of-pairings equality checks)."""
    return final_exponentiation(miller_loop(p_g1, q_g2))


def pairing_product_is_one(pairs):
    f = Fq12.one()
    for p, q in pairs:
        f = f * miller_loop(p, q)
    return final_exponentiation(f) == Fq12.one()
