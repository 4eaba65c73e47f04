// Host-side BLS12-381 arithmetic for the parts of the path that are O(1) per proof and run on
// the CPU by design: the Groth16 verifier's pairing check, point (de)compression of the
// 192-byte proof and of parameter files.  The reference reaches these through
//   verify_seal -> StackedCompound::verify   (libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:339-395,
//   libs/storage/.../core/proof/compound_proof.hpp:97-102 -- a stub; the arithmetic lives in the
//   un-vendored crypto3-algebra `pairing/bls12.hpp`, included at libs/filecoin/.../caches.hpp:30-31).
// Not a port of any of them: 6x64-bit Montgomery limbs with unsigned __int128, the textbook tower
// Fq2 = Fq[u]/(u^2+1), Fq6 = Fq2[v]/(v^3-(1+u)), Fq12 = Fq6[w]/(w^2-v), an affine Miller loop on the
// sextic twist for |x| and a final exponentiation as conj(f)/f raised to (p^2+1)(p^4-p^2+1)/r.
// Constants are checked against Python big-int arithmetic by tests/test_verify_host.py.
#pragma once
#include <stdint.h>
#include <string.h>
#include <vector>

namespace b2h {

typedef unsigned __int128 u128;

static const uint64_t FQ_P[6] = {0xb9feffffffffaaabull, 0x1eabfffeb153ffffull, 0x6730d2a0f6b0f624ull,
                                 0x64774b84f38512bfull, 0x4b1ba7b6434bacd7ull, 0x1a0111ea397fe69aull};
static const uint64_t FQ_INV = 0x89f3fffcfffcfffdull;            // -p^-1 mod 2^64
static const uint64_t FQ_ONE[6] = {0x760900000002fffdull, 0xebf4000bc40c0002ull, 0x5f48985753c758baull,
                                   0x77ce585370525745ull, 0x5c071a97a256ec6dull, 0x15f65ec3fa80e493ull};   // 2^384 mod p
static const uint64_t FQ_R2[6] = {0xf4df1f341c341746ull, 0x0a76e6a609d104f1ull, 0x8de5476c4c95b6d5ull,
                                  0x67eb88a9939d83c0ull, 0x9a793e85b519952dull, 0x11988fe592cae3aaull};    // 2^768 mod p
static const uint64_t FQ_PM2[6] = {0xb9feffffffffaaa9ull, 0x1eabfffeb153ffffull, 0x6730d2a0f6b0f624ull,
                                   0x64774b84f38512bfull, 0x4b1ba7b6434bacd7ull, 0x1a0111ea397fe69aull};   // p - 2
static const uint64_t FQ_SQRT_EXP[6] = {0xee7fbfffffffeaabull, 0x07aaffffac54ffffull, 0xd9cc34a83dac3d89ull,
                                        0xd91dd2e13ce144afull, 0x92c6e9ed90d2eb35ull, 0x0680447a8e5ff9a6ull};  // (p+1)/4
static const uint64_t FQ_HALF[6] = {0xdcff7fffffffd555ull, 0x0f55ffff58a9ffffull, 0xb39869507b587b12ull,
                                    0xb23ba5c279c2895full, 0x258dd3db21a5d66bull, 0x0d0088f51cbff34dull};  // (p-1)/2
static const uint64_t FR_MOD[4] = {0xffffffff00000001ull, 0x53bda402fffe5bfeull, 0x3339d80809a1d805ull, 0x73eda753299d7d48ull};
static const uint64_t BLS_X_ABS = 0xd201000000010000ull;         // |x|, x < 0
// (p^2+1)(p^4-p^2+1)/r : the part of the final exponent left after f -> conj(f)/f = f^(p^6-1)
static const uint64_t FINAL_EXP_TAIL[32] = {
    0x8739e1cdc0705d6aull, 0x09a5256de0381a16ull, 0x9cf0f70a61c791e2ull, 0x3a09c4497903f76eull, 0x2d7271563890f133ull,
    0x224741b36fec7760ull, 0x338259c22a12bd40ull, 0x38ee1cd4778e0de7ull, 0xc3b5ef4b188a20b0ull, 0x1d615d49e2764d7bull,
    0x816101ddd076117dull, 0xf007c01e7ebe3afcull, 0x27d7bd90935021c3ull, 0xc3b5e2f557c0b15full, 0x5e886c94c4f82384ull,
    0xee6a95db11e63f56ull, 0x2b822f514a9c4f6full, 0x12d6a874d21b73daull, 0x1304275ef499dffbull, 0x967878febcb95d1full,
    0x4744497f8b2f2922ull, 0x85a2e707f0841855ull, 0x9f0c50126c802eecull, 0xfb46e197bd2fa489ull, 0x548ce0809bc5f61aull,
    0xcf56fb1573beaa8cull, 0xad7375a3763bdf7cull, 0xe0ec9031179bdeccull, 0x6579aea83c48c1daull, 0xdbf85ae664cf5bb3ull,
    0x7b6f235c55ca7566ull, 0x000028b314877503ull};

// ------------------------------------------------------------------------------- Fq
struct Fq {
    uint64_t v[6];
    static Fq zero() { Fq r; memset(r.v, 0, sizeof r.v); return r; }
    static Fq one() { Fq r; memcpy(r.v, FQ_ONE, sizeof r.v); return r; }
    bool is_zero() const { uint64_t t = 0; for (int i = 0; i < 6; ++i) t |= v[i]; return t == 0; }
    bool operator==(const Fq& o) const { return memcmp(v, o.v, sizeof v) == 0; }
    bool operator!=(const Fq& o) const { return !(*this == o); }
    static bool geq(const uint64_t* a, const uint64_t* b) {          // a >= b on raw limbs
        for (int i = 5; i >= 0; --i) if (a[i] != b[i]) return a[i] > b[i];
        return true;
    }
    static void sub_raw(uint64_t* r, const uint64_t* a, const uint64_t* b) {
        u128 borrow = 0;
        for (int i = 0; i < 6; ++i) { u128 t = (u128)a[i] - b[i] - borrow; r[i] = (uint64_t)t; borrow = (t >> 64) & 1; }
    }
    Fq operator+(const Fq& o) const {
        Fq r; u128 c = 0;
        for (int i = 0; i < 6; ++i) { c += (u128)v[i] + o.v[i]; r.v[i] = (uint64_t)c; c >>= 64; }
        if (geq(r.v, FQ_P)) sub_raw(r.v, r.v, FQ_P);                  // p < 2^381: no carry out of 384 bits
        return r;
    }
    Fq operator-(const Fq& o) const {
        Fq r;
        if (geq(v, o.v)) { sub_raw(r.v, v, o.v); return r; }
        uint64_t t[6]; sub_raw(t, o.v, v); sub_raw(r.v, FQ_P, t); return r;
    }
    Fq neg() const { return is_zero() ? *this : zero() - *this; }
    Fq dbl() const { return *this + *this; }
    Fq operator*(const Fq& o) const {                                 // CIOS Montgomery product
        uint64_t t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 6; ++i) {
            u128 c = 0;
            for (int j = 0; j < 6; ++j) { c += (u128)v[j] * o.v[i] + t[j]; t[j] = (uint64_t)c; c >>= 64; }
            c += t[6]; t[6] = (uint64_t)c; t[7] = (uint64_t)(c >> 64);
            const uint64_t m = t[0] * FQ_INV;
            c = ((u128)m * FQ_P[0] + t[0]) >> 64;
            for (int j = 1; j < 6; ++j) { c += (u128)m * FQ_P[j] + t[j]; t[j - 1] = (uint64_t)c; c >>= 64; }
            c += t[6]; t[5] = (uint64_t)c; t[6] = t[7] + (uint64_t)(c >> 64);
        }
        Fq r; memcpy(r.v, t, sizeof r.v);
        if (t[6] || geq(r.v, FQ_P)) sub_raw(r.v, r.v, FQ_P);
        return r;
    }
    Fq sqr() const { return *this * *this; }
    Fq pow(const uint64_t* e, int n) const {
        Fq r = one();
        for (int i = 64 * n - 1; i >= 0; --i) { r = r.sqr(); if ((e[i >> 6] >> (i & 63)) & 1) r = r * *this; }
        return r;
    }
    Fq inv() const { return pow(FQ_PM2, 6); }                         // inv(0) = 0
    // canonical little-endian 48 bytes (the C ABI's coordinate format) <-> Montgomery; false if >= p
    static bool from_le48(const uint8_t* b, Fq* out) {
        Fq c; memcpy(c.v, b, 48);
        if (geq(c.v, FQ_P)) return false;
        Fq r2; memcpy(r2.v, FQ_R2, sizeof r2.v);
        *out = c * r2; return true;
    }
    void to_le48(uint8_t* b) const { Fq o = zero(); o.v[0] = 1; Fq c = *this * o; memcpy(b, c.v, 48); }
    static bool from_be48(const uint8_t* b, Fq* out) { uint8_t le[48]; for (int i = 0; i < 48; ++i) le[i] = b[47 - i]; return from_le48(le, out); }
    void to_be48(uint8_t* b) const { uint8_t le[48]; to_le48(le); for (int i = 0; i < 48; ++i) b[i] = le[47 - i]; }
    bool larger_half() const {                                        // canonical value > (p-1)/2
        uint8_t le[48]; to_le48(le); uint64_t c[6]; memcpy(c, le, 48);
        return !geq(FQ_HALF, c);
    }
    bool sqrt(Fq* out) const {                                        // p = 3 mod 4
        Fq s = pow(FQ_SQRT_EXP, 6);
        if (s.sqr() != *this) return false;
        *out = s; return true;
    }
    static Fq small(unsigned k) { Fq r = zero(), o = one(); for (unsigned i = 0; i < k; ++i) r = r + o; return r; }
};

// ------------------------------------------------------------------------------ Fq2
struct Fq2 {
    Fq c0, c1;
    static Fq2 zero() { return {Fq::zero(), Fq::zero()}; }
    static Fq2 one() { return {Fq::one(), Fq::zero()}; }
    bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    bool operator==(const Fq2& o) const { return c0 == o.c0 && c1 == o.c1; }
    bool operator!=(const Fq2& o) const { return !(*this == o); }
    Fq2 operator+(const Fq2& o) const { return {c0 + o.c0, c1 + o.c1}; }
    Fq2 operator-(const Fq2& o) const { return {c0 - o.c0, c1 - o.c1}; }
    Fq2 neg() const { return {c0.neg(), c1.neg()}; }
    Fq2 dbl() const { return {c0.dbl(), c1.dbl()}; }
    Fq2 conj() const { return {c0, c1.neg()}; }
    Fq2 operator*(const Fq2& o) const {
        Fq ac = c0 * o.c0, bd = c1 * o.c1, t = (c0 + c1) * (o.c0 + o.c1);
        return {ac - bd, t - ac - bd};
    }
    Fq2 sqr() const { Fq ab = c0 * c1; return {(c0 + c1) * (c0 - c1), ab.dbl()}; }
    Fq2 scale(const Fq& k) const { return {c0 * k, c1 * k}; }
    Fq2 mul_xi() const { return {c0 - c1, c0 + c1}; }                 // * (1 + u)
    Fq2 inv() const { Fq n = (c0.sqr() + c1.sqr()).inv(); return {c0 * n, (c1 * n).neg()}; }
    bool lex_larger() const { return c1.is_zero() ? c0.larger_half() : c1.larger_half(); }   // compare c1 first, then c0
    bool sqrt(Fq2* out) const {                                       // complex method, p = 3 mod 4
        if (is_zero()) { *out = zero(); return true; }
        if (c1.is_zero()) {
            Fq s;
            if (c0.sqrt(&s)) { *out = {s, Fq::zero()}; return true; }
            if (c0.neg().sqrt(&s)) { *out = {Fq::zero(), s}; return true; }
            return false;
        }
        Fq n;
        if (!(c0.sqr() + c1.sqr()).sqrt(&n)) return false;
        const Fq inv2 = Fq::small(2).inv();
        for (int k = 0; k < 2; ++k) {
            Fq t = (c0 + (k ? n.neg() : n)) * inv2, x0;
            if (!t.sqrt(&x0) || x0.is_zero()) continue;
            Fq x1 = c1 * x0.dbl().inv();
            Fq2 cand = {x0, x1};
            if (cand.sqr() == *this) { *out = cand; return true; }
        }
        return false;
    }
};

// ------------------------------------------------------------------------ Fq6, Fq12
struct Fq6 {                                                          // a0 + a1 v + a2 v^2, v^3 = 1 + u
    Fq2 a0, a1, a2;
    static Fq6 zero() { return {Fq2::zero(), Fq2::zero(), Fq2::zero()}; }
    static Fq6 one() { return {Fq2::one(), Fq2::zero(), Fq2::zero()}; }
    bool operator==(const Fq6& o) const { return a0 == o.a0 && a1 == o.a1 && a2 == o.a2; }
    Fq6 operator+(const Fq6& o) const { return {a0 + o.a0, a1 + o.a1, a2 + o.a2}; }
    Fq6 operator-(const Fq6& o) const { return {a0 - o.a0, a1 - o.a1, a2 - o.a2}; }
    Fq6 neg() const { return {a0.neg(), a1.neg(), a2.neg()}; }
    Fq6 operator*(const Fq6& o) const {
        Fq2 t00 = a0 * o.a0, t11 = a1 * o.a1, t22 = a2 * o.a2;
        Fq2 c0 = t00 + ((a1 + a2) * (o.a1 + o.a2) - t11 - t22).mul_xi();
        Fq2 c1 = (a0 + a1) * (o.a0 + o.a1) - t00 - t11 + t22.mul_xi();
        Fq2 c2 = (a0 + a2) * (o.a0 + o.a2) - t00 - t22 + t11;
        return {c0, c1, c2};
    }
    Fq6 mul_v() const { return {a2.mul_xi(), a0, a1}; }
    Fq6 inv() const {
        Fq2 t0 = a0.sqr() - (a1 * a2).mul_xi();
        Fq2 t1 = a2.sqr().mul_xi() - a0 * a1;
        Fq2 t2 = a1.sqr() - a0 * a2;
        Fq2 d = (a0 * t0 + (a2 * t1 + a1 * t2).mul_xi()).inv();
        return {t0 * d, t1 * d, t2 * d};
    }
};

struct Fq12 {                                                         // c0 + c1 w, w^2 = v
    Fq6 c0, c1;
    static Fq12 one() { return {Fq6::one(), Fq6::zero()}; }
    bool operator==(const Fq12& o) const { return c0 == o.c0 && c1 == o.c1; }
    Fq12 operator*(const Fq12& o) const {
        Fq6 aa = c0 * o.c0, bb = c1 * o.c1;
        return {aa + bb.mul_v(), (c0 + c1) * (o.c0 + o.c1) - aa - bb};
    }
    Fq12 sqr() const { return *this * *this; }
    Fq12 conj() const { return {c0, c1.neg()}; }                      // = f^(p^6)
    Fq12 inv() const {
        Fq6 d = (c0 * c0 - (c1 * c1).mul_v()).inv();
        return {c0 * d, (c1 * d).neg()};
    }
    Fq12 pow(const uint64_t* e, int n) const {
        Fq12 r = one(); bool started = false;
        for (int i = 64 * n - 1; i >= 0; --i) {
            if (started) r = r.sqr();
            if ((e[i >> 6] >> (i & 63)) & 1) { r = started ? r * *this : *this; started = true; }
        }
        return r;
    }
};

// --------------------------------------------------------------------- curve points
// affine (x, y) with `inf`; XYZZ accumulator (formulas: SURVEY App. E) for scalar products
template <class F> struct Aff {
    F x, y; bool inf;
    static Aff identity() { return {F::zero(), F::zero(), true}; }
    Aff neg() const { return {x, y.neg(), inf}; }
    bool operator==(const Aff& o) const { return inf == o.inf && (inf || (x == o.x && y == o.y)); }
};
template <class F> struct Pt {
    F X, Y, ZZ, ZZZ;
    static Pt identity() { return {F::zero(), F::zero(), F::zero(), F::zero()}; }
    static Pt from(const Aff<F>& a) { return a.inf ? identity() : Pt{a.x, a.y, F::one(), F::one()}; }
    bool is_identity() const { return ZZ.is_zero(); }
    Pt dbl() const {
        if (is_identity() || Y.is_zero()) return identity();
        F U = Y.dbl(), V = U.sqr(), W = U * V, S = X * V, X2 = X.sqr(), M = X2.dbl() + X2;
        F X3 = M.sqr() - S.dbl();
        return {X3, M * (S - X3) - W * Y, V * ZZ, W * ZZZ};
    }
    Pt add(const Pt& o) const {
        if (o.is_identity()) return *this;
        if (is_identity()) return o;
        F U1 = X * o.ZZ, U2 = o.X * ZZ, S1 = Y * o.ZZZ, S2 = o.Y * ZZZ, P = U2 - U1, R = S2 - S1;
        if (P.is_zero()) return R.is_zero() ? dbl() : identity();
        F PP = P.sqr(), PPP = P * PP, Q = U1 * PP, X3 = R.sqr() - PPP - Q.dbl();
        return {X3, R * (Q - X3) - S1 * PPP, ZZ * o.ZZ * PP, ZZZ * o.ZZZ * PPP};
    }
    Pt mul(const uint64_t* k, int n) const {                          // MSB-first double-and-add
        Pt r = identity();
        for (int i = 64 * n - 1; i >= 0; --i) { r = r.dbl(); if ((k[i >> 6] >> (i & 63)) & 1) r = r.add(*this); }
        return r;
    }
    Aff<F> affine() const {
        if (is_identity()) return Aff<F>::identity();
        F I = (ZZ * ZZZ).inv();
        return {X * I * ZZZ, Y * I * ZZ, false};
    }
};
typedef Aff<Fq> G1;
typedef Aff<Fq2> G2;

inline Fq g1_b() { return Fq::small(4); }
inline Fq2 g2_b() { return {Fq::small(4), Fq::small(4)}; }             // 4 (1 + u)
inline bool on_curve(const G1& p) { return p.inf || p.y.sqr() == p.x.sqr() * p.x + g1_b(); }
inline bool on_curve(const G2& p) { return p.inf || p.y.sqr() == p.x.sqr() * p.x + g2_b(); }
template <class F> inline bool in_subgroup(const Aff<F>& p) { return p.inf || Pt<F>::from(p).mul(FR_MOD, 4).is_identity(); }

// ---- C-ABI encodings: little-endian canonical coordinates, identity = all zero (include/b2g16.h)
inline bool g1_from_abi(const uint8_t* b, G1* o) {
    bool z = true; for (int i = 0; i < 96; ++i) z = z && b[i] == 0;
    if (z) { *o = G1::identity(); return true; }
    o->inf = false; return Fq::from_le48(b, &o->x) && Fq::from_le48(b + 48, &o->y);
}
inline bool g2_from_abi(const uint8_t* b, G2* o) {
    bool z = true; for (int i = 0; i < 192; ++i) z = z && b[i] == 0;
    if (z) { *o = G2::identity(); return true; }
    o->inf = false;
    return Fq::from_le48(b, &o->x.c0) && Fq::from_le48(b + 48, &o->x.c1) && Fq::from_le48(b + 96, &o->y.c0) && Fq::from_le48(b + 144, &o->y.c1);
}
inline void g1_to_abi(const G1& p, uint8_t* b) { if (p.inf) { memset(b, 0, 96); return; } p.x.to_le48(b); p.y.to_le48(b + 48); }
inline void g2_to_abi(const G2& p, uint8_t* b) {
    if (p.inf) { memset(b, 0, 192); return; }
    p.x.c0.to_le48(b); p.x.c1.to_le48(b + 48); p.y.c0.to_le48(b + 96); p.y.c1.to_le48(b + 144);
}

// ---- zcash/bellman encodings: big-endian, flag bits in byte 0:
//      bit7 compressed, bit6 infinity, bit5 y is the lexicographically larger root.
//      G2 coordinates are written c1 then c0.
inline void g1_compress(const G1& p, uint8_t* o) {
    if (p.inf) { memset(o, 0, 48); o[0] = 0xC0; return; }
    p.x.to_be48(o); o[0] |= 0x80; if (p.y.larger_half()) o[0] |= 0x20;
}
inline void g2_compress(const G2& p, uint8_t* o) {
    if (p.inf) { memset(o, 0, 96); o[0] = 0xC0; return; }
    p.x.c1.to_be48(o); p.x.c0.to_be48(o + 48); o[0] |= 0x80; if (p.y.lex_larger()) o[0] |= 0x20;
}
inline bool rest_zero(const uint8_t* b, int n) { if (b[0] & 0x3F) return false; for (int i = 1; i < n; ++i) if (b[i]) return false; return true; }
inline bool g1_decompress(const uint8_t* b, G1* o) {
    if (!(b[0] & 0x80)) return false;
    if (b[0] & 0x40) { if (!rest_zero(b, 48)) return false; *o = G1::identity(); return true; }
    uint8_t t[48]; memcpy(t, b, 48); t[0] &= 0x1F;
    Fq x, y;
    if (!Fq::from_be48(t, &x) || !(x.sqr() * x + g1_b()).sqrt(&y)) return false;
    if (y.larger_half() != (bool)(b[0] & 0x20)) y = y.neg();
    *o = {x, y, false}; return true;
}
inline bool g2_decompress(const uint8_t* b, G2* o) {
    if (!(b[0] & 0x80)) return false;
    if (b[0] & 0x40) { if (!rest_zero(b, 96)) return false; *o = G2::identity(); return true; }
    uint8_t t[48]; memcpy(t, b, 48); t[0] &= 0x1F;
    Fq2 x, y;
    if (!Fq::from_be48(t, &x.c1) || !Fq::from_be48(b + 48, &x.c0) || !(x.sqr() * x + g2_b()).sqrt(&y)) return false;
    if (y.lex_larger() != (bool)(b[0] & 0x20)) y = y.neg();
    *o = {x, y, false}; return true;
}
// uncompressed (parameter files): G1 96 B = x | y, G2 192 B = x.c1 x.c0 y.c1 y.c0; bit7 = 0, bit6 = infinity
inline void g1_write_uncompressed(const G1& p, uint8_t* o) { if (p.inf) { memset(o, 0, 96); o[0] = 0x40; return; } p.x.to_be48(o); p.y.to_be48(o + 48); }
inline void g2_write_uncompressed(const G2& p, uint8_t* o) {
    if (p.inf) { memset(o, 0, 192); o[0] = 0x40; return; }
    p.x.c1.to_be48(o); p.x.c0.to_be48(o + 48); p.y.c1.to_be48(o + 96); p.y.c0.to_be48(o + 144);
}
inline bool g1_read_uncompressed(const uint8_t* b, G1* o) {
    if (b[0] & 0x80) return false;
    if (b[0] & 0x40) { if (!rest_zero(b, 96)) return false; *o = G1::identity(); return true; }
    o->inf = false; return Fq::from_be48(b, &o->x) && Fq::from_be48(b + 48, &o->y);
}
inline bool g2_read_uncompressed(const uint8_t* b, G2* o) {
    if (b[0] & 0x80) return false;
    if (b[0] & 0x40) { if (!rest_zero(b, 192)) return false; *o = G2::identity(); return true; }
    o->inf = false;
    return Fq::from_be48(b, &o->x.c1) && Fq::from_be48(b + 48, &o->x.c0) && Fq::from_be48(b + 96, &o->y.c1) && Fq::from_be48(b + 144, &o->y.c0);
}

// -------------------------------------------------------------------------- pairing
// Untwist E'(Fq2) -> E(Fq12): (x', y') -> (x'/w^2, y'/w^3), w^6 = 1 + u.  The line through the
// untwisted T with twist-slope lambda, evaluated at P = (xp, yp) in E(Fq) and scaled by w^3 (an
// element of the proper subfield Fq4, killed by the final exponentiation), is
//     (yt - lambda xt)  +  (lambda xp) w^2  -  yp w^3     with w^2 = v, w^3 = v w.
struct PairTerm { G1 p; G2 q; };

inline Fq12 line_value(const Fq2& lambda, const Fq2& xt, const Fq2& yt, const G1& p) {
    Fq12 l;
    l.c0 = {yt - lambda * xt, lambda.scale(p.x), Fq2::zero()};
    l.c1 = {Fq2::zero(), Fq2{p.y.neg(), Fq::zero()}, Fq2::zero()};
    return l;
}
// prod_i f_{|x|,Q_i}(P_i), before the final exponentiation.  Terms with an identity are skipped.
// Returns false only on a degenerate step (cannot happen for points of order r).
inline bool miller_loop(const std::vector<PairTerm>& terms, Fq12* out) {
    std::vector<PairTerm> t;
    for (const auto& e : terms) if (!e.p.inf && !e.q.inf) t.push_back(e);
    std::vector<G2> T;
    for (const auto& e : t) T.push_back(e.q);
    Fq12 f = Fq12::one();
    for (int i = 62; i >= 0; --i) {                                   // bit 63 of |x| is the leading one
        f = f.sqr();
        for (size_t k = 0; k < t.size(); ++k) {
            Fq2 x2 = T[k].x.sqr(), den = T[k].y.dbl();
            if (den.is_zero()) return false;
            Fq2 lambda = (x2.dbl() + x2) * den.inv();
            f = f * line_value(lambda, T[k].x, T[k].y, t[k].p);
            Fq2 x3 = lambda.sqr() - T[k].x.dbl();
            T[k] = {x3, lambda * (T[k].x - x3) - T[k].y, false};
        }
        if ((BLS_X_ABS >> i) & 1) {
            for (size_t k = 0; k < t.size(); ++k) {
                Fq2 den = t[k].q.x - T[k].x;
                if (den.is_zero()) return false;
                Fq2 lambda = (t[k].q.y - T[k].y) * den.inv();
                f = f * line_value(lambda, T[k].x, T[k].y, t[k].p);
                Fq2 x3 = lambda.sqr() - T[k].x - t[k].q.x;
                T[k] = {x3, lambda * (T[k].x - x3) - T[k].y, false};
            }
        }
    }
    *out = f;
    return true;
}
inline Fq12 final_exponentiation(const Fq12& f) { return (f.conj() * f.inv()).pow(FINAL_EXP_TAIL, 32); }
inline bool pairing_product_is_one(const std::vector<PairTerm>& terms) {
    Fq12 f;
    return miller_loop(terms, &f) && final_exponentiation(f) == Fq12::one();
}

// scalars: 32-byte little-endian canonical Fr
inline bool fr_canonical(const uint8_t* b32) {
    uint64_t c[4]; memcpy(c, b32, 32);
    for (int i = 3; i >= 0; --i) if (c[i] != FR_MOD[i]) return c[i] < FR_MOD[i];
    return false;
}
template <class F> inline Pt<F> scalar_mul(const Aff<F>& p, const uint8_t* k32) { uint64_t k[4]; memcpy(k, k32, 32); return Pt<F>::from(p).mul(k, 4); }

}  // namespace b2h
