// Double-width products and a separate Montgomery reduction for Fq / Fq2.
//
// fp.cuh's operator* interleaves the product rows with the reduction rows (CIOS): 2 N^2 + N multiplier slots
// per product.  Splitting the two halves costs the same for one product but opens three savings:
//   * Karatsuba on the product half (N = 12 -> three 6 x 6 products: 108 wide MADs instead of 144; the extra
//     additions run on the ALU pipe, which the multiplier-bound kernels leave idle);
//   * a dedicated squaring (N diagonal + N (N - 1) / 2 off-diagonal products: 78 instead of 144);
//   * lazy reduction in Fq2: (a0 + a1 u)(b0 + b1 u) needs three double-width products but only TWO reductions
//     (c0 = redc(a0 b0 - a1 b1), c1 = redc((a0 + a1)(b0 + b1) - a0 b0 - a1 b1)) instead of three.
// Everything is written with the carry-chain primitives of fp.cuh (ptx::), so tests/host_emu executes the
// identical limb algorithm on the CPU.  Replaces the same un-vendored crypto3-multiprecision / crypto3-algebra
// arithmetic as fp.cuh.
#pragma once
#include "fp.cuh"

namespace b2 {
namespace wide {

// (acc[j], acc[j+1]) += a[j] * b for even j < n, one carry chain (carry out left in CF)
template <int n> B2_HD void mad_pairs_n(uint32_t* acc, const uint32_t* a, uint32_t b) {
    acc[0] = ptx::mad_lo_cc(a[0], b, acc[0]);
    acc[1] = ptx::madc_hi_cc(a[0], b, acc[1]);
#pragma unroll
    for (int j = 2; j < n; j += 2) {
        acc[j] = ptx::madc_lo_cc(a[j], b, acc[j]);
        acc[j + 1] = ptx::madc_hi_cc(a[j], b, acc[j + 1]);
    }
}

// t[0 .. 2n) = a[0 .. n) * b[0 .. n), n even.  Two column frames: in row i the "even" frame holds the 64-bit
// columns at limb positions (i, i+1), (i+2, i+3), .. and the "odd" frame (i+1, i+2), ..; the frames swap roles
// from row to row (the old even frame, shifted down by the limb it has just emitted, becomes the odd one).
// n^2 wide MADs.
template <int n> B2_HD void mul_wide(uint32_t* t, const uint32_t* a, const uint32_t* b) {
    static_assert(n % 2 == 0, "even limb count");
    uint32_t x[n], y[n];
#pragma unroll
    for (int j = 0; j < n; j += 2) { y[j] = ptx::mul_lo(a[j + 1], b[0]); y[j + 1] = ptx::mul_hi(a[j + 1], b[0]); }
#pragma unroll
    for (int j = 0; j < n; j += 2) { x[j] = ptx::mul_lo(a[j], b[0]); x[j + 1] = ptx::mul_hi(a[j], b[0]); }
#pragma unroll
    for (int i = 1; i < n; ++i) {
        uint32_t* E = (i & 1) ? x : y;      // even frame of row i - 1: emits limb i - 1, then shifts into the odd role
        uint32_t* O = (i & 1) ? y : x;      // odd frame of row i - 1 = even frame of row i
        const uint32_t bi = b[i];
        t[i - 1] = E[0];
        O[0] = ptx::add_cc(O[0], E[1]);
#pragma unroll
        for (int j = 0; j < n - 2; j += 2) {
            E[j] = ptx::madc_lo_cc(a[j + 1], bi, E[j + 2]);
            E[j + 1] = ptx::madc_hi_cc(a[j + 1], bi, E[j + 3]);
        }
        E[n - 2] = ptx::madc_lo_cc(a[n - 1], bi, 0);
        E[n - 1] = ptx::madc_hi(a[n - 1], bi, 0);
        mad_pairs_n<n>(O, a, bi);
        E[n - 1] = ptx::addc(E[n - 1], 0);
    }
    // after row n - 1 (n - 1 is odd): even frame = y (limbs n-1 ..), odd frame = x (limbs n ..)
    uint32_t* Ev = ((n - 1) & 1) ? y : x;
    uint32_t* Od = ((n - 1) & 1) ? x : y;
    t[n - 1] = Ev[0];
    t[n] = ptx::add_cc(Ev[1], Od[0]);
#pragma unroll
    for (int k = 2; k < n; ++k) t[n - 1 + k] = ptx::addc_cc(Ev[k], Od[k - 1]);
    t[2 * n - 1] = ptx::addc(Od[n - 1], 0);
}

// r = a - b (n limbs), returns the borrow as 0 / 0xffffffff
template <int n> B2_HD uint32_t sub_n(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    r[0] = ptx::sub_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < n; ++i) r[i] = ptx::subc_cc(a[i], b[i]);
    return ptx::subc(0, 0);
}
// r = a + b (n limbs), returns the carry (0 / 1)
template <int n> B2_HD uint32_t add_n(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    r[0] = ptx::add_cc(a[0], b[0]);
#pragma unroll
    for (int i = 1; i < n; ++i) r[i] = ptx::addc_cc(a[i], b[i]);
    return ptx::addc(0, 0);
}
// |a - b| and the sign mask (0xffffffff when a < b)
template <int n> B2_HD uint32_t abs_diff_n(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t d[n];
    const uint32_t neg = sub_n<n>(d, a, b);
    // negate when the difference wrapped: -d = ~d + 1
    r[0] = ptx::add_cc(d[0] ^ neg, neg & 1u);
#pragma unroll
    for (int i = 1; i < n; ++i) r[i] = ptx::addc_cc(d[i] ^ neg, 0);
    return neg;
}

// 12 x 12 product by one Karatsuba level over 6-limb halves (subtractive form: no 193-bit operands):
//   aL bH + aH bL = aL bL + aH bH - (aL - aH)(bL - bH)
// 3 * 36 = 108 wide MADs + ~100 ALU additions.
B2_HD void mul_wide_karatsuba12(uint32_t* t, const uint32_t* a, const uint32_t* b) {
    constexpr int h = 6;
    uint32_t p0[2 * h], p2[2 * h], pm[2 * h], da[h], db[h];
    mul_wide<h>(p0, a, b);
    mul_wide<h>(p2, a + h, b + h);
    const uint32_t sa = abs_diff_n<h>(da, a, a + h);
    const uint32_t sb = abs_diff_n<h>(db, b, b + h);
    mul_wide<h>(pm, da, db);
    const uint32_t neg = ~(sa ^ sb);            // all ones: (aL - aH)(bL - bH) = +pm, to be subtracted; zero: it is -pm, add pm
    // mid = p0 + p2 -+ pm  (13 limbs, never negative)
    uint32_t mid[2 * h];
    uint32_t top = add_n<2 * h>(mid, p0, p2);
    // mid += (pm ^ neg) + (neg & 1), i.e. -pm in two's complement when neg: the wrap 2^(32*12) is cancelled by top -= 1.
    // The + 1 enters as the carry flag: neg + neg carries exactly when neg is all ones.
    (void)ptx::add_cc(neg, neg);
#pragma unroll
    for (int i = 0; i < 2 * h; ++i) mid[i] = ptx::addc_cc(mid[i], pm[i] ^ neg);
    top += ptx::addc(0, 0);
    top -= neg & 1u;
    // t = p0 + mid * 2^(32 h) + p2 * 2^(64 h)
#pragma unroll
    for (int i = 0; i < h; ++i) t[i] = p0[i];
    t[h] = ptx::add_cc(p0[h], mid[0]);
#pragma unroll
    for (int i = 1; i < h; ++i) t[h + i] = ptx::addc_cc(p0[h + i], mid[i]);
#pragma unroll
    for (int i = 0; i < h; ++i) t[2 * h + i] = ptx::addc_cc(p2[i], mid[h + i]);
    t[3 * h] = ptx::addc_cc(p2[h], top);
#pragma unroll
    for (int i = 1; i < h - 1; ++i) t[3 * h + i] = ptx::addc_cc(p2[h + i], 0);
    t[4 * h - 1] = ptx::addc(p2[2 * h - 1], 0);
}

// Montgomery reduction of a double-width value: r = T / 2^(32 N) mod p, fully reduced.
// Requires T < p * 2^(32 N) (so that the high half is < p).  The low half seeds the even frame, N reduction
// rows (m = T_i * inv, T += m p 2^(32 i)) run exactly like the reduction half of fp.cuh's CIOS, and the high
// half is added at the end: (T_low + M) / 2^(32 N) <= p, + T_high < p  =>  < 2 p, one conditional subtraction.
// N^2 wide MADs + N low MULs.
template <class PR> B2_HD void redc_wide(uint32_t* r, const uint32_t* t) {
    constexpr int N = PR::N;
    uint32_t x[N], y[N], mod[N];
#pragma unroll
    for (int i = 0; i < N; ++i) { mod[i] = PR::mod_limb(i); x[i] = t[i]; }
    const uint32_t inv = PR::inv_limb();
    {   // row 0: x is the even frame, y the (empty) odd frame
        const uint32_t m = ptx::mul_lo(x[0], inv);
#pragma unroll
        for (int j = 0; j < N; j += 2) { y[j] = ptx::mul_lo(mod[j + 1], m); y[j + 1] = ptx::mul_hi(mod[j + 1], m); }
        mad_pairs_n<N>(x, mod, m);
        y[N - 1] = ptx::addc(y[N - 1], 0);
    }
#pragma unroll
    for (int i = 1; i < N; ++i) {
        uint32_t* E = (i & 1) ? x : y;      // even frame of row i - 1 (its limb 0 is now zero)
        uint32_t* O = (i & 1) ? y : x;      // becomes the even frame of row i
        O[0] = ptx::add_cc(O[0], E[1]);
        const uint32_t m = ptx::mul_lo(O[0], inv);       // mul.lo leaves the carry flag alone
#pragma unroll
        for (int j = 0; j < N - 2; j += 2) {
            E[j] = ptx::madc_lo_cc(mod[j + 1], m, E[j + 2]);
            E[j + 1] = ptx::madc_hi_cc(mod[j + 1], m, E[j + 3]);
        }
        E[N - 2] = ptx::madc_lo_cc(mod[N - 1], m, 0);
        E[N - 1] = ptx::madc_hi(mod[N - 1], m, 0);
        mad_pairs_n<N>(O, mod, m);
        E[N - 1] = ptx::addc(E[N - 1], 0);
    }
    uint32_t* Ev = ((N - 1) & 1) ? y : x;   // limb 0 zeroed by the last row; limbs 1.. are positions N ..
    uint32_t* Od = ((N - 1) & 1) ? x : y;   // positions N ..
    uint32_t s[N];
    s[0] = ptx::add_cc(Ev[1], Od[0]);
#pragma unroll
    for (int k = 1; k < N - 1; ++k) s[k] = ptx::addc_cc(Ev[k + 1], Od[k]);
    s[N - 1] = ptx::addc(Od[N - 1], 0);
    s[0] = ptx::add_cc(s[0], t[N]);
#pragma unroll
    for (int k = 1; k < N - 1; ++k) s[k] = ptx::addc_cc(s[k], t[N + k]);
    s[N - 1] = ptx::addc(s[N - 1], t[2 * N - 1]);
    Fp<PR>::final_sub(r, s, 0);
}

// a * b through the split path (KARATSUBA: 264 multiplier slots for Fq, else 300 like operator*)
template <bool KARATSUBA> B2_HD Fq fq_mul(const Fq& a, const Fq& b) {
    uint32_t t[24];
    if (KARATSUBA) mul_wide_karatsuba12(t, a.v, b.v); else mul_wide<12>(t, a.v, b.v);
    Fq r; redc_wide<FqParams>(r.v, t); return r;
}

// a^2 for a 12-limb value through its 6-limb halves: a^2 = aL^2 + 2 aL aH 2^192 + aH^2 2^384 -- three 6 x 6 products
// (108 wide MADs instead of 144) and ~40 additions; no sign handling, unlike the Karatsuba product.
B2_HD void sqr_wide_halves12(uint32_t* t, const uint32_t* a) {
    constexpr int h = 6;
    uint32_t p0[2 * h], p2[2 * h], pm[2 * h + 1];
    mul_wide<h>(p0, a, a);
    mul_wide<h>(p2, a + h, a + h);
    mul_wide<h>(pm, a, a + h);
    // pm <- 2 pm (13 limbs)
    pm[2 * h] = pm[2 * h - 1] >> 31;
#pragma unroll
    for (int i = 2 * h - 1; i > 0; --i) pm[i] = (pm[i] << 1) | (pm[i - 1] >> 31);
    pm[0] <<= 1;
    // t = p0 + pm 2^(32 h) + p2 2^(64 h)
#pragma unroll
    for (int i = 0; i < h; ++i) t[i] = p0[i];
    t[h] = ptx::add_cc(p0[h], pm[0]);
#pragma unroll
    for (int i = 1; i < h; ++i) t[h + i] = ptx::addc_cc(p0[h + i], pm[i]);
#pragma unroll
    for (int i = 0; i < h; ++i) t[2 * h + i] = ptx::addc_cc(p2[i], pm[h + i]);
    t[3 * h] = ptx::addc_cc(p2[h], pm[2 * h]);
#pragma unroll
    for (int i = 1; i < h - 1; ++i) t[3 * h + i] = ptx::addc_cc(p2[h + i], 0);
    t[4 * h - 1] = ptx::addc(p2[2 * h - 1], 0);
}
B2_HD Fq fq_sqr_halves(const Fq& a) {
    uint32_t t[24];
    sqr_wide_halves12(t, a.v);
    Fq r; redc_wide<FqParams>(r.v, t); return r;
}

// ---- Fq2 = Fq[u] / (u^2 + 1) with lazy reduction -------------------------------------------------
// t -= s over 24 limbs; returns the borrow mask
B2_HD uint32_t sub_wide(uint32_t* t, const uint32_t* s) { return sub_n<24>(t, t, s); }

template <bool KARATSUBA> B2_HD Fq2 fq2_mul_lazy(const Fq2& a, const Fq2& b) {
    uint32_t t0[24], t1[24], t2[24], sa[12], sb[12];
    if (KARATSUBA) { mul_wide_karatsuba12(t0, a.c0.v, b.c0.v); mul_wide_karatsuba12(t1, a.c1.v, b.c1.v); }
    else { mul_wide<12>(t0, a.c0.v, b.c0.v); mul_wide<12>(t1, a.c1.v, b.c1.v); }
    add_n<12>(sa, a.c0.v, a.c1.v);            // < 2p < 2^382: no carry, no reduction
    add_n<12>(sb, b.c0.v, b.c1.v);
    if (KARATSUBA) mul_wide_karatsuba12(t2, sa, sb); else mul_wide<12>(t2, sa, sb);
    // c1 = (a0 + a1)(b0 + b1) - a0 b0 - a1 b1 = a0 b1 + a1 b0 in [0, 2 p^2)
    sub_wide(t2, t0);
    sub_wide(t2, t1);
    // c0 = a0 b0 - a1 b1 in (-p^2, p^2): add p * 2^384 when negative
    const uint32_t neg = sub_wide(t0, t1);
    t0[12] = ptx::add_cc(t0[12], FqParams::mod_limb(0) & neg);
#pragma unroll
    for (int i = 1; i < 11; ++i) t0[12 + i] = ptx::addc_cc(t0[12 + i], FqParams::mod_limb(i) & neg);
    t0[23] = ptx::addc(t0[23], FqParams::mod_limb(11) & neg);
    Fq2 r;
    redc_wide<FqParams>(r.c0.v, t0);
    redc_wide<FqParams>(r.c1.v, t2);
    return r;
}

}  // namespace wide
}  // namespace b2
