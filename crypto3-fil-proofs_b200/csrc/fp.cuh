// Montgomery prime-field arithmetic on 32-bit limbs for sm_100a.
//
// Replaces what the reference delegates to crypto3-multiprecision `modular_adaptor`
// and crypto3-algebra `fields::bls12_{base,scalar}_field<381>` (un-vendored; SURVEY.md
// §8 a15).  Values are little-endian arrays of N uint32 limbs in Montgomery form
// (x * 2^(32N) mod p), always fully reduced to [0, p).
//
// The multiplier keeps two interleaved accumulators ("even" and "odd" columns) so that
// every 32x32->64 product is one `mad.lo.cc / madc.hi.cc` pair on a two-register
// aligned pair -- ptxas fuses each pair into a single IMAD.WIDE(.X) on the FMA pipe.
// Cost: 2*N^2 wide MADs + N low MULs per product (N=12: 288+12, N=8: 128+8).
//
// Every primitive has a host emulation (carry flag in a thread-local) so the exact
// limb algorithm is unit-tested on the CPU (tests/test_host_emulation.py) before it
// is run on a B200.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
// the field product: inlined into device code, an ordinary call in host code (the host
// passes emulate the carry chain instruction by instruction; inlining 10+ of them per
// group operation only costs compile time)
#if defined(__CUDA_ARCH__)
#define B2_HD_MUL __host__ __device__ __forceinline__
#else
#define B2_HD_MUL __host__ __device__ __attribute__((noinline))
#endif
#else
#define B2_HD inline
#define B2_D inline
#define B2_HD_MUL inline
#endif

#if !defined(__CUDACC__)
struct uint4 { uint32_t x, y, z, w; };      // host-emulation builds only
#endif

namespace b2 {

// ---------------------------------------------------------------- PTX primitives
namespace ptx {
#if defined(__CUDA_ARCH__)
B2_D uint32_t add_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t addc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t addc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t sub_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t subc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t subc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
B2_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
B2_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
B2_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
B2_D uint32_t mul_lo(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
B2_D uint32_t mul_hi(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
#else
// Host emulation: one carry flag per thread, same semantics as PTX CC.CF.
inline uint32_t& cf() { static thread_local uint32_t f = 0; return f; }
inline uint32_t add_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b; cf() = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b + cf(); cf() = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc(uint32_t a, uint32_t b) { return a + b + cf(); }
inline uint32_t sub_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b; cf() = (uint32_t)(t >> 32) & 1; return (uint32_t)t; }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b - cf(); cf() = (uint32_t)(t >> 32) & 1; return (uint32_t)t; }
inline uint32_t subc(uint32_t a, uint32_t b) { return a - b - cf(); }
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)(uint32_t)((uint64_t)a * b) + c; cf() = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)(uint32_t)((uint64_t)a * b) + c + cf(); cf() = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (((uint64_t)a * b) >> 32) + c + cf(); cf() = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return (uint32_t)((((uint64_t)a * b) >> 32) + c + cf()); }
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return (uint32_t)((uint64_t)a * b); }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
#endif
}  // namespace ptx

// ------------------------------------------------------------------ field params
// All constants below are re-derived and asserted by tests/test_host_emulation.py
// (p, r from the BLS parameter; R, R^2, -p^-1 by big-int arithmetic).
struct FrParams {
    static constexpr int N = 8;
    // r = .. ffffffff 00000001: -r^-1 = -1 (mod 2^32), so the Montgomery quotient digit is m = -T[0], m * r[0] = m and
    // m * r[1] = (m << 32) - m -- the two low limbs of every reduction row need no multiplier (mad_redc below)
    static constexpr bool LOW_LIMBS_ARE_1_FFFFFFFF = true;
    // r = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
    B2_HD static constexpr uint32_t P(int i) {
        constexpr uint32_t t[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u,
                                      0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u};
        return t[i];
    }
    static constexpr uint32_t INV = 0xffffffffu;          // -r^-1 mod 2^32
    B2_HD static uint32_t mod_limb(int i);                // modulus / INV as the multiplier reads them
    B2_HD static uint32_t inv_limb();                     // (constant memory on the device)
    B2_HD static constexpr uint32_t ONE(int i) {
        constexpr uint32_t t[8] = {0xfffffffeu, 0x00000001u, 0x00034802u, 0x5884b7fau,
                                        0xecbc4ff5u, 0x998c4fefu, 0xacc5056fu, 0x1824b159u};
        return t[i];
    }   // 2^256 mod r
    B2_HD static constexpr uint32_t R2(int i) {
        constexpr uint32_t t[8] = {0xf3f29c6du, 0xc999e990u, 0x87925c23u, 0x2b6cedcbu,
                                       0x7254398fu, 0x05d31496u, 0x9f59ff11u, 0x0748d9d9u};
        return t[i];
    }    // 2^512 mod r
};
struct FqParams {
    static constexpr int N = 12;
    static constexpr bool LOW_LIMBS_ARE_1_FFFFFFFF = false;
    // p = 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
    B2_HD static constexpr uint32_t P(int i) {
        constexpr uint32_t t[12] = {0xffffaaabu, 0xb9feffffu, 0xb153ffffu, 0x1eabfffeu, 0xf6b0f624u, 0x6730d2a0u,
                                       0xf38512bfu, 0x64774b84u, 0x434bacd7u, 0x4b1ba7b6u, 0x397fe69au, 0x1a0111eau};
        return t[i];
    }
    static constexpr uint32_t INV = 0xfffcfffdu;          // -p^-1 mod 2^32
    B2_HD static uint32_t mod_limb(int i);
    B2_HD static uint32_t inv_limb();
    B2_HD static constexpr uint32_t ONE(int i) {
        constexpr uint32_t t[12] = {0x0002fffdu, 0x76090000u, 0xc40c0002u, 0xebf4000bu, 0x53c758bau, 0x5f489857u,
                                         0x70525745u, 0x77ce5853u, 0xa256ec6du, 0x5c071a97u, 0xfa80e493u, 0x15f65ec3u};
        return t[i];
    }  // 2^384 mod p
    B2_HD static constexpr uint32_t R2(int i) {
        constexpr uint32_t t[12] = {0x1c341746u, 0xf4df1f34u, 0x09d104f1u, 0x0a76e6a6u, 0x4c95b6d5u, 0x8de5476cu,
                                        0x939d83c0u, 0x67eb88a9u, 0xb519952du, 0x9a793e85u, 0x92cae3aau, 0x11988fe5u};
        return t[i];
    }  // 2^768 mod p
};

#if defined(__CUDACC__)
// Runtime-visible copies: ptxas must not constant-fold the modulus into the multiplier.
static __device__ __constant__ uint32_t k_fr_mod[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u,
                                                       0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u};
static __device__ __constant__ uint32_t k_fq_mod[12] = {0xffffaaabu, 0xb9feffffu, 0xb153ffffu, 0x1eabfffeu, 0xf6b0f624u, 0x6730d2a0u,
                                                        0xf38512bfu, 0x64774b84u, 0x434bacd7u, 0x4b1ba7b6u, 0x397fe69au, 0x1a0111eau};
static __device__ __constant__ uint32_t k_fr_inv = 0xffffffffu;
static __device__ __constant__ uint32_t k_fq_inv = 0xfffcfffdu;
#endif
B2_HD uint32_t FrParams::mod_limb(int i) {
#if defined(__CUDA_ARCH__)
    return k_fr_mod[i];
#else
    return P(i);
#endif
}
B2_HD uint32_t FrParams::inv_limb() {
#if defined(__CUDA_ARCH__)
    return k_fr_inv;
#else
    return INV;
#endif
}
B2_HD uint32_t FqParams::mod_limb(int i) {
#if defined(__CUDA_ARCH__)
    return k_fq_mod[i];
#else
    return P(i);
#endif
}
B2_HD uint32_t FqParams::inv_limb() {
#if defined(__CUDA_ARCH__)
    return k_fq_inv;
#else
    return INV;
#endif
}

// ----------------------------------------------------------------------- Fp<P>
template <class PR> struct alignas(16) Fp {
    static constexpr int N = PR::N;
    uint32_t v[N];

    B2_HD static Fp zero() { Fp r; for (int i = 0; i < N; ++i) r.v[i] = 0; return r; }
    B2_HD static Fp one() { Fp r; for (int i = 0; i < N; ++i) r.v[i] = PR::ONE(i); return r; }
    B2_HD static Fp r2() { Fp r; for (int i = 0; i < N; ++i) r.v[i] = PR::R2(i); return r; }
    B2_HD bool is_zero() const { uint32_t t = 0; for (int i = 0; i < N; ++i) t |= v[i]; return t == 0; }
    B2_HD bool operator==(const Fp& o) const { uint32_t t = 0; for (int i = 0; i < N; ++i) t |= v[i] ^ o.v[i]; return t == 0; }
    B2_HD bool operator!=(const Fp& o) const { return !(*this == o); }

    // t >= p ?  (raw limb compare; used for canonical-input validation)
    B2_HD static bool geq_p(const uint32_t* t) {
        for (int i = N - 1; i >= 0; --i) {
            if (t[i] != PR::P(i)) return t[i] > PR::P(i);
        }
        return true;
    }

    // r = t - p if t >= p else t   (t < 2p, optional extra top bit `hi`)
    B2_HD static void final_sub(uint32_t* r, const uint32_t* t, uint32_t hi) {
        uint32_t d[N];
        d[0] = ptx::sub_cc(t[0], PR::P(0));
#pragma unroll
        for (int i = 1; i < N; ++i) d[i] = ptx::subc_cc(t[i], PR::P(i));
        uint32_t b = ptx::subc(hi, 0);          // 0 if t >= p (no borrow), 0xffffffff if borrow
        bool keep = (b != 0);
#pragma unroll
        for (int i = 0; i < N; ++i) r[i] = keep ? t[i] : d[i];
    }

    B2_HD Fp operator+(const Fp& o) const {
        uint32_t t[N];
        t[0] = ptx::add_cc(v[0], o.v[0]);
#pragma unroll
        for (int i = 1; i < N; ++i) t[i] = ptx::addc_cc(v[i], o.v[i]);
        uint32_t hi = ptx::addc(0, 0);          // p < 2^(32N-1) so hi is always 0, kept for safety
        Fp r; final_sub(r.v, t, hi); return r;
    }
    B2_HD Fp operator-(const Fp& o) const {
        Fp r;
        r.v[0] = ptx::sub_cc(v[0], o.v[0]);
#pragma unroll
        for (int i = 1; i < N; ++i) r.v[i] = ptx::subc_cc(v[i], o.v[i]);
        uint32_t mask = ptx::subc(0, 0);        // all ones if borrow
        r.v[0] = ptx::add_cc(r.v[0], PR::P(0) & mask);
#pragma unroll
        for (int i = 1; i < N - 1; ++i) r.v[i] = ptx::addc_cc(r.v[i], PR::P(i) & mask);
        r.v[N - 1] = ptx::addc(r.v[N - 1], PR::P(N - 1) & mask);
        return r;
    }
    B2_HD Fp neg() const { return zero() - *this; }
    B2_HD Fp dbl() const { return *this + *this; }

    // acc pairs (acc[j], acc[j+1]) += a[j] * b for j = 0, 2, .., N-2, one carry chain.
    // Each lo/hi pair is one IMAD.WIDE.U32(.X) in SASS.
    B2_HD static void mad_pairs(uint32_t* acc, const uint32_t* a, uint32_t b) {
        acc[0] = ptx::mad_lo_cc(a[0], b, acc[0]);
        acc[1] = ptx::madc_hi_cc(a[0], b, acc[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            acc[j] = ptx::madc_lo_cc(a[j], b, acc[j]);
            acc[j + 1] = ptx::madc_hi_cc(a[j], b, acc[j + 1]);
        }
    }
    // One CIOS step on the split accumulator.  In the frame of this step `ev` holds the
    // 64-bit columns at limb positions (0,1),(2,3),.. and `od` the columns (1,2),(3,4),..;
    // on entry (unless FIRST) `od` is still the previous step's `ev`, i.e. not yet shifted
    // down by the limb that the previous reduction zeroed.  T += a*bi; T += m*p; T >>= 32
    // (the shift is the role swap done by the caller).
    template <bool FIRST>
    B2_HD static void mad_redc(uint32_t* ev, uint32_t* od, const uint32_t* a, uint32_t bi, const uint32_t* mod, uint32_t inv) {
        if (FIRST) {
#pragma unroll
            for (int j = 0; j < N; j += 2) { od[j] = ptx::mul_lo(a[j + 1], bi); od[j + 1] = ptx::mul_hi(a[j + 1], bi); }
#pragma unroll
            for (int j = 0; j < N; j += 2) { ev[j] = ptx::mul_lo(a[j], bi); ev[j + 1] = ptx::mul_hi(a[j], bi); }
        } else {
            ev[0] = ptx::add_cc(ev[0], od[1]);              // limb left over by the shift
#pragma unroll
            for (int j = 0; j < N - 2; j += 2) {            // shifted columns + odd limbs of a
                od[j] = ptx::madc_lo_cc(a[j + 1], bi, od[j + 2]);
                od[j + 1] = ptx::madc_hi_cc(a[j + 1], bi, od[j + 3]);
            }
            od[N - 2] = ptx::madc_lo_cc(a[N - 1], bi, 0);
            od[N - 1] = ptx::madc_hi(a[N - 1], bi, 0);
            mad_pairs(ev, a, bi);
            od[N - 1] = ptx::addc(od[N - 1], 0);
        }
        if (PR::LOW_LIMBS_ARE_1_FFFFFFFF) {
            // Fr: m = -T[0].  Row m * r added by hand for the limbs r[0] = 1 and r[1] = 2^32 - 1:
            //   odd chain  (columns 1,2): + m * (2^32 - 1) = { lo = -m = T[0], hi = m - [m != 0] }
            //   even chain (columns 0,1): + m            -> T[0] + m = 2^32 [m != 0]: only the carry [T[0] != 0] survives
            // and six IMAD.WIDE.X for r[2..7] instead of eight + the mul.lo for m: 112 multiplier slots per product, not 136.
            const uint32_t t0 = ev[0];
            const uint32_t m = 0u - t0;
            const uint32_t nz = t0 != 0u ? 1u : 0u;
            od[0] = ptx::add_cc(od[0], t0);
            od[1] = ptx::addc_cc(od[1], m - nz);
#pragma unroll
            for (int j = 2; j < N; j += 2) {
                od[j] = ptx::madc_lo_cc(mod[j + 1], m, od[j]);
                od[j + 1] = ptx::madc_hi_cc(mod[j + 1], m, od[j + 1]);
            }
            ev[0] = 0u;
            ev[1] = ptx::add_cc(ev[1], nz);
#pragma unroll
            for (int j = 2; j < N; j += 2) {
                ev[j] = ptx::madc_lo_cc(mod[j], m, ev[j]);
                ev[j + 1] = ptx::madc_hi_cc(mod[j], m, ev[j + 1]);
            }
            od[N - 1] = ptx::addc(od[N - 1], 0);
            return;
        }
        // m is computed through an opaque mul.lo with `inv` read from constant memory (a visible m = -T[0] lets ptxas
        // rewrite the reduction rows into unpaired lo/hi products)
        const uint32_t m = ptx::mul_lo(ev[0], inv);
        mad_pairs(od, mod + 1, m);
        mad_pairs(ev, mod, m);
        od[N - 1] = ptx::addc(od[N - 1], 0);
    }

    // Montgomery product (CIOS, 32-bit limbs, even/odd 64-bit columns; the classic
    // carry-chain layout of GPU big-number code): N^2 + N^2 wide MADs + N low MULs.
    B2_HD_MUL Fp operator*(const Fp& o) const {
        static_assert(N % 2 == 0, "even limb count");
        uint32_t ev[N], od[N], mod[N], a[N], b[N];
#pragma unroll
        for (int i = 0; i < N; ++i) { mod[i] = PR::mod_limb(i); a[i] = v[i]; b[i] = o.v[i]; }   // operands in registers: the
        // asm statements below are volatile, a memory-resident operand would be re-read per limb
        const uint32_t inv = PR::inv_limb();
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            if (i == 0) mad_redc<true>(ev, od, a, b[0], mod, inv); else mad_redc<false>(ev, od, a, b[i], mod, inv);
            mad_redc<false>(od, ev, a, b[i + 1], mod, inv);
        }
        // T = ev (columns at 0,2,..) + od shifted by the dropped limb; T < 2p < 2^(32N)
        uint32_t t[N];
        t[0] = ptx::add_cc(ev[0], od[1]);
#pragma unroll
        for (int k = 1; k < N - 1; ++k) t[k] = ptx::addc_cc(ev[k], od[k + 1]);
        t[N - 1] = ptx::addc(ev[N - 1], 0);
        Fp r; final_sub(r.v, t, 0); return r;
    }
    B2_HD Fp sqr() const { return *this * *this; }

    // canonical (non-Montgomery) limbs <-> Montgomery
    B2_HD static Fp from_canonical(const uint32_t* c) { Fp t; for (int i = 0; i < N; ++i) t.v[i] = c[i]; return t * r2(); }
    B2_HD void to_canonical(uint32_t* c) const {
        Fp o = zero(); o.v[0] = 1;
        Fp t = *this * o;
        for (int i = 0; i < N; ++i) c[i] = t.v[i];
    }

    // a^e for a public exponent given as limbs (square-and-multiply, MSB first)
    B2_HD Fp pow_limbs(const uint32_t* e, int n) const {
        Fp r = one();
        for (int i = n * 32 - 1; i >= 0; --i) {
            r = r.sqr();
            if ((e[i >> 5] >> (i & 31)) & 1) r = r * *this;
        }
        return r;
    }
    // Fermat inverse a^(p-2); inv(0) = 0
    B2_HD Fp inv() const {
        uint32_t e[N];
        uint32_t borrow = 2;                    // e = p - 2 with borrow propagation (r[0] == 1)
        for (int i = 0; i < N; ++i) {
            uint32_t pi = PR::P(i);
            e[i] = pi - borrow;
            borrow = (pi < borrow) ? 1u : 0u;
        }
        return pow_limbs(e, N);
    }
    // Inverse by the binary extended Euclidean algorithm on the limbs: ~2 log2(p) halvings and log2(p) subtractions of
    // N-limb integers instead of the ~1.5 log2(p) full products of the Fermat ladder -- about ten times shorter as a
    // *serial* chain, which is what matters where one thread inverts the product of a whole batch (msm_tree.cuh).
    // Invariant: x1 * a == u, x2 * a == v (mod p).  inv(0) = 0.
    B2_HD Fp inv_gcd() const {
        if (is_zero()) return zero();
        uint32_t u[N], w[N];
        Fp x1 = zero(), x2 = zero();
        x1.v[0] = 1;
        for (int i = 0; i < N; ++i) { u[i] = v[i]; w[i] = PR::P(i); }
        auto is_one = [](const uint32_t* t) { uint32_t r = t[0] ^ 1u; for (int i = 1; i < N; ++i) r |= t[i]; return r == 0; };
        auto shr1 = [](uint32_t* t, uint32_t top) { for (int i = 0; i < N - 1; ++i) t[i] = (t[i] >> 1) | (t[i + 1] << 31); t[N - 1] = (t[N - 1] >> 1) | (top << 31); };
        auto halve_mod = [&](Fp& x) {                    // x / 2 mod p for x in [0, p)
            uint32_t top = 0;
            if (x.v[0] & 1u) {
                x.v[0] = ptx::add_cc(x.v[0], PR::P(0));
                for (int i = 1; i < N; ++i) x.v[i] = ptx::addc_cc(x.v[i], PR::P(i));
                top = ptx::addc(0, 0);
            }
            shr1(x.v, top);
        };
        auto geq = [](const uint32_t* a, const uint32_t* b) { for (int i = N - 1; i >= 0; --i) if (a[i] != b[i]) return a[i] > b[i]; return true; };
        auto sub_n = [](uint32_t* a, const uint32_t* b) {  // a -= b, a >= b
            a[0] = ptx::sub_cc(a[0], b[0]);
            for (int i = 1; i < N - 1; ++i) a[i] = ptx::subc_cc(a[i], b[i]);
            a[N - 1] = ptx::subc(a[N - 1], b[N - 1]);
        };
        while (!is_one(u) && !is_one(w)) {
            while (!(u[0] & 1u)) { shr1(u, 0); halve_mod(x1); }
            while (!(w[0] & 1u)) { shr1(w, 0); halve_mod(x2); }
            if (geq(u, w)) { sub_n(u, w); x1 = x1 - x2; }
            else { sub_n(w, u); x2 = x2 - x1; }
        }
        // y = (a R)^-1 as a plain residue; a^-1 R = y R^2: two Montgomery products with R^2
        const Fp y = is_one(u) ? x1 : x2;
        return (y * r2()) * r2();
    }
};

typedef Fp<FrParams> Fr;
typedef Fp<FqParams> Fq;

// ------------------------------------------------------------------------- Fq2
// Fq[u]/(u^2+1).  mul = 3 Fq products (Karatsuba), sqr = 2.
struct Fq2 {
    Fq c0, c1;
    B2_HD static Fq2 zero() { return {Fq::zero(), Fq::zero()}; }
    B2_HD static Fq2 one() { return {Fq::one(), Fq::zero()}; }
    B2_HD bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    B2_HD bool operator==(const Fq2& o) const { return c0 == o.c0 && c1 == o.c1; }
    B2_HD bool operator!=(const Fq2& o) const { return !(*this == o); }
    B2_HD Fq2 operator+(const Fq2& o) const { return {c0 + o.c0, c1 + o.c1}; }
    B2_HD Fq2 operator-(const Fq2& o) const { return {c0 - o.c0, c1 - o.c1}; }
    B2_HD Fq2 neg() const { return {c0.neg(), c1.neg()}; }
    B2_HD Fq2 dbl() const { return {c0.dbl(), c1.dbl()}; }
    B2_HD Fq2 mul_inline(const Fq2& o) const {
        Fq ac = c0 * o.c0, bd = c1 * o.c1;
        Fq t = (c0 + c1) * (o.c0 + o.c1);
        return {ac - bd, t - ac - bd};
    }
    B2_HD Fq2 sqr_inline() const {
        Fq ab = c0 * c1;
        return {(c0 + c1) * (c0 - c1), ab.dbl()};
    }
    // On the device the Fq2 product is an out-of-line call: three inlined 12-limb products
    // per call site made G2 kernels too large to compile (and to fit the instruction cache).
    B2_HD Fq2 operator*(const Fq2& o) const;
    B2_HD Fq2 sqr() const;
    B2_HD Fq2 inv() const {
        Fq n = (c0.sqr() + c1.sqr()).inv();
        return {c0 * n, (c1 * n).neg()};
    }
    B2_HD Fq2 inv_gcd() const {
        Fq n = (c0.sqr() + c1.sqr()).inv_gcd();
        return {c0 * n, (c1 * n).neg()};
    }
};

}  // namespace b2

#include "fp_wide.cuh"      // double-width products, separate reduction, lazy-reduction Fq2 product (needs Fq, Fq2 above)

namespace b2 {

// The Fq2 product everybody uses: three double-width products, two reductions (fp_wide.cuh) -- 744 multiplier slots
// instead of the 900 of three full Montgomery products; measured 1.17e10 against 0.97e10 products/s on B200.
#if defined(__CUDA_ARCH__)
static __device__ __noinline__ void fq2_mul_call(Fq2* r, const Fq2* a, const Fq2* b) { *r = wide::fq2_mul_lazy<false>(*a, *b); }
static __device__ __noinline__ void fq2_sqr_call(Fq2* r, const Fq2* a) { *r = a->sqr_inline(); }
B2_HD Fq2 Fq2::operator*(const Fq2& o) const { Fq2 r; fq2_mul_call(&r, this, &o); return r; }
B2_HD Fq2 Fq2::sqr() const { Fq2 r; fq2_sqr_call(&r, this); return r; }
#else
B2_HD Fq2 Fq2::operator*(const Fq2& o) const { return wide::fq2_mul_lazy<false>(*this, o); }
B2_HD Fq2 Fq2::sqr() const { return sqr_inline(); }
#endif

}  // namespace b2
