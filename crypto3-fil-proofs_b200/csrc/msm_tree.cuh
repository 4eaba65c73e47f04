// Bucket accumulation by a pairwise tree of *affine* additions with one shared inversion per level
// ("batched affine"): the part of the Pippenger multiexp (msm_impl.cuh step 4) that costs
// n * windows group additions.  Replaces crypto3-algebra `multiexp_with_mixed_addition`'s inner loop
// (un-vendored) for the queries of
// libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:46-66.
//
// A mixed XYZZ addition costs 8M + 2S = 10 field products.  An affine addition costs
// lambda = (y1 - y0) / (x1 - x0), x2 = lambda^2 - x0 - x1, y2 = lambda (x0 - x2) - y0: 2M + 1S and one
// inversion -- and Montgomery's trick turns k inversions into one inversion + 3 (k - 1) products.  So
// if millions of independent additions are available at once, an addition costs ~6 products instead
// of 10.  The sorted bucket layout provides them: every bucket is reduced by a binary tree, level l
// turns the entries (e_0, e_1, e_2, ..) of a bucket into (e_0 + e_1, e_2 + e_3, ..) (an odd last entry
// is copied), all buckets and all windows of a level form ONE batch:
//
//   up0   : thread t owns K consecutive outputs; d_j = x1 - x0 (2 y0 for a doubling, 1 for a copy);
//           writes the exclusive prefix products pp[o] and its total product tot1[t]
//   up    : the same prefix-product step on the totals, K2 per thread, until <= TOP values remain
//   top   : one thread inverts the remaining product (Fermat) and back-substitutes
//   down  : inverse of every total from the level above:  inv[i] = I * pp[i];  I *= v[i]
//   down0 : inv(d_j) the same way, then the addition itself; writes the affine output
//
// After L levels the buckets hold ~ 2^-L of their entries, which the XYZZ accumulate kernel
// (msm_accumulate_chunked_kernel, direct form) finishes; the bucket reduction is unchanged.
// Results are canonical affine points, so the summation order does not show in the output bytes.
//
// Every kernel body is a B2_HD function of the thread index so that tests/host_emu runs the identical
// index arithmetic and special-case handling on the CPU against the oracle.
#pragma once
#include "ec.cuh"

namespace b2 {

template <class F> B2_HD Affine<F> tree_ld_affine(const Affine<F>* p) {
#if defined(__CUDA_ARCH__)
    Affine<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(Affine<F>) / 16); ++i) { uint4 t = __ldg(q + i); d[4 * i] = t.x; d[4 * i + 1] = t.y; d[4 * i + 2] = t.z; d[4 * i + 3] = t.w; }
    return r;
#else
    return *p;
#endif
}
template <class F> B2_HD F tree_ld_field(const F* p) {
#if defined(__CUDA_ARCH__)
    F r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(F) / 16); ++i) { uint4 t = __ldg(q + i); d[4 * i] = t.x; d[4 * i + 1] = t.y; d[4 * i + 2] = t.z; d[4 * i + 3] = t.w; }
    return r;
#else
    return *p;
#endif
}
template <class F> B2_HD void tree_st_field(F* p, const F& v) {
#if defined(__CUDA_ARCH__)
    uint4* q = reinterpret_cast<uint4*>(p);
    const uint32_t* s = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(F) / 16); ++i) q[i] = make_uint4(s[4 * i], s[4 * i + 1], s[4 * i + 2], s[4 * i + 3]);
#else
    *p = v;
#endif
}

// Entries of a level.  Level 0: entry e is the key point bases[sorted[e] & 0x7fffffff], negated when bit 31 is
// set; off holds absolute entry indices.  Level >= 1: entry e is pts[e].
template <class F> struct TreeIn {
    const Affine<F>* bases;
    const uint32_t* sorted;     // null at levels >= 1
    const Affine<F>* pts;
    const uint32_t* off;        // n_buckets + 1 offsets of this level's entries
};
template <class F> B2_HD Affine<F> tree_entry(const TreeIn<F>& in, uint32_t e) {
    if (in.sorted) {
        const uint32_t ix = in.sorted[e];
        Affine<F> p = tree_ld_affine(in.bases + (ix & 0x7fffffffu));
        if ((ix >> 31) && !p.is_identity()) p.y = p.y.neg();
        return p;
    }
    return tree_ld_affine(in.pts + e);
}
// (An explicit prefetch.global.L2 of the next output's entries was measured and dropped: the prefix pass is DRAM-bound --
// 62 % of peak in ncu -- and the addition pass multiplier-bound, so the extra requests only cost: G1 up0 3.9 -> 5.2 ms,
// down0 13.8 -> 14.2 ms, G2 down0 50.5 -> 51.5 ms per 2^22 multiexp.)
// x coordinate of entry e alone (the sign bit of a level-0 reference does not touch x)
template <class F> B2_HD F tree_entry_x(const TreeIn<F>& in, uint32_t e) {
    if (in.sorted) return tree_ld_field(&in.bases[in.sorted[e] & 0x7fffffffu].x);
    return tree_ld_field(&in.pts[e].x);
}

// Products of the fast path of tree_down0: inlined for both fields.  Fq2::operator* is an out-of-line call whose
// operands and result travel through local memory -- fine for the rare branches, too slow for the loop that does
// ~90 % of a G2 multiexp's arithmetic (-DB2_TREE_INLINE_FQ2=0 restores the calls for comparison).
#ifndef B2_TREE_INLINE_FQ2
#define B2_TREE_INLINE_FQ2 1
#endif
B2_HD Fq tree_mul(const Fq& a, const Fq& b) { return a * b; }
#ifndef B2_TREE_FQ_SQR_HALVES
#define B2_TREE_FQ_SQR_HALVES 0     // measured neutral (13.856 vs 13.852 ms for the five addition passes of a 2^22 multiexp): the pass is not bound by its multiply count
#endif
#if B2_TREE_FQ_SQR_HALVES
B2_HD Fq tree_sqr(const Fq& a) { return wide::fq_sqr_halves(a); }     // 108 + 156 multiplier slots instead of 300
#else
B2_HD Fq tree_sqr(const Fq& a) { return a.sqr(); }
#endif
// -DB2_TREE_FQ2_CALLS=1: the lazy product built from two out-of-line pieces (12 x 12 wide product, Montgomery reduction)
// instead of fully inlined -- a ~4x smaller loop body for the instruction cache at the price of operands passing
// through local memory; measured against the inlined form before choosing.
#ifndef B2_TREE_FQ2_CALLS
#define B2_TREE_FQ2_CALLS 0
#endif
#if defined(__CUDA_ARCH__) && B2_TREE_FQ2_CALLS
static __device__ __noinline__ void tree_mul_wide12_call(uint32_t* t, const uint32_t* a, const uint32_t* b) { wide::mul_wide<12>(t, a, b); }
static __device__ __noinline__ void tree_redc_call(uint32_t* r, const uint32_t* t) { wide::redc_wide<FqParams>(r, t); }
B2_D Fq2 tree_mul(const Fq2& a, const Fq2& b) {
    uint32_t t0[24], t1[24], t2[24], sa[12], sb[12];
    tree_mul_wide12_call(t0, a.c0.v, b.c0.v);
    tree_mul_wide12_call(t1, a.c1.v, b.c1.v);
    wide::add_n<12>(sa, a.c0.v, a.c1.v);
    wide::add_n<12>(sb, b.c0.v, b.c1.v);
    tree_mul_wide12_call(t2, sa, sb);
    wide::sub_wide(t2, t0);
    wide::sub_wide(t2, t1);
    const uint32_t neg = wide::sub_wide(t0, t1);
    t0[12] = ptx::add_cc(t0[12], FqParams::mod_limb(0) & neg);
#pragma unroll
    for (int i = 1; i < 11; ++i) t0[12 + i] = ptx::addc_cc(t0[12 + i], FqParams::mod_limb(i) & neg);
    t0[23] = ptx::addc(t0[23], FqParams::mod_limb(11) & neg);
    Fq2 r;
    tree_redc_call(r.c0.v, t0);
    tree_redc_call(r.c1.v, t2);
    return r;
}
B2_D Fq2 tree_sqr(const Fq2& a) { return a.sqr_inline(); }
#elif B2_TREE_INLINE_FQ2
B2_HD Fq2 tree_mul(const Fq2& a, const Fq2& b) { return wide::fq2_mul_lazy<false>(a, b); }
B2_HD Fq2 tree_sqr(const Fq2& a) { return a.sqr_inline(); }
#else
B2_HD Fq2 tree_mul(const Fq2& a, const Fq2& b) { return a * b; }
B2_HD Fq2 tree_sqr(const Fq2& a) { return a.sqr(); }
#endif

// y coordinate of entry e, sign of a level-0 reference applied (the entry is known not to be the identity)
template <class F> B2_HD F tree_entry_y(const TreeIn<F>& in, uint32_t e) {
    if (in.sorted) {
        const uint32_t ix = in.sorted[e];
        const F y = tree_ld_field(&in.bases[ix & 0x7fffffffu].y);
        return (ix >> 31) ? y.neg() : y;
    }
    return tree_ld_field(&in.pts[e].y);
}

// What one output of a level is.  KIND_ADD / KIND_DBL need the inverse of `d`.
enum { TREE_COPY0 = 0, TREE_COPY1 = 1, TREE_ADD = 2, TREE_DBL = 3, TREE_ZERO = 4 };
template <class F> B2_HD int tree_classify(const Affine<F>& p0, const Affine<F>& p1, bool has_pair, F& d) {
    if (!has_pair || p1.is_identity()) return TREE_COPY0;
    if (p0.is_identity()) return TREE_COPY1;
    if (p0.x != p1.x) { d = p1.x - p0.x; return TREE_ADD; }
    if (p0.y == p1.y && !p0.y.is_zero()) { d = p0.y.dbl(); return TREE_DBL; }
    return TREE_ZERO;                                   // P + (-P), or doubling a point of order two
}

// largest b in [0, n_buckets) with off[b] <= o (off[n_buckets] > o)
B2_HD uint32_t tree_bucket_of(const uint32_t* off, uint32_t n_buckets, uint32_t o) {
    uint32_t lo = 0, hi = n_buckets;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (off[mid] <= o) lo = mid; else hi = mid; }
    return lo;
}
// number of values at depth `depth` of the product tree over e_out outputs: depth 1 = thread totals of up0
B2_HD uint32_t tree_count_at(uint32_t e_out, int depth, uint32_t K, uint32_t K2) {
    uint32_t n = (e_out + K - 1) / K;
    for (int i = 1; i < depth; ++i) n = (n + K2 - 1) / K2;
    return n;
}

// ---- up0: prefix products of the denominators of thread t's K outputs
template <class F> B2_HD void tree_up0_body(const TreeIn<F>& in, const uint32_t* off_out, uint32_t n_buckets, uint32_t K, uint32_t t,
                                            F* pp, F* tot) {
    const uint32_t e_out = off_out[n_buckets];
    const uint64_t o0 = (uint64_t)t * K;
    if (o0 >= e_out) return;
    const uint32_t o_end = (e_out - (uint32_t)o0 < K) ? e_out : (uint32_t)o0 + K;
    uint32_t b = tree_bucket_of(off_out, n_buckets, (uint32_t)o0);
    uint32_t next = off_out[b + 1];
    F prod = F::one();
    bool any = false;
    for (uint32_t o = (uint32_t)o0; o < o_end; ++o) {
        while (o >= next) { ++b; next = off_out[b + 1]; }
        const uint32_t i0 = in.off[b] + 2 * (o - off_out[b]);
        const bool has_pair = i0 + 1 < in.off[b + 1];
        tree_st_field(pp + o, prod);
        if (!has_pair) continue;
        // The denominator of an ordinary addition needs the two x coordinates only (half the bytes of the gather):
        // x != 0 rules out the identity, x0 != x1 the doubling / cancellation cases; everything else takes the full path.
        F d;
        int kind;
        const F x0 = tree_entry_x(in, i0), x1 = tree_entry_x(in, i0 + 1);
        if (!x0.is_zero() && !x1.is_zero() && x0 != x1) { d = x1 - x0; kind = TREE_ADD; }
        else {
            const Affine<F> p0 = tree_entry(in, i0), p1 = tree_entry(in, i0 + 1);
            kind = tree_classify(p0, p1, true, d);
        }
        if (kind == TREE_ADD || kind == TREE_DBL) { prod = any ? prod * d : d; any = true; }
    }
    tree_st_field(tot + t, prod);
}
// ---- up: the same on an array of values (the totals of the level below), K2 per thread
template <class F> B2_HD void tree_up_body(const F* v, uint32_t n, uint32_t K2, uint32_t t, F* pp, F* tot) {
    const uint64_t i0 = (uint64_t)t * K2;
    if (i0 >= n) return;
    const uint32_t i_end = (n - (uint32_t)i0 < K2) ? n : (uint32_t)i0 + K2;
    F prod = tree_ld_field(v + i0);
    tree_st_field(pp + i0, F::one());
    for (uint32_t i = (uint32_t)i0 + 1; i < i_end; ++i) {
        tree_st_field(pp + i, prod);
        prod = prod * tree_ld_field(v + i);
    }
    tree_st_field(tot + t, prod);
}
// ---- top: v[i] <- 1 / v[i] for the n <= TREE_TOP remaining values, one thread, one inversion
constexpr uint32_t TREE_TOP = 32;
template <class F> B2_HD void tree_top_body(F* v, uint32_t n, F* scratch /* n */) {
    if (n == 0) return;
    F prod = v[0];
    scratch[0] = F::one();
    for (uint32_t i = 1; i < n; ++i) { scratch[i] = prod; prod = prod * v[i]; }
    F I = prod.inv_gcd();
    for (uint32_t i = n; i-- > 0;) { const F x = v[i]; v[i] = I * scratch[i]; I = I * x; }
}
// ---- down: v[i] <- 1 / v[i] given inv_tot[t] = 1 / (product of thread t's values)
template <class F> B2_HD void tree_down_body(F* v, uint32_t n, uint32_t K2, uint32_t t, const F* pp, const F* inv_tot) {
    const uint64_t i0 = (uint64_t)t * K2;
    if (i0 >= n) return;
    const uint32_t i_end = (n - (uint32_t)i0 < K2) ? n : (uint32_t)i0 + K2;
    F I = tree_ld_field(inv_tot + t);
    for (uint32_t i = i_end; i-- > (uint32_t)i0;) {
        const F x = tree_ld_field(v + i);
        tree_st_field(v + i, I * tree_ld_field(pp + i));
        I = I * x;
    }
}
// ---- down0: the additions of thread t's K outputs
template <class F> B2_HD void tree_down0_body(const TreeIn<F>& in, const uint32_t* off_out, uint32_t n_buckets, uint32_t K, uint32_t t,
                                              const F* pp, const F* inv_tot, Affine<F>* out) {
    const uint32_t e_out = off_out[n_buckets];
    const uint64_t o0 = (uint64_t)t * K;
    if (o0 >= e_out) return;
    const uint32_t o_end = (e_out - (uint32_t)o0 < K) ? e_out : (uint32_t)o0 + K;
    uint32_t b = tree_bucket_of(off_out, n_buckets, o_end - 1);
    uint32_t first = off_out[b];
    F I = tree_ld_field(inv_tot + t);
    for (uint32_t o = o_end; o-- > (uint32_t)o0;) {
        while (o < first) { --b; first = off_out[b]; }
        const uint32_t i0 = in.off[b] + 2 * (o - first);
        const bool has_pair = i0 + 1 < in.off[b + 1];
        Affine<F> r;
        bool done = false;
        if (has_pair) {
            // ordinary addition (both x non-zero and different -- the same test as in up0): the coordinates are fetched
            // when they are needed and dropped as soon as they are used, which keeps the G2 form's live state at five
            // field elements instead of nine
            const F x0 = tree_entry_x(in, i0), x1 = tree_entry_x(in, i0 + 1);
            if (!x0.is_zero() && !x1.is_zero() && x0 != x1) {
                F lam = tree_mul(I, tree_ld_field(pp + o));  // 1 / (x1 - x0)
                I = tree_mul(I, x1 - x0);
                lam = tree_mul(tree_entry_y(in, i0 + 1) - tree_entry_y(in, i0), lam);
                r.x = tree_sqr(lam) - x0 - x1;
                r.y = tree_mul(lam, x0 - r.x) - tree_entry_y(in, i0);
                done = true;
            }
        }
        if (!done) {
            const Affine<F> p0 = tree_entry(in, i0);
            r = p0;
            if (has_pair) {
                const Affine<F> p1 = tree_entry(in, i0 + 1);
                F d;
                const int kind = tree_classify(p0, p1, true, d);
                if (kind == TREE_COPY1) r = p1;
                else if (kind == TREE_ZERO) r = Affine<F>::identity();
                else if (kind != TREE_COPY0) {
                    const F inv_d = I * tree_ld_field(pp + o);
                    I = I * d;
                    F lam;
                    if (kind == TREE_ADD) lam = (p1.y - p0.y) * inv_d;
                    else { const F xx = p0.x.sqr(); lam = (xx.dbl() + xx) * inv_d; }
                    r.x = lam.sqr() - p0.x - p1.x;           // doubling: p1.x == p0.x
                    r.y = lam * (p0.x - r.x) - p0.y;
                }
            }
        }
        out[o] = r;
    }
}

#if defined(__CUDACC__)
// ---- kernels: one thread per body call.  Grids are sized from host-side upper bounds of the entry counts (the true
// count is device data: off_out[n_buckets]); surplus threads return at once.
// CTAs per SM the big level kernels are compiled for.  The G2 forms sit at the 255-register cap (two warps per scheduler,
// 58 % of the multiplier rate in tree_down0); forcing 3 or 4 CTAs per SM (-DB2_TREE_G2_MINB=3 / 4: 168 / 128 registers,
// the rest spilled) was measured slower: 64.7 / 66.4 ms against 54.4 ms for the five down0 launches of a 2^22 multiexp.
#ifndef B2_TREE_G2_MINB
#define B2_TREE_G2_MINB 1
#endif
template <class F> struct TreeMinBlocks { static constexpr int value = sizeof(F) > 48 ? B2_TREE_G2_MINB : 4; };   // G1: 128 registers, four CTAs per SM
template <class F> __global__ void __launch_bounds__(128, TreeMinBlocks<F>::value) tree_up0_kernel(TreeIn<F> in, const uint32_t* __restrict__ off_out, uint32_t n_buckets, uint32_t K,
                                                                          F* __restrict__ pp, F* __restrict__ tot) {
    tree_up0_body<F>(in, off_out, n_buckets, K, blockIdx.x * blockDim.x + threadIdx.x, pp, tot);
}
template <class F> __global__ void __launch_bounds__(128, TreeMinBlocks<F>::value) tree_down0_kernel(TreeIn<F> in, const uint32_t* __restrict__ off_out, uint32_t n_buckets, uint32_t K,
                                                                            const F* __restrict__ pp, const F* __restrict__ inv_tot, Affine<F>* __restrict__ out) {
    tree_down0_body<F>(in, off_out, n_buckets, K, blockIdx.x * blockDim.x + threadIdx.x, pp, inv_tot, out);
}
template <class F> __global__ void __launch_bounds__(128) tree_up_kernel(const F* __restrict__ v, const uint32_t* __restrict__ e_out, int depth, uint32_t K, uint32_t K2,
                                                                         F* __restrict__ pp, F* __restrict__ tot) {
    tree_up_body<F>(v, tree_count_at(*e_out, depth, K, K2), K2, blockIdx.x * blockDim.x + threadIdx.x, pp, tot);
}
template <class F> __global__ void __launch_bounds__(128) tree_down_kernel(F* __restrict__ v, const uint32_t* __restrict__ e_out, int depth, uint32_t K, uint32_t K2,
                                                                           const F* __restrict__ pp, const F* __restrict__ inv_tot) {
    tree_down_body<F>(v, tree_count_at(*e_out, depth, K, K2), K2, blockIdx.x * blockDim.x + threadIdx.x, pp, inv_tot);
}
template <class F> __global__ void __launch_bounds__(32) tree_top_kernel(F* v, const uint32_t* __restrict__ e_out, int depth, uint32_t K, uint32_t K2, F* scratch) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const uint32_t n = tree_count_at(*e_out, depth, K, K2);
    tree_top_body<F>(v, n < TREE_TOP ? n : TREE_TOP, scratch);
}
#endif

}  // namespace b2
