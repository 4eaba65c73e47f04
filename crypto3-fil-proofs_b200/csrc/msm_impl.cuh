// Pippenger multi-scalar multiplication for G1 (F = Fq) and G2 (F = Fq2) on sm_100a.
// Replaces crypto3-algebra `multiexp<BDLO12>` / `multiexp_with_mixed_addition`
// (un-vendored) for the a / b_g1 / b_g2 / l / h queries of
// libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:46-66.
//
// Pipeline (all windows at once, one launch per step):
//   1. signed-digit decomposition (window c bits, W = ceil(256/c) windows, digit in
//      [-2^(c-1), 2^(c-1)], zero digits dropped) + histogram of (window, |digit|) buckets
//   2. exclusive scan of the histogram -> bucket offsets; a second scan of ceil(count/S)
//      -> segment offsets (a bucket with more than S entries is split, so a skewed
//      witness -- most Filecoin scalars are 0/1 -- cannot serialise on one thread)
//   3. counting-sort scatter of point indices (sign in bit 31) into bucket order
//   4. bucket accumulation: one thread per segment, XYZZ accumulator, mixed-affine adds
//      (8M + 2S each; INT32-multiply bound -- SURVEY §8d counts 10 Fq products per entry)
//   5. bucket reduction: each thread running-sums a chunk of adjacent buckets of one
//      window and lifts it by its chunk offset; a warp per window folds the chunks with
//      register shuffles (warp-cooperative reduction)
//   6. Horner over the windows (c doublings each), batch-free inversion, affine ABI bytes.
// Results are canonical affine points, hence bit-exact whatever the summation order.
#pragma once
#include "common.cuh"
#include <cstdlib>
#include <cstdio>
#include <utility>
#include <vector>
#include "msm_digits.cuh"
#include "msm_tree.cuh"

namespace b2 {

#if defined(__CUDACC__)
B2_D void load_scalar(const uint32_t* scalars, size_t i, uint32_t* s) {
    const uint4* q = reinterpret_cast<const uint4*>(scalars + 8 * i);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
}

// ---- 1. histogram ------------------------------------------------------------------
// The dense key layout (scheme_params.hpp:46-66) holds the identity for every variable whose A / B
// polynomial is zero -- a third or more of the a / b_g1 / b_g2 queries.  bellman filters those points out
// of its keys; here they are dropped with the zero digits, so they take no slot in the sorted lists and
// no lane in the accumulate kernel.  q = sizeof(Affine<F>) / 16.
B2_D bool base_is_identity(const uint4* __restrict__ bases_raw, size_t i, int q) {
    const uint4* p = bases_raw + i * (size_t)q;
    uint32_t acc = 0;
    for (int k = 0; k < q; ++k) { const uint4 v = __ldg(p + k); acc |= v.x | v.y | v.z | v.w; }
    return acc == 0;
}
static __global__ void msm_hist_kernel(const uint32_t* __restrict__ scalars, const uint4* __restrict__ bases_raw, int q, size_t n, int c, int W,
                                       uint32_t* __restrict__ hist) {
    const uint32_t B = 1u << (c - 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t s[8]; load_scalar(scalars, i, s);
        if ((s[0] | s[1] | s[2] | s[3] | s[4] | s[5] | s[6] | s[7]) == 0 || base_is_identity(bases_raw, i, q)) continue;   // zero scalar: base not read
        for (int w = 0; w < W; ++w) {
            int d = msm_signed_digit(s, w, c);
            if (d) atomicAdd(hist + (size_t)w * B + (uint32_t)((d < 0 ? -d : d) - 1), 1u);
        }
    }
}
// ---- 3. scatter --------------------------------------------------------------------
static __global__ void msm_scatter_kernel(const uint32_t* __restrict__ scalars, const uint4* __restrict__ bases_raw, int q, size_t n, int c, int W,
                                          uint32_t* __restrict__ cursor, uint32_t* __restrict__ sorted) {
    const uint32_t B = 1u << (c - 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t s[8]; load_scalar(scalars, i, s);
        if ((s[0] | s[1] | s[2] | s[3] | s[4] | s[5] | s[6] | s[7]) == 0 || base_is_identity(bases_raw, i, q)) continue;   // zero scalar: base not read
        for (int w = 0; w < W; ++w) {
            int d = msm_signed_digit(s, w, c);
            if (d) {
                uint32_t slot = atomicAdd(cursor + (size_t)w * B + (uint32_t)((d < 0 ? -d : d) - 1), 1u);
                sorted[slot] = (uint32_t)i | (d < 0 ? 0x80000000u : 0u);
            }
        }
    }
}
// ---- 2. scans (uint32, exclusive, with the total appended at out[n]) -----------------
constexpr int SCAN_BLOCK = 1024, SCAN_ITEMS = 8, SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;
// mode 0: out = in ; mode 1: out = ceil(in / seg) ; mode 2: `in` is an offsets array (n + 1 entries), out = scan of
// ceil((in[i+1] - in[i]) / 2) -- the entry counts of the next level of the pairwise tree (msm_tree.cuh)
static __global__ void __launch_bounds__(SCAN_BLOCK) scan_tile_kernel(const uint32_t* in, uint32_t* out, uint32_t* tile_sums, size_t n, int mode, uint32_t seg) {
    __shared__ uint32_t warp_sums[32];
    const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        uint32_t x = 0u;
        if (base + k < n) x = mode == 2 ? (in[base + k + 1] - in[base + k] + 1) / 2 : in[base + k];
        if (mode == 1) x = (x + seg - 1) / seg;
        v[k] = sum; sum += x;
    }
    uint32_t incl = sum;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = warp_sums[lane], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, d); if (lane >= d) wi += t; }
        warp_sums[lane] = wi - w;
        if (lane == 31) tile_sums[blockIdx.x] = wi;
    }
    __syncthreads();
    const uint32_t off = warp_sums[wid] + incl - sum;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) out[base + k] = off + v[k];
}
static __global__ void __launch_bounds__(1024) scan_top_kernel(uint32_t* tile_sums, int n_tiles, uint32_t* total_out) {
    // single block; n_tiles <= 1024 * 64
    __shared__ uint32_t sh[1024];
    const int per = (n_tiles + 1023) / 1024;
    uint32_t sum = 0;
    for (int k = 0; k < per; ++k) { int i = threadIdx.x * per + k; if (i < n_tiles) sum += tile_sums[i]; }
    sh[threadIdx.x] = sum;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        uint32_t t = threadIdx.x >= d ? sh[threadIdx.x - d] : 0u;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    uint32_t run = sh[threadIdx.x] - sum;
    for (int k = 0; k < per; ++k) { int i = threadIdx.x * per + k; if (i < n_tiles) { uint32_t x = tile_sums[i]; tile_sums[i] = run; run += x; } }
    if (threadIdx.x == 1023) *total_out = sh[1023];
}
static __global__ void scan_add_kernel(uint32_t* out, const uint32_t* tile_sums, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] += tile_sums[i / SCAN_TILE];
}

// ---- 4. bucket accumulation ----------------------------------------------------------
template <class F> B2_D Affine<F> load_affine(const Affine<F>* p) {
    Affine<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(Affine<F>) / 16); ++i) { uint4 t = __ldg(q + i); d[4 * i] = t.x; d[4 * i + 1] = t.y; d[4 * i + 2] = t.z; d[4 * i + 3] = t.w; }
    return r;
}
template <class F> B2_D Xyzz<F> xyzz_shfl_down(const Xyzz<F>& p, int delta) {
    Xyzz<F> r;
    const uint32_t* s = reinterpret_cast<const uint32_t*>(&p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(Xyzz<F>) / 4); ++i) d[i] = __shfl_down_sync(0xffffffffu, s[i], delta);
    return r;
}
B2_D void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <class F, int THREADS, int MINB, int PF> __global__ void __launch_bounds__(THREADS, MINB)
msm_accumulate_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ hist,
                      const uint32_t* __restrict__ off, const uint32_t* __restrict__ seg_off, uint32_t n_buckets, uint32_t S, Xyzz<F>* __restrict__ seg_acc) {
    const uint32_t seg_total = seg_off[n_buckets];
    for (uint32_t seg = blockIdx.x * blockDim.x + threadIdx.x; seg < seg_total; seg += gridDim.x * blockDim.x) {
        // bucket owning this segment: largest b with seg_off[b] <= seg
        uint32_t lo = 0, hi = n_buckets;
        while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (seg_off[mid] <= seg) lo = mid; else hi = mid; }
        const uint32_t b = lo;
        const uint32_t begin = off[b] + (seg - seg_off[b]) * S;
        const uint32_t bucket_end = off[b] + hist[b];
        const uint32_t end = begin + S < bucket_end ? begin + S : bucket_end;
        Xyzz<F> acc = Xyzz<F>::identity();
        if (PF == 0) {
            for (uint32_t e = begin; e < end; ++e) {
                const uint32_t ix = sorted[e];
                Affine<F> p = load_affine(bases + (ix & 0x7fffffffu));
                if (ix >> 31) p.y = p.y.neg();
                acc.add_mixed(p);
            }
        } else if (PF == 1) {
            // index look-ahead + L2 prefetch of the next point: the 96/192-byte gathers are
            // random, their DRAM latency otherwise sits on the critical path of every add
            uint32_t nxt = begin < end ? sorted[begin] : 0u;
            for (uint32_t e = begin; e < end; ++e) {
                const uint32_t ix = nxt;
                if (e + 1 < end) {
                    nxt = sorted[e + 1];
                    const char* q = reinterpret_cast<const char*>(bases + (nxt & 0x7fffffffu));
                    prefetch_l2(q);
                    prefetch_l2(q + sizeof(Affine<F>) - 4);
                }
                Affine<F> p = load_affine(bases + (ix & 0x7fffffffu));
                if (ix >> 31) p.y = p.y.neg();
                acc.add_mixed(p);
            }
        } else {
            // register double-buffer: the next point is loaded while the current add runs
            uint32_t ix = begin < end ? sorted[begin] : 0u;
            Affine<F> p = load_affine(bases + (ix & 0x7fffffffu));
            for (uint32_t e = begin; e < end; ++e) {
                const uint32_t cix = ix;
                Affine<F> cur = p;
                if (e + 1 < end) { ix = sorted[e + 1]; p = load_affine(bases + (ix & 0x7fffffffu)); }
                if (cix >> 31) cur.y = cur.y.neg();
                acc.add_mixed(cur);
            }
        }
        seg_acc[seg] = acc;
    }
}
// ---- 4'. equal-work bucket accumulation ------------------------------------------------
// The sorted entry array is cut into one contiguous chunk of L entries per thread, whatever the
// bucket boundaries: every lane of a warp performs the same number of mixed additions (the
// one-thread-per-bucket kernel above loses ~16 % of its SIMD slots to Poisson bucket sizes at 128
// entries per bucket and ~45 % at 32, which is what made wide windows unprofitable).
// A bucket that lies inside one chunk is written straight to bucket_acc[b]; a bucket cut by chunk
// boundaries leaves one "tail" piece (its first part, at the end of a chunk) and "head" pieces (its
// continuations, at the start of later chunks) that msm_fix_pieces_kernel folds afterwards.
constexpr uint32_t NO_BUCKET = 0xffffffffu;
constexpr uint32_t RED_BITS_MAX = 16384;      // items per window at which the bucket reduction switches to the bit-decomposition tail

// DIRECT: the entries are the points themselves (bases[e], the output of the affine tree levels, identities possible)
// instead of (index, sign) references into the key.
template <class F, bool DIRECT> B2_D Affine<F> chunk_entry_point(const Affine<F>* __restrict__ bases, uint32_t ix, uint32_t e) {
    if (DIRECT) return load_affine(bases + e);
    Affine<F> p = load_affine(bases + (ix & 0x7fffffffu));
    if (ix >> 31) p.y = p.y.neg();
    return p;
}
template <class F, int THREADS, int MINB, bool DIRECT = false> __global__ void __launch_bounds__(THREADS, MINB)
msm_accumulate_chunked_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ off,
                              uint32_t n_buckets, Xyzz<F>* __restrict__ bucket_acc, Xyzz<F>* __restrict__ head_part, Xyzz<F>* __restrict__ tail_part,
                              uint32_t* __restrict__ head_bucket, uint32_t* __restrict__ tail_bucket) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, n_thr = gridDim.x * blockDim.x;
    const uint32_t E = off[n_buckets];
    uint32_t L = (E + n_thr - 1) / n_thr;
    if (L == 0) L = 1;
    const uint64_t begin64 = (uint64_t)tid * L;
    uint32_t hb = NO_BUCKET, tb = NO_BUCKET;
    if (begin64 < E) {
        const uint32_t begin = (uint32_t)begin64;
        const uint32_t end = (E - begin < L) ? E : begin + L;
        uint32_t lo = 0, hi = n_buckets;                 // largest b with off[b] <= begin
        while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (off[mid] <= begin) lo = mid; else hi = mid; }
        uint32_t b = lo, next = off[b + 1];
        bool head_open = off[b] < begin;                 // first bucket started in an earlier chunk
        Xyzz<F> acc = Xyzz<F>::identity();
        uint32_t e = begin;
        uint32_t ix = DIRECT ? 0u : sorted[e];
        while (e < end) {
            if (e == next) {                             // bucket b is complete: flush, move on
                if (head_open) { head_part[tid] = acc; hb = b; head_open = false; } else bucket_acc[b] = acc;
                acc = Xyzz<F>::identity();
                do { ++b; next = off[b + 1]; } while (next == e);
            }
            Affine<F> p = chunk_entry_point<F, DIRECT>(bases, ix, e);
            ++e;
            if (!DIRECT && e < end) {
                ix = sorted[e];
                const char* q = reinterpret_cast<const char*>(bases + (ix & 0x7fffffffu));
                prefetch_l2(q); prefetch_l2(q + sizeof(Affine<F>) - 4);
            }
            if (acc.is_identity() && !p.is_identity()) {
                // opening a bucket costs no addition slot: take the point and, if the bucket goes on
                // inside this chunk, fall through to add the next one in the same trip
                acc.X = p.x; acc.Y = p.y; acc.ZZ = F::one(); acc.ZZZ = F::one();
                if (e >= end || e == next) continue;
                p = chunk_entry_point<F, DIRECT>(bases, ix, e);
                ++e;
                if (!DIRECT && e < end) {
                    ix = sorted[e];
                    const char* q = reinterpret_cast<const char*>(bases + (ix & 0x7fffffffu));
                    prefetch_l2(q); prefetch_l2(q + sizeof(Affine<F>) - 4);
                }
            }
            acc.add_mixed(p);
        }
        // last bucket of the chunk: complete iff it ends exactly here
        const bool goes_on = next > end;
        if (head_open) { head_part[tid] = acc; hb = b; }             // chunk lies inside one bucket, or its only piece so far
        else if (goes_on) { tail_part[tid] = acc; tb = b; }
        else bucket_acc[b] = acc;
        if (head_open && !goes_on) { /* the bucket that began earlier ends here: still a head piece */ }
    }
    head_bucket[tid] = hb;
    tail_bucket[tid] = tb;
}
// One warp per chunk t whose last bucket continues: bucket = tail piece of t + head pieces of t+1..
template <class F> __global__ void __launch_bounds__(128)
msm_fix_pieces_kernel(Xyzz<F>* __restrict__ bucket_acc, const Xyzz<F>* __restrict__ head_part, const Xyzz<F>* __restrict__ tail_part,
                      const uint32_t* __restrict__ head_bucket, const uint32_t* __restrict__ tail_bucket, uint32_t n_chunks) {
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < n_chunks; t += (gridDim.x * blockDim.x) >> 5) {
        const uint32_t b = tail_bucket[t];
        if (b == NO_BUCKET) continue;                    // warp-uniform
        Xyzz<F> acc = Xyzz<F>::identity();
        {   // the usual case -- a bucket cut once or twice -- needs one or two additions, not a 32-lane tree of six
            const bool mine0 = t + 1 + lane < n_chunks && head_bucket[t + 1 + lane] == b;
            const uint32_t mask = __ballot_sync(0xffffffffu, mine0);
            const int m = mask == 0xffffffffu ? 32 : __ffs((int)~mask) - 1;      // continuation pieces are contiguous from chunk t + 1
            if (m <= 3) {
                if (lane == 0) {
                    acc = tail_part[t];
                    for (int j = 0; j < m; ++j) { Xyzz<F> q = head_part[t + 1 + j]; xyzz_add_call(&acc, &q); }
                    bucket_acc[b] = acc;
                }
                continue;
            }
        }
        if (lane == 0) acc = tail_part[t];
        // continuation pieces are contiguous chunks t+1, t+2, .. carrying the same bucket id
        for (uint32_t u = t + 1 + lane; ; u += 32) {
            const bool mine = u < n_chunks && head_bucket[u] == b;
            if (mine) { Xyzz<F> q = head_part[u]; xyzz_add_call(&acc, &q); }
            if (!__any_sync(0xffffffffu, mine)) break;
            if (!__all_sync(0xffffffffu, mine)) break;   // pieces are contiguous: a gap ends the run
        }
        for (int d = 16; d >= 1; d >>= 1) {
            Xyzz<F> o = xyzz_shfl_down(acc, d);
            xyzz_add_call(&acc, &o);
        }
        if (lane == 0) bucket_acc[b] = acc;
    }
}

// a bucket that was split into several segments: one warp folds them into its first slot
template <class F> __global__ void __launch_bounds__(128)
msm_fold_segments_kernel(Xyzz<F>* __restrict__ seg_acc, const uint32_t* __restrict__ seg_off, uint32_t n_buckets) {
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < n_buckets; b += (gridDim.x * blockDim.x) >> 5) {
        const uint32_t s0 = seg_off[b], s1 = seg_off[b + 1];
        if (s1 - s0 <= 1) continue;                      // warp-uniform
        Xyzz<F> acc = Xyzz<F>::identity();
        for (uint32_t s = s0 + lane; s < s1; s += 32) { Xyzz<F> t = seg_acc[s]; xyzz_add_call(&acc, &t); }
        for (int d = 16; d >= 1; d >>= 1) {
            Xyzz<F> o = xyzz_shfl_down(acc, d);
            xyzz_add_call(&acc, &o);
        }
        __syncwarp();
        if (lane == 0) seg_acc[s0] = acc;
    }
}

// ---- 5. bucket reduction -------------------------------------------------------------
template <class F> __global__ void __launch_bounds__(128)
msm_reduce_kernel(const Xyzz<F>* __restrict__ seg_acc, const uint32_t* __restrict__ seg_off, int c, int W, uint32_t T, Xyzz<F>* __restrict__ partial) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (uint32_t)W * T) return;
    const uint32_t w = gid / T, t = gid % T, B = 1u << (c - 1), per = B / T;
    const uint32_t lo = t * per;
    Xyzz<F> run = Xyzz<F>::identity(), acc = Xyzz<F>::identity();
    for (uint32_t j = lo + per; j-- > lo;) {                    // weights j+1, descending
        const uint32_t b = w * B + j;
        if (seg_off == nullptr) { Xyzz<F> t = seg_acc[b]; xyzz_add_call(&run, &t); }
        else if (seg_off[b + 1] > seg_off[b]) { Xyzz<F> t = seg_acc[seg_off[b]]; xyzz_add_call(&run, &t); }
        xyzz_add_call(&acc, &run);
    }
    // acc = sum (j - lo + 1) B_j ;  add lo * run (double-and-add over the bits of lo)
    Xyzz<F> m = Xyzz<F>::identity();
    for (int bit = 31; bit >= 0; --bit) {
        xyzz_dbl_call(&m);
        if ((lo >> bit) & 1u) xyzz_add_call(&m, &run);
    }
    xyzz_add_call(&acc, &m);
    partial[gid] = acc;
}
// one warp per window: fold the T chunk results
template <class F> __global__ void __launch_bounds__(32)
msm_window_sum_kernel(const Xyzz<F>* __restrict__ partial, uint32_t T, Xyzz<F>* __restrict__ win) {
    const uint32_t w = blockIdx.x, lane = threadIdx.x;
    Xyzz<F> acc = Xyzz<F>::identity();
    for (uint32_t i = lane; i < T; i += 32) { Xyzz<F> t = partial[w * T + i]; xyzz_add_call(&acc, &t); }
    for (int d = 16; d >= 1; d >>= 1) {
        Xyzz<F> o = xyzz_shfl_down(acc, d);
        xyzz_add_call(&acc, &o);
    }
    if (lane == 0) win[w] = acc;
}
// ---- 5'. bucket reduction, recursive chunked running sums -------------------------------
// Window sum = sum_{d=1..B} d * bucket[d-1].  One level turns `count` items (a_i, r_i) with the value
//   V = sum_i a_i + 2^log_s * sum_i i * r_i
// into T = ceil(count / per) items of the same form with log_s' = log_s + log2(per):
//   a'_t = sum_j a_{t,j} + 2^log_s * sum_j j * r_{t,j}   (descending running sum inside the chunk)
//   r'_t = sum_j r_{t,j}
// because i = t*per + j.  Level 1 reads the buckets as items 1..B (item 0 = identity, no a_i); the
// recursion ends with one chunk per window, whose a' is the window sum.  Every level is fully
// parallel (one thread per chunk, all windows in one launch): 2 additions per bucket at level 1,
// 1/per of that at each further level.
template <class F> __global__ void __launch_bounds__(128)
msm_reduce_level_kernel(const Xyzz<F>* __restrict__ a_in, const Xyzz<F>* __restrict__ r_in, uint32_t count, uint32_t in_stride, int shift_one,
                        uint32_t per, int log_s, Xyzz<F>* __restrict__ a_out, Xyzz<F>* __restrict__ r_out, uint32_t T, int W) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (uint32_t)W * T) return;
    const uint32_t w = gid / T, t = gid % T;
    const uint32_t lo = t * per, hi = lo + per < count ? lo + per : count;
    const Xyzz<F>* rw = r_in + (size_t)w * in_stride;
    Xyzz<F> run = Xyzz<F>::identity(), acc = Xyzz<F>::identity(), asum = Xyzz<F>::identity();
    for (uint32_t i = hi; i-- > lo;) {
        if (!shift_one) { Xyzz<F> r = rw[i]; xyzz_add_call(&run, &r); }
        else if (i > 0) { Xyzz<F> r = rw[i - 1]; xyzz_add_call(&run, &r); }
        if (i > lo) xyzz_add_call(&acc, &run);
        if (a_in) { Xyzz<F> a = a_in[(size_t)w * in_stride + i]; xyzz_add_call(&asum, &a); }
    }
    for (int k = 0; k < log_s; ++k) xyzz_dbl_call(&acc);
    xyzz_add_call(&asum, &acc);
    a_out[(size_t)w * T + t] = asum;
    r_out[(size_t)w * T + t] = run;
}

// ---- 5''. tail of the bucket reduction by bit decomposition ---------------------------------
// V = sum_t a_t + 2^log_s sum_t t r_t over `count` items per window.  sum_t t r_t = sum_b 2^b S_b with
// S_b = sum of the r_t whose index has bit b set: NB independent plain sums (plus one over the a_t), each a
// CTA-wide tree -- ~16 dependent group operations -- where every further level of the recursion above is a
// chain of ~30 on a handful of threads.  The work doubles (NB / 2 + 1 additions per item instead of 3), which
// is why the big first level(s) stay with the running-sum form.
// grid (NB + 1, W); block 256.  CTA (b, w): b < NB -> S_b of window w, b == NB -> sum of the a_t.
template <class F> __global__ void __launch_bounds__(256)
msm_reduce_bits_kernel(const Xyzz<F>* __restrict__ a_in, const Xyzz<F>* __restrict__ r_in, uint32_t count, uint32_t stride, int NB, Xyzz<F>* __restrict__ sums) {
    __shared__ Xyzz<F> warp_part[8];
    const int b = blockIdx.x, w = blockIdx.y;
    const Xyzz<F>* src = (b == NB ? a_in : r_in) + (size_t)w * stride;
    Xyzz<F> acc = Xyzz<F>::identity();
    if (b == NB) {
        for (uint32_t t = threadIdx.x; t < count; t += blockDim.x) { Xyzz<F> v = src[t]; xyzz_add_call(&acc, &v); }
    } else {
        for (uint32_t j = threadIdx.x;; j += blockDim.x) {
            const uint32_t t = msm_bits_index(j, b);                       // j-th index with bit b set
            if (t >= count) break;                       // t grows with j
            Xyzz<F> v = src[t]; xyzz_add_call(&acc, &v);
        }
    }
    for (int d = 16; d >= 1; d >>= 1) { Xyzz<F> o = xyzz_shfl_down(acc, d); xyzz_add_call(&acc, &o); }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) warp_part[wid] = acc;
    __syncthreads();
    if (wid == 0) {
        acc = lane < 8 ? warp_part[lane] : Xyzz<F>::identity();
        for (int d = 4; d >= 1; d >>= 1) { Xyzz<F> o = xyzz_shfl_down(acc, d); xyzz_add_call(&acc, &o); }
        if (lane == 0) sums[(size_t)w * (NB + 1) + b] = acc;
    }
}
// one warp per window: lane b doubles S_b b times (all lanes at once), a shuffle tree adds them, then the 2^log_s
// factor and the sum of the a_t
template <class F> __global__ void __launch_bounds__(32)
msm_reduce_bits_final_kernel(const Xyzz<F>* __restrict__ sums, int NB, int log_s, Xyzz<F>* __restrict__ win) {
    const int w = blockIdx.x, lane = threadIdx.x;
    Xyzz<F> acc = lane < NB ? sums[(size_t)w * (NB + 1) + lane] : Xyzz<F>::identity();
    for (int k = 0; k < NB; ++k) if (k < lane) xyzz_dbl_call(&acc);
    for (int d = 16; d >= 1; d >>= 1) { Xyzz<F> o = xyzz_shfl_down(acc, d); xyzz_add_call(&acc, &o); }
    if (lane == 0) {
        for (int k = 0; k < log_s; ++k) xyzz_dbl_call(&acc);
        Xyzz<F> a = sums[(size_t)w * (NB + 1) + NB];
        xyzz_add_call(&acc, &a);
        win[w] = acc;
    }
}

// ---- 6. Horner over windows + affine output --------------------------------------------
B2_D void egress_fq(uint8_t* p, const Fq& m) {
    Fq c; m.to_canonical(c.v);
    uint32_t* q = reinterpret_cast<uint32_t*>(p);
#pragma unroll
    for (int i = 0; i < 12; ++i) q[i] = c.v[i];
}
B2_D void egress_affine(uint8_t* p, const G1Affine& a) { egress_fq(p, a.x); egress_fq(p + 48, a.y); }
B2_D void egress_affine(uint8_t* p, const G2Affine& a) { egress_fq(p, a.x.c0); egress_fq(p + 48, a.x.c1); egress_fq(p + 96, a.y.c0); egress_fq(p + 144, a.y.c1); }

template <class F> __global__ void __launch_bounds__(32)
msm_final_kernel(const Xyzz<F>* __restrict__ win, int c, int W, uint8_t* __restrict__ out_bytes) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    // ~256 dependent doublings by one thread are the fixed cost of every standalone multiexp: the doubling is inlined here
    // (one call site in a loop; the out-of-line form moves the accumulator through local memory on every call) and the
    // final inversion uses the Euclidean algorithm instead of the Fermat ladder
    Xyzz<F> acc = win[W - 1];
    for (int w = W - 2; w >= 0; --w) {
#pragma unroll 1
        for (int k = 0; k < c; ++k) acc = acc.dbl();
        Xyzz<F> t = win[w];
        xyzz_add_call(&acc, &t);
    }
    const Affine<F> aff = acc.to_affine_gcd();
    egress_affine(out_bytes, aff);
}
// sum of n affine points (combining per-rank partial results), single thread
template <class F> __global__ void __launch_bounds__(32)
sum_points_kernel(const Affine<F>* __restrict__ pts, size_t n, uint8_t* __restrict__ out_bytes) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    Xyzz<F> acc = Xyzz<F>::identity();
    for (size_t i = 0; i < n; ++i) { Affine<F> t = pts[i]; xyzz_add_mixed_call(&acc, &t); }
    const Affine<F> aff = acc.to_affine_gcd();
    egress_affine(out_bytes, aff);
}

// prime-order subgroup membership r * P == O for each point (verifying-key IC points, as bellman's VerifyingKey::read
// checks them); sets bit 2 of *bad.  Double-and-add over the 255 bits of r, one thread per point.
template <class F> __global__ void __launch_bounds__(64) subgroup_check_kernel(const Affine<F>* __restrict__ pts, size_t n, int* bad) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const Affine<F> p = pts[i];
        Xyzz<F> acc = Xyzz<F>::identity();
        for (int bit = 254; bit >= 0; --bit) {
            xyzz_dbl_call(&acc);
            if ((FrParams::P(bit >> 5) >> (bit & 31)) & 1u) xyzz_add_mixed_call(&acc, &p);
        }
        if (!acc.is_identity()) atomicOr(bad, 4);
    }
}

// ---- fixed-base multiples of the group generator (synthetic bases, Groth16 setup) -------
// table[w * 255 + (d-1)] = d * 2^(8w) * G, affine, w < 32
template <class F> __global__ void __launch_bounds__(32) fixed_base_roots_kernel(Affine<F> gen, Affine<F>* roots /*32*/) {
    const int w = threadIdx.x;
    Xyzz<F> p = Xyzz<F>::from_affine(gen);
    for (int k = 0; k < 8 * w; ++k) xyzz_dbl_call(&p);
    Affine<F> aff; xyzz_to_affine_call(&aff, &p);
    roots[w] = aff;
}
template <class F> __global__ void __launch_bounds__(64) fixed_base_table_kernel(const Affine<F>* roots, Affine<F>* table) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= 32 * 255) return;
    const int w = id / 255, d = id % 255 + 1;
    const Affine<F> base = roots[w];
    Xyzz<F> p = Xyzz<F>::identity();
    for (int bit = 7; bit >= 0; --bit) {
        xyzz_dbl_call(&p);
        if ((d >> bit) & 1) xyzz_add_mixed_call(&p, &base);
    }
    Affine<F> aff; xyzz_to_affine_call(&aff, &p);
    table[id] = aff;
}
template <class F> __global__ void __launch_bounds__(128) gen_mul_kernel(const Affine<F>* __restrict__ table, const uint32_t* __restrict__ scalars, size_t n, Affine<F>* __restrict__ out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t s[8]; load_scalar(scalars, i, s);
        Xyzz<F> acc = Xyzz<F>::identity();
        for (int w = 0; w < 32; ++w) {
            const uint32_t d = (s[w >> 2] >> (8 * (w & 3))) & 0xffu;
            if (d) { Affine<F> t = load_affine(table + w * 255 + (d - 1)); xyzz_add_mixed_call(&acc, &t); }
        }
        Affine<F> aff; xyzz_to_affine_call(&aff, &acc);
        out[i] = aff;
    }
}

// ---- host driver ---------------------------------------------------------------------
// Window width by size (measured on B200, uniform scalars, chunked accumulate): every window costs one mixed
// addition per point, every bucket two full additions in the reduction, and the counting-sort scatter gets less
// local as buckets multiply -- 16 bits up to 2^22 points, 20 bits from 2^23.  Standalone, 17 bits is 1.5 % faster
// for 2^20 .. 2^23 points (profiles/r01_window_sweep2_g{1,2}.jsonl), but inside a proof it moves the G2 multiexp
// onto the chunked kernel, whose few long CTAs co-schedule badly with the other four multiexps running on
// sibling streams (2^22 proof: 0.261 s against 0.241 s); 17 bits for G1 alone (B2G16_G1_C17=1) makes each
// accumulate launch 1.6 ms shorter and the proof 7 ms longer (0.248 s).  The proof-friendly choice stays.
inline bool g1_17bit_windows() { static const bool on = getenv("B2G16_G1_C17") != nullptr; return on; }
inline int msm_pick_window(size_t n, bool g2 = false) {
    int lg = 0;
    while (((size_t)1 << (lg + 1)) <= n) ++lg;
    if (lg >= 23) return 20;
    if (!g2 && lg >= 20 && g1_17bit_windows()) return 17;          // experiment knob B2G16_G1_C17 (see above)
    int c = lg - 4;
    if (c < 8) c = 8;
    if (c > 16) c = 16;
    return c;
}

static void scan_u32(Ctx* ctx, cudaStream_t st, const uint32_t* in, uint32_t* out /*n+1*/, size_t n, int mode, uint32_t seg, uint32_t* tile_sums) {
    const int n_tiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
    B2_REQUIRE(n_tiles <= 1024 * 64, B2G16_E_INVALID_ARG, "msm: too many buckets for the scan");
    scan_tile_kernel<<<n_tiles, SCAN_BLOCK, 0, st>>>(in, out, tile_sums, n, mode, seg);
    scan_top_kernel<<<1, 1024, 0, st>>>(tile_sums, n_tiles, out + n);
    scan_add_kernel<<<grid_for(n, 256), 256, 0, st>>>(out, tile_sums, n);
    ctx->launches += 3;
    B2_CUDA(cudaGetLastError());
}

template <class F, int MINB, int PF>
static void launch_accumulate_v(cudaStream_t st, size_t max_segs, const Affine<F>* bases, const uint32_t* sorted, const uint32_t* hist, const uint32_t* off,
                                const uint32_t* seg_off, uint32_t nb, uint32_t S, Xyzz<F>* seg_acc) {
    constexpr int THREADS = sizeof(F) > 48 ? 64 : 128;
    msm_accumulate_kernel<F, THREADS, MINB, PF><<<grid_for(max_segs, THREADS, 148 * 64), THREADS, 0, st>>>(bases, sorted, hist, off, seg_off, nb, S, seg_acc);
}
// one thread per bucket segment (the path of G2 at <= 16-bit windows when the tree levels are switched off, and of
// acc_variant < 100).  The prefetch / occupancy variants of round 1 (PF 1 / 2, 3-4 CTAs per SM) were measured slower
// and are no longer instantiated.
template <class F>
static void launch_accumulate(Ctx* ctx, cudaStream_t st, size_t max_segs, const Affine<F>* bases, const uint32_t* sorted, const uint32_t* hist, const uint32_t* off,
                              const uint32_t* seg_off, uint32_t nb, uint32_t S, Xyzz<F>* seg_acc) {
    ctx->launches++;
    launch_accumulate_v<F, 1, 0>(st, max_segs, bases, sorted, hist, off, seg_off, nb, S, seg_acc);
}

// d_win_out == nullptr: the affine result goes to d_out_bytes (Horner over the windows by one device thread).
// d_win_out != nullptr: the W window sums are copied there instead and *plan_out = {c, W}; the caller combines
// them (the prover does it on the host: 240 dependent doublings are 3 ms for a lone GPU thread and 0.3 ms
// for a CPU core).  W = 0 means the empty sum.
struct MsmPlan { int c, W; };
template <class F>
void msm_run(Ctx* ctx, const Affine<F>* d_bases, const uint32_t* d_scalars, size_t n, uint8_t* d_out_bytes, cudaStream_t st, int window_bits,
             Xyzz<F>* d_win_out = nullptr, MsmPlan* plan_out = nullptr) {
    constexpr int OUT_BYTES = (int)sizeof(Affine<F>);
    if (n == 0) {
        if (d_win_out) { plan_out->c = 0; plan_out->W = 0; return; }
        B2_CUDA(cudaMemsetAsync(d_out_bytes, 0, OUT_BYTES, st)); return;
    }
    B2_REQUIRE(n < ((size_t)1 << 31), B2G16_E_INVALID_ARG, "msm: at most 2^31-1 points per call");
    int c = window_bits ? window_bits : (ctx->msm_window_override ? ctx->msm_window_override : msm_pick_window(n, sizeof(F) > 48));
    B2_REQUIRE(c >= 8 && c <= 23, B2G16_E_INVALID_ARG, "msm: window bits must be in [8, 23]");
    const int W = msm_windows(c);
    const uint32_t B = 1u << (c - 1);
    const size_t nb = (size_t)W * B;
    B2_REQUIRE((double)n * W < 4.0e9, B2G16_E_INVALID_ARG, "msm: n * windows exceeds 32-bit entry offsets");
    const size_t max_entries = n * (size_t)W;
    uint32_t mean = (uint32_t)(n >> (c - 1));
    const uint32_t S = 4 * mean + 256;
    const size_t max_segs = nb + max_entries / S + 1;
    const int n_tiles = (int)((nb + SCAN_TILE - 1) / SCAN_TILE);

    // Levels of the batched-affine pairwise tree that run before the XYZZ accumulate (msm_tree.cuh): each level halves
    // the entries of every bucket at ~6 field products per addition instead of 10; stop when ~4 entries per bucket are
    // left (the tail levels would be launch- and inversion-latency bound).  B2G16_TREE_LEVELS / "tree_levels": -1 = by
    // mean bucket occupancy, 0 = off.
    int tree_levels = ctx->tree_levels;
    if (tree_levels < 0) {
        tree_levels = 0;
        for (uint32_t m = mean; m >= 8; m >>= 1) ++tree_levels;
    }
    if (tree_levels > 12) tree_levels = 12;
    // G2 at 16-bit windows: the one-thread-per-bucket kernel is faster (measured 86 vs 99 ms at 2^22)
    const bool chunked = (ctx->acc_variant >= 100 && !(sizeof(F) > 48 && c <= 16)) || tree_levels > 0;
    // chunked-accumulate launch shape (threads per CTA, CTAs per SM).  G1 at 194 registers per thread: 2 x 128 threads per
    // SM (variant 102, the default); 101 / 103 = 1 / 3 CTAs per SM.  160- and 96-thread CTAs and 4 CTAs per SM were
    // measured slower in round 1 and are gone.  G2: 64-thread CTAs at the 255-register cap, 4 per SM.
    int ch_threads = 0, ch_minb = 0;
    if (chunked) {
        if (sizeof(F) > 48) { ch_threads = 64; ch_minb = 4; }
        else if (ctx->acc_variant < 100 || tree_levels > 0) { ch_threads = 128; ch_minb = 2; }
        else { ch_threads = 128; ch_minb = ctx->acc_variant == 101 ? 1 : ctx->acc_variant == 103 ? 3 : 2; }
    }
    const uint32_t n_chunks = chunked ? (uint32_t)(148 * ch_minb * ch_threads) : 0;
    DevBuf<uint32_t> hist(nb + 1, st), off(nb + 1, st), cursor(nb + 1, st), seg_off(chunked ? 0 : nb + 1, st), tile_sums((size_t)n_tiles + 1, st), sorted(max_entries, st);
    DevBuf<Xyzz<F>> seg_acc(chunked ? nb : max_segs, st), pieces(2 * (size_t)n_chunks, st);
    DevBuf<uint32_t> piece_bucket(2 * (size_t)n_chunks, st);
    // items per thread and level in the bucket reduction (tuning knob B2G16_RED_LOG, default 2^3): the reduction is
    // latency-bound (a chain of 2 * RED_PER dependent additions per thread and level), so short chains on many
    // threads beat long chains on few
    const int red_log = ctx->red_log;
    const uint32_t RED_PER = 1u << red_log;
    const uint32_t T1 = (B + 1 + RED_PER - 1) / RED_PER;
    DevBuf<Xyzz<F>> red_a[2] = {DevBuf<Xyzz<F>>((size_t)W * T1, st), DevBuf<Xyzz<F>>((size_t)W * T1, st)};
    DevBuf<Xyzz<F>> red_r[2] = {DevBuf<Xyzz<F>>((size_t)W * T1, st), DevBuf<Xyzz<F>>((size_t)W * T1, st)};

    cudaEvent_t ev[6] = {};
    const bool timed = ctx->time_msm;
    if (timed) for (auto& e : ev) B2_CUDA(cudaEventCreate(&e));
    auto mark = [&](int i) { if (timed) B2_CUDA(cudaEventRecord(ev[i], st)); };
    // B2G16_MSM_TRACE=1 (with timing on): a CUDA event after every step of the accumulation / reduction, printed to
    // stderr once the multiexp is done -- the step-by-step split that ncu's serialised per-launch times cannot give
    static const bool trace_on = getenv("B2G16_MSM_TRACE") != nullptr;
    std::vector<std::pair<const char*, cudaEvent_t>> tr;
    auto tmark = [&](const char* label) {
        if (!(timed && trace_on)) return;
        cudaEvent_t e; B2_CUDA(cudaEventCreate(&e)); B2_CUDA(cudaEventRecord(e, st)); tr.push_back({label, e});
    };

    mark(0);
    B2_CUDA(cudaMemsetAsync(hist.get(), 0, (nb + 1) * sizeof(uint32_t), st));
    static_assert(sizeof(Affine<F>) % 16 == 0, "bases are read as uint4");
    const uint4* bases_raw = reinterpret_cast<const uint4*>(d_bases);
    const int base_q = (int)(sizeof(Affine<F>) / 16);
    msm_hist_kernel<<<grid_for(n, 256, 148 * 8), 256, 0, st>>>(d_scalars, bases_raw, base_q, n, c, W, hist.get()); ctx->launches++;
    mark(1);
    scan_u32(ctx, st, hist.get(), off.get(), nb, 0, 1, tile_sums.get());
    if (!chunked) scan_u32(ctx, st, hist.get(), seg_off.get(), nb, 1, S, tile_sums.get());
    B2_CUDA(cudaMemcpyAsync(cursor.get(), off.get(), (nb + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    mark(2);
    msm_scatter_kernel<<<grid_for(n, 256, 148 * 8), 256, 0, st>>>(d_scalars, bases_raw, base_q, n, c, W, cursor.get(), sorted.get()); ctx->launches++;
    mark(3);
    Ctx::KernelLogEntry kl{sizeof(F) > 48 ? 2 : 1, c, (uint64_t)n, nullptr, nullptr};
    if (ctx->klog_on) { B2_CUDA(cudaEventCreate(&kl.e0)); B2_CUDA(cudaEventCreate(&kl.e1)); }
    if (chunked) {
        B2_CUDA(cudaMemsetAsync(seg_acc.get(), 0, nb * sizeof(Xyzz<F>), st));            // empty bucket = identity (ZZ = 0)
        Xyzz<F>* head = pieces.get(); Xyzz<F>* tail = pieces.get() + n_chunks;
        uint32_t* hb = piece_bucket.get(); uint32_t* tb = piece_bucket.get() + n_chunks;
        if (kl.e0) B2_CUDA(cudaEventRecord(kl.e0, st));
        bool tree_fell_back = false;
        if (tree_levels > 0) {
            // ---- affine tree levels, then the XYZZ accumulate over what is left; windows in groups that fit the memory
            const uint32_t K = ctx->tree_k, K2 = ctx->tree_k2;
            auto caps = [&](size_t entries, uint32_t nbg, size_t* cap /*levels*/) {
                size_t e = entries;
                for (int l = 0; l < tree_levels; ++l) { e = (e + nbg) / 2 + 1; cap[l] = e; }
            };
            auto group_bytes = [&](int wg) {
                size_t cap[16]; caps(n * (size_t)wg, (uint32_t)wg * B, cap);
                const size_t c0 = cap[0], c1 = tree_levels > 1 ? cap[1] : 0;
                return c0 * (sizeof(Affine<F>) + sizeof(F)) + c1 * sizeof(Affine<F>) + 3 * (c0 / K + 64) * sizeof(F) + 2 * ((size_t)wg * B + 1) * sizeof(uint32_t);
            };
            size_t free_b = 0, total_b = 0;
            B2_CUDA(cudaMemGetInfo(&free_b, &total_b));
            cudaMemPool_t pool; uint64_t reserved = 0, used = 0;
            B2_CUDA(cudaDeviceGetDefaultMemPool(&pool, ctx->device));
            B2_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved));
            B2_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used));
            // 75 % of what the device and the pool's cache can still give (the rest is left to the transforms and multiexps
            // of the same proof that allocate after this point on other streams); if the allocation fails all the same --
            // a fragmented pool, another process on the device -- halve the window group and try again, and in the end run
            // the whole accumulation on the XYZZ kernel, which needs no scratch of its own
            const size_t avail = (size_t)((free_b + (reserved - used)) * 0.75);
            int wg = W;
            while (wg > 1 && group_bytes(wg) > avail) wg = (wg + 1) / 2;
            if (ctx->tree_wgroup > 0 && ctx->tree_wgroup < wg) wg = ctx->tree_wgroup;
            size_t cap[16];
            DevBuf<Affine<F>> ptsA, ptsB;
            DevBuf<F> pp0, top_scratch;
            DevBuf<uint32_t> offA, offB;
            std::vector<DevBuf<F>> val, ppd;
            int depth_top = 1;
            auto try_alloc = [&](int g) -> bool {
                const uint32_t nbg_max = (uint32_t)g * B;
                caps(n * (size_t)g, nbg_max, cap);
                depth_top = 1;                                   // product tree over the thread totals: depth d holds count_at(cap0, d) values
                while (tree_count_at((uint32_t)cap[0], depth_top, K, K2) > TREE_TOP) ++depth_top;
                try {
                    ptsA.alloc(cap[0], st); ptsB.alloc(tree_levels > 1 ? cap[1] : 0, st);
                    pp0.alloc(cap[0], st);
                    offA.alloc((size_t)nbg_max + 1, st); offB.alloc((size_t)nbg_max + 1, st);
                    val.clear(); ppd.clear();
                    val.resize(depth_top + 1); ppd.resize(depth_top + 1);
                    for (int d = 1; d <= depth_top; ++d) {
                        val[d].alloc(tree_count_at((uint32_t)cap[0], d, K, K2) + 1, st);
                        if (d < depth_top) ppd[d].alloc(tree_count_at((uint32_t)cap[0], d, K, K2) + 1, st);
                    }
                    top_scratch.alloc(TREE_TOP, st);
                    return true;
                } catch (const CudaError& e) {
                    if (e.e != cudaErrorMemoryAllocation) throw;
                    cudaGetLastError();
                    ptsA.free(); ptsB.free(); pp0.free(); offA.free(); offB.free(); val.clear(); ppd.clear(); top_scratch.free();
                    return false;
                }
            };
            bool have_scratch = try_alloc(wg);
            while (!have_scratch && wg > 1) {
                B2_CUDA(cudaMemPoolTrimTo(pool, 0));             // hand the pool's idle blocks back before the next attempt
                wg = (wg + 1) / 2;
                have_scratch = try_alloc(wg);
            }
            if (!have_scratch) {
                tree_levels = 0;                                 // no room for even one window: plain XYZZ accumulation below
                tree_fell_back = true;
            } else {
            for (int w0 = 0; w0 < W; w0 += wg) {
                const int w1 = w0 + wg < W ? w0 + wg : W;
                const uint32_t nbg = (uint32_t)(w1 - w0) * B, b0 = (uint32_t)w0 * B;
                caps(n * (size_t)(w1 - w0), nbg, cap);
                TreeIn<F> in{d_bases, sorted.get(), nullptr, off.get() + b0};
                tmark("begin");
                for (int l = 0; l < tree_levels; ++l) {
                    uint32_t* off_o = (l & 1) ? offB.get() : offA.get();
                    Affine<F>* pts_o = (l & 1) ? ptsB.get() : ptsA.get();
                    scan_u32(ctx, st, in.off, off_o, nbg, 2, 2, tile_sums.get());
                    const uint32_t* e_out = off_o + nbg;
                    const uint32_t t1 = tree_count_at((uint32_t)cap[l], 1, K, K2);
                    tmark("scan");
                    tree_up0_kernel<F><<<(t1 + 127) / 128, 128, 0, st>>>(in, off_o, nbg, K, pp0.get(), val[1].get());
                    tmark("up0");
                    int dt = 1;
                    while (tree_count_at((uint32_t)cap[l], dt, K, K2) > TREE_TOP) {
                        const uint32_t tn = tree_count_at((uint32_t)cap[l], dt + 1, K, K2);
                        tree_up_kernel<F><<<(tn + 127) / 128, 128, 0, st>>>(val[dt].get(), e_out, dt, K, K2, ppd[dt].get(), val[dt + 1].get());
                        ++dt;
                    }
                    tmark("up");
                    tree_top_kernel<F><<<1, 32, 0, st>>>(val[dt].get(), e_out, dt, K, K2, top_scratch.get());
                    tmark("top");
                    for (int d = dt - 1; d >= 1; --d) {
                        const uint32_t tn = tree_count_at((uint32_t)cap[l], d + 1, K, K2);
                        tree_down_kernel<F><<<(tn + 127) / 128, 128, 0, st>>>(val[d].get(), e_out, d, K, K2, ppd[d].get(), val[d + 1].get());
                    }
                    tmark("down");
                    tree_down0_kernel<F><<<(t1 + 127) / 128, 128, 0, st>>>(in, off_o, nbg, K, pp0.get(), val[1].get(), pts_o);
                    tmark("down0");
                    ctx->launches += 3 + 2 * (dt - 1);
                    in = TreeIn<F>{nullptr, nullptr, pts_o, off_o};
                }
                if (sizeof(F) > 48) msm_accumulate_chunked_kernel<F, 64, 4, true><<<148 * 4, 64, 0, st>>>(in.pts, nullptr, in.off, nbg, seg_acc.get() + b0, head, tail, hb, tb);
                else msm_accumulate_chunked_kernel<F, 128, 2, true><<<148 * 2, 128, 0, st>>>(in.pts, nullptr, in.off, nbg, seg_acc.get() + b0, head, tail, hb, tb);
                tmark("xyzz_tail");
                msm_fix_pieces_kernel<F><<<grid_for((size_t)n_chunks * 32, 128, 148 * 16), 128, 0, st>>>(seg_acc.get() + b0, head, tail, hb, tb, n_chunks);
                tmark("fix_pieces");
                ctx->launches += 2;
            }
            if (kl.e1) B2_CUDA(cudaEventRecord(kl.e1, st));
            }
        }
        if (tree_levels == 0) {
#define B2_CH(THREADS, MINB) msm_accumulate_chunked_kernel<F, THREADS, MINB><<<148 * MINB, THREADS, 0, st>>>(d_bases, sorted.get(), off.get(), (uint32_t)nb, seg_acc.get(), head, tail, hb, tb)
        if (sizeof(F) > 48) B2_CH(64, 4);
        else if (tree_fell_back) B2_CH(128, 2);              // the launch shape n_chunks was sized for
        else switch (ctx->acc_variant) {
            case 101: B2_CH(128, 1); break;
            case 103: B2_CH(128, 3); break;
            default: B2_CH(128, 2);                          // 102: two CTAs of 128 threads per SM (194 registers, no spills)
        }
#undef B2_CH
        if (kl.e1) B2_CUDA(cudaEventRecord(kl.e1, st));
        msm_fix_pieces_kernel<F><<<grid_for((size_t)n_chunks * 32, 128, 148 * 16), 128, 0, st>>>(seg_acc.get(), head, tail, hb, tb, n_chunks);
        ctx->launches += 2;
        }
    } else {
        if (kl.e0) B2_CUDA(cudaEventRecord(kl.e0, st));
        launch_accumulate<F>(ctx, st, max_segs, d_bases, sorted.get(), hist.get(), off.get(), seg_off.get(), (uint32_t)nb, S, seg_acc.get());
        if (kl.e1) B2_CUDA(cudaEventRecord(kl.e1, st));
        msm_fold_segments_kernel<F><<<grid_for(nb * 32, 128, 148 * 16), 128, 0, st>>>(seg_acc.get(), seg_off.get(), (uint32_t)nb); ctx->launches++;
    }
    if (kl.e0) {
        if (ctx->d_klog_entries && ctx->klog.size() < 64)
            B2_CUDA(cudaMemcpyAsync(ctx->d_klog_entries + ctx->klog.size(), off.get() + nb, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        ctx->klog.push_back(kl);
    }
    mark(4);
    const Xyzz<F>* win = nullptr;
    if (chunked) {
        // recursive chunked running sums over bucket_acc (= seg_acc), all windows per launch
        uint32_t count = B + 1, stride = B;
        int log_s = 0, side = 0, shift_one = 1;
        const Xyzz<F>* a_in = nullptr; const Xyzz<F>* r_in = seg_acc.get();
        for (;;) {
            const uint32_t T = (count + RED_PER - 1) / RED_PER;
            msm_reduce_level_kernel<F><<<(W * T + 127) / 128, 128, 0, st>>>(a_in, r_in, count, stride, shift_one, RED_PER, log_s, red_a[side].get(), red_r[side].get(), T, W);
            tmark("reduce_level");
            ctx->launches++;
            a_in = red_a[side].get(); r_in = red_r[side].get();
            if (T == 1) break;
            count = T; stride = T; log_s += red_log; shift_one = 0; side ^= 1;
            if (ctx->red_bits && count <= RED_BITS_MAX) {
                // few enough items left: finish by bit decomposition (two short launches instead of 3-5 latency-bound levels)
                int NB = 0;
                while ((1u << NB) < count) ++NB;                        // indices 0 .. count-1 fit NB bits
                Xyzz<F>* sums = red_a[side].get();                      // W * (NB + 1) <= W * T1 slots
                Xyzz<F>* out = red_r[side].get();
                msm_reduce_bits_kernel<F><<<dim3(NB + 1, W), 256, 0, st>>>(a_in, r_in, count, stride, NB, sums);
                tmark("reduce_bits");
                msm_reduce_bits_final_kernel<F><<<W, 32, 0, st>>>(sums, NB, log_s, out);
                tmark("reduce_bits_final");
                ctx->launches += 2;
                a_in = out;
                break;
            }
        }
        win = a_in;
    } else {
        const uint32_t T = msm_reduce_chunks(c);
        DevBuf<Xyzz<F>> partial((size_t)W * T, st);
        msm_reduce_kernel<F><<<(W * T + 127) / 128, 128, 0, st>>>(seg_acc.get(), seg_off.get(), c, W, T, partial.get()); ctx->launches++;
        msm_window_sum_kernel<F><<<W, 32, 0, st>>>(partial.get(), T, red_a[0].get()); ctx->launches++;
        win = red_a[0].get();
    }
    if (d_win_out) {
        B2_REQUIRE(W <= 32, B2G16_E_INTERNAL, "msm: window count");
        B2_CUDA(cudaMemcpyAsync(d_win_out, win, (size_t)W * sizeof(Xyzz<F>), cudaMemcpyDeviceToDevice, st));
        plan_out->c = c; plan_out->W = W;
    } else {
        msm_final_kernel<F><<<1, 32, 0, st>>>(win, c, W, d_out_bytes); ctx->launches++;
    }
    mark(5);
    B2_CUDA(cudaGetLastError());
    if (timed) {
        B2_CUDA(cudaEventSynchronize(ev[5]));
        for (int i = 0; i < 5; ++i) B2_CUDA(cudaEventElapsedTime(&ctx->msm_ms[i], ev[i], ev[i + 1]));
        for (auto& e : ev) cudaEventDestroy(e);
        for (size_t i = 1; i < tr.size(); ++i) {
            float ms = 0.f; cudaEventElapsedTime(&ms, tr[i - 1].second, tr[i].second);
            fprintf(stderr, "msm_trace g%d c=%d %s %.3f ms\n", sizeof(F) > 48 ? 2 : 1, c, tr[i].first, ms);
        }
        for (auto& t : tr) cudaEventDestroy(t.second);
    }
}

template <class F>
Affine<F>* fixed_base_table(Ctx* ctx, int slot, const Affine<F>& gen) {
    if (ctx->fixed_base[slot]) return (Affine<F>*)ctx->fixed_base[slot];
    Affine<F>* table = nullptr;
    B2_CUDA(cudaMalloc((void**)&table, sizeof(Affine<F>) * (32 * 255 + 32)));
    Affine<F>* roots = table + 32 * 255;
    fixed_base_roots_kernel<F><<<1, 32, 0, ctx->stream>>>(gen, roots);
    fixed_base_table_kernel<F><<<(32 * 255 + 63) / 64, 64, 0, ctx->stream>>>(roots, table);
    ctx->launches += 2;
    B2_CUDA(cudaGetLastError());
    ctx->fixed_base[slot] = table;
    return table;
}
#endif  // __CUDACC__

}  // namespace b2
