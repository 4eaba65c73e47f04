// Fr radix-2 NTT as a sequence of shared-memory passes (multi-step / "four-step"
// decomposition).  Replaces crypto3-math `evaluation_domain<Fr>::fft / inverse_fft` and
// the coset multiplies used by the r1cs_to_qap witness map (un-vendored).
//
// A transform of size N = 2^log_n is split into passes p = 0..P-1 with k_p index bits
// each (sum k_p = log_n).  Pass p sees the vector as contiguous blocks of size 2^L_p
// (L_0 = log_n, L_{p+1} = L_p - k_p); inside a block it runs 2^(L_p-k_p) independent
// size-2^k_p NTTs along stride S_p = 2^(L_p-k_p) with *pure* 2^k_p-th roots, and
// multiplies element (k1, i2) by the inter-pass twiddle w_N^((i2*k1) << (log_n-L_p)).
//   NR order ("natural -> digit-reversed"): passes 0..P-1, twiddle AFTER the small NTT.
//   RN order ("digit-reversed -> natural"): passes P-1..0, twiddle BEFORE the small NTT
//   (the exact transpose of NR).  Memory position with pass digits (d_0,..,d_{P-1})
//   (top to bottom) holds spectrum/coefficient index d_0 + 2^k_0 d_1 + 2^(k_0+k_1) d_2 ...
// The witness map chains NR(inverse) -> RN(forward, coset) -> pointwise -> NR(inverse)
// so no reordering pass is ever needed; only the public natural-order API permutes.
//
// One CTA owns a tile of 2^(k+logc) <= 2048 elements (64 KB of shared memory as two
// uint4 planes -> conflict-free 128-bit accesses): 2^k rows x 2^logc columns.  Strided
// passes take 2^logc *adjacent* columns so every global row access is 2^logc*32 B
// contiguous; the last pass is fully contiguous.  HBM traffic: one read + one write of
// the vector per pass (2*N*32 B), everything else stays on chip.
//
// The kernel body is written as barrier-free *phases* (load / stage / store) so that
// tests/host_emu can execute the identical index arithmetic on the CPU.
#pragma once
#include "fp.cuh"

namespace b2 {

constexpr int NTT_TILE_LOG = 11;           // 2048 elements = 64 KB
constexpr int NTT_ROOT_LOG = 11;           // in-tile root table: w_{2^11}^j, j < 1024
constexpr int NTT_MAX_PASSES = 4;

struct NttDigits {                         // pass split, top to bottom
    int n_pass;
    int k[NTT_MAX_PASSES];
    int log_n;
};

// spectrum/coefficient index stored at memory position `pos` in digit-reversed layouts
B2_HD uint64_t ntt_digit_index(const NttDigits& d, uint64_t pos) {
    int L = d.log_n, shift = 0;
    uint64_t idx = 0;
    for (int p = 0; p < d.n_pass; ++p) {
        L -= d.k[p];
        idx |= ((pos >> L) & ((1ull << d.k[p]) - 1)) << shift;
        shift += d.k[p];
    }
    return idx;
}

B2_HD int ntt_swz_raw(int e, bool on) { return on ? (e ^ ((e >> 3) & 7)) : e; }

struct NttPass {
    int log_n, L, k, logc;      // block log-size, row bits, column bits of the tile
    int contiguous;             // 1: last pass (rows contiguous, columns = adjacent blocks)
    int tw_mode;                // 0 none, 1 after small NTT (NR), 2 before (RN): inter-pass twiddle rebuilt as tlo * thi (fallback);
                                // 3: no inter-pass twiddle -- every butterfly of the pass takes its *combined* twiddle
                                //    w_{2^L}^((j S + i2) << (k-1-hs)) from the table `tw` (see ntt_stage_twiddle)
    int dit;                    // 1: decimation-in-time stages (rows bit-reversed on load, half-size ascending, twiddle before
                                //    the butterfly): the transpose of the DIF pass, used by strided passes of RN transforms
    const Fr* tw;               // tw_mode 3: w_{2^L}^x, x < 2^(L-1), forward roots; the inverse transform reads -tw[2^(L-1) - x]
    uint32_t tw_half;           // 2^(L-1)
    int tw_inv;                 // 1: inverse roots wanted
    int scale_mode;             // 0 none, 1 on load, 2 on store
    int scale_kind;             // 0: constant cscale; 1: g-power table at natural position;
                                // 2: g-power table at digit-reversed index of the position
    const Fr* roots;            // w_{2^11}^j  (forward or inverse set)
    const Fr* tlo; const Fr* thi; int lo_bits;      // w_N^e = tlo[e & mask] * thi[e >> lo_bits]
    const Fr* glo; const Fr* ghi; int g_lo_bits;    // scale(e) = glo[e & mask] * ghi[e >> g_lo_bits]
    Fr cscale;
    NttDigits digits;
    // fusions at the two ends of a transform chain
    int in_canon;               // input holds canonical 32-byte scalars: convert (x * R^2) on load
    int check_canon;            // input holds canonical scalars that are used as they are (see ntt_plan.cuh): range check only
    int out_canon;              // write canonical scalars on store
    int perm_store;             // store position pos -> ntt_digit_index(digits, pos) (digit-reversed -> natural)
    int fuse_abc;               // load computes (x * b[pos] - c[pos]) * zinv   (witness map)
    const Fr* src_b; const Fr* src_c; Fr zinv;
    int* bad;                   // device flag, set when an in_canon input is >= r (device build only)
    int no_swz;                 // 1: tile slots are not swizzled (tiles loaded by TMA)
};
B2_HD int ntt_swz(int e, const NttPass& p) { return ntt_swz_raw(e, !p.no_swz && p.k + p.logc >= 6); }

struct Fr2x { uint4 lo, hi; };
B2_HD Fr fr_from_planes(const uint4& lo, const uint4& hi) {
    Fr r; r.v[0] = lo.x; r.v[1] = lo.y; r.v[2] = lo.z; r.v[3] = lo.w; r.v[4] = hi.x; r.v[5] = hi.y; r.v[6] = hi.z; r.v[7] = hi.w; return r;
}
B2_HD void fr_to_planes(const Fr& f, uint4& lo, uint4& hi) {
    lo.x = f.v[0]; lo.y = f.v[1]; lo.z = f.v[2]; lo.w = f.v[3]; hi.x = f.v[4]; hi.y = f.v[5]; hi.z = f.v[6]; hi.w = f.v[7];
}
B2_HD Fr fr_load(const Fr* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 lo = q[0], hi = q[1];
    return fr_from_planes(lo, hi);
}
B2_HD void fr_store(Fr* p, const Fr& f) {
    uint4 lo, hi; fr_to_planes(f, lo, hi);
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = lo; q[1] = hi;
}

B2_HD uint32_t bitrev_u32(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; ++i) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

// Shared-memory slot of tile element e.  The register rounds below read a tile with lane strides of 8 or 32 elements
// (the rows a thread owns are contiguous in the last round), which would put a quarter-warp's 16-byte accesses on the same
// banks; XOR-ing the low three index bits with the next three spreads those strides over all 32 banks and only permutes
// elements inside 128-byte blocks, so unit-stride accesses stay conflict-free.  Tiles smaller than 64 elements are left alone, and so are tiles that a
// tensor map fills (NttPass::no_swz: the TMA swizzle modes want a box row as wide as their span, the plane rows are 16 B).

// ---- tile geometry -------------------------------------------------------------
// tile element e in [0, 2^(k+logc)):  (row, col) and its global position.
struct NttTile {
    int k, logc, contiguous;
    uint64_t base;          // global position of (row 0, col 0)
    uint64_t row_stride;    // global stride between rows
    uint64_t col0;          // first column index inside the block (i2 of col 0), strided passes
    B2_HD int row_of(int e) const { return contiguous ? (e & ((1 << k) - 1)) : (e >> logc); }
    B2_HD int col_of(int e) const { return contiguous ? (e >> k) : (e & ((1 << logc) - 1)); }
    B2_HD int elem(int row, int col) const { return contiguous ? ((col << k) | row) : ((row << logc) | col); }
    B2_HD uint64_t gpos(int row, int col) const {
        return contiguous ? base + ((uint64_t)col << k) + row : base + (uint64_t)row * row_stride + col;
    }
};

B2_HD NttTile ntt_make_tile(const NttPass& p, uint64_t tile_id) {
    NttTile t; t.k = p.k; t.logc = p.logc; t.contiguous = p.contiguous;
    if (p.contiguous) {
        t.base = tile_id << (p.k + p.logc); t.row_stride = 1; t.col0 = 0;
    } else {
        int s_log = p.L - p.k;                         // columns per block = 2^s_log
        uint64_t tiles_per_block = 1ull << (s_log - p.logc);
        uint64_t blk = tile_id / tiles_per_block, cg = tile_id % tiles_per_block;
        t.col0 = cg << p.logc;
        t.row_stride = 1ull << s_log;
        t.base = (blk << p.L) + t.col0;
    }
    return t;
}

B2_HD Fr ntt_twiddle(const NttPass& p, uint64_t i2, uint32_t k1) {
    uint64_t e = (i2 * k1) << (p.log_n - p.L);
    uint32_t lo = (uint32_t)(e & ((1ull << p.lo_bits) - 1)), hi = (uint32_t)(e >> p.lo_bits);
    Fr t = fr_load(p.tlo + lo);
    if (hi) t = t * fr_load(p.thi + hi);
    return t;
}
B2_HD Fr ntt_scale(const NttPass& p, uint64_t pos) {
    if (p.scale_kind == 0) return p.cscale;
    uint64_t e = (p.scale_kind == 1) ? pos : ntt_digit_index(p.digits, pos);
    uint32_t lo = (uint32_t)(e & ((1ull << p.g_lo_bits) - 1)), hi = (uint32_t)(e >> p.g_lo_bits);
    Fr t = fr_load(p.glo + lo);
    if (hi) t = t * fr_load(p.ghi + hi);
    return t;
}

// ---- phases (each is data-parallel over `e`; barriers only between phases) ------
B2_HD void ntt_phase_load(const NttPass& p, const NttTile& t, const Fr* data, uint4* s_lo, uint4* s_hi, int e) {
    int row = t.row_of(e), col = t.col_of(e);
    uint64_t g = t.gpos(row, col);
    Fr x = fr_load(data + g);
    if (p.in_canon || p.check_canon) {
#if defined(__CUDA_ARCH__)
        if (Fr::geq_p(x.v)) atomicOr(p.bad, 1);
#endif
        if (p.in_canon) x = x * Fr::r2();
    }
    if (p.fuse_abc) x = (x * fr_load(p.src_b + g) - fr_load(p.src_c + g)) * p.zinv;
    if (p.scale_mode == 1) x = x * ntt_scale(p, g);
    if (p.tw_mode == 2 && row != 0) x = x * ntt_twiddle(p, t.col0 + col, (uint32_t)row);
    const int dst = ntt_swz(p.dit ? t.elem((int)bitrev_u32((uint32_t)row, p.k), col) : e, p);
    fr_to_planes(x, s_lo[dst], s_hi[dst]);
}
// Twiddle of the butterfly (row j of the half-block, column col) in the stage with half-size h = 2^hs.
// Contiguous pass / fallback modes: the pure root w_{2^(hs+1)}^j from the 2^11-th root table.
// tw_mode 3 (strided pass of a block of 2^L positions, S = 2^(L-k) columns): the pass is the first k stages of the
// radix-2 transform of the whole block, whose butterfly (q, q + h S), q = j S + i2, carries w_{2 h S}^q =
// w_{2^L}^(q << (k-1-hs)); using it directly leaves nothing for an inter-pass twiddle -- one product per butterfly
// and none per element, where the fallback pays 1-2 more products per element and pass.  `neg` is set when the
// value returned is the negated twiddle (inverse roots: w^-x = -w^(2^(L-1) - x)); the caller folds the sign into
// its add / sub.  Returns false when the twiddle is 1.
B2_HD bool ntt_stage_twiddle(const NttPass& p, const NttTile& t, int hs, int j, int col, Fr& w, bool& neg) {
    neg = false;
    if (p.tw_mode == 3) {
        const uint64_t x = ((((uint64_t)j) << (p.L - p.k)) + t.col0 + (uint64_t)col) << (p.k - 1 - hs);
        if (x == 0) return false;
        if (p.tw_inv) { w = fr_load(p.tw + (p.tw_half - x)); neg = true; }
        else w = fr_load(p.tw + x);
        return true;
    }
    if (j == 0) return false;
    w = fr_load(p.roots + ((uint32_t)j << (NTT_ROOT_LOG - 1 - hs)));
    return true;
}
// butterfly `b` of the stage with half-size h = 2^hs: DIF (u + v, (u - v) w) or, for p.dit, DIT (u + v w, u - v w)
B2_HD void ntt_phase_stage(const NttPass& p, const NttTile& t, uint4* s_lo, uint4* s_hi, int hs, int b) {
    int half_rows = 1 << (p.k - 1);
    int rb, col;
    if (t.contiguous) { rb = b & (half_rows - 1); col = b >> (p.k - 1); }
    else { col = b & ((1 << p.logc) - 1); rb = b >> p.logc; }
    int h = 1 << hs;
    int j = rb & (h - 1), grp = rb >> hs;
    int r0 = (grp << (hs + 1)) | j;
    int e0 = ntt_swz(t.elem(r0, col), p), e1 = ntt_swz(t.elem(r0 + h, col), p);
    Fr u = fr_from_planes(s_lo[e0], s_hi[e0]);
    Fr v = fr_from_planes(s_lo[e1], s_hi[e1]);
    Fr w; bool neg;
    const bool has_w = ntt_stage_twiddle(p, t, hs, j, col, w, neg);
    Fr s, d;
    if (p.dit) {
        if (has_w) v = v * w;
        if (neg) { s = u - v; d = u + v; } else { s = u + v; d = u - v; }
    } else {
        s = u + v;
        d = neg ? v - u : u - v;
        if (has_w) d = d * w;
    }
    fr_to_planes(s, s_lo[e0], s_hi[e0]);
    fr_to_planes(d, s_lo[e1], s_hi[e1]);
}
// One butterfly on two values held in registers (same arithmetic as ntt_phase_stage), twiddle already fetched
B2_HD void ntt_butterfly(bool dit, bool has_w, bool neg, const Fr& w, Fr& u, Fr& v) {
    if (dit) {
        if (has_w) v = v * w;
        const Fr s = neg ? u - v : u + v, d = neg ? u + v : u - v;
        u = s; v = d;
    } else {
        const Fr s = u + v;
        Fr d = neg ? v - u : u - v;
        if (has_w) d = d * w;
        u = s; v = d;
    }
}
// R consecutive stages (half-sizes 2^hmin .. 2^(hmin+R-1)) of the tile's row transform on the 2^R elements one thread
// keeps in registers: rows base | (m << hmin), m < 2^R, one column.  A tile has 2^(k+logc-R) such slots; consecutive
// slots are adjacent in shared memory.  One shared-memory round trip and one barrier per R stages instead of per stage,
// and the index arithmetic is paid once per 2^R elements.
template <int R, bool DIT> B2_HD void ntt_phase_round(const NttPass& p, const NttTile& t, uint4* s_lo, uint4* s_hi, int hmin, int slot) {
    constexpr int M = 1 << R;
    const int hmax = hmin + R - 1;
    int col, low, high;
    if (t.contiguous) {
        low = slot & ((1 << hmin) - 1);
        const int q = slot >> hmin, nh = p.k - 1 - hmax;
        high = q & ((1 << nh) - 1); col = q >> nh;
    } else {
        col = slot & ((1 << p.logc) - 1);
        const int q = slot >> p.logc;
        low = q & ((1 << hmin) - 1); high = q >> hmin;
    }
    const int base = (high << (hmax + 1)) | low;
    Fr v[M];
#pragma unroll
    for (int m = 0; m < M; ++m) { const int e = ntt_swz(t.elem(base | (m << hmin), col), p); v[m] = fr_from_planes(s_lo[e], s_hi[e]); }
    // all 2^R - 1 twiddles of the round first (the stage with half-size bit b has 2^b distinct ones: the row's bits below
    // b), so that their global-memory latencies overlap instead of queueing in front of each stage
    Fr w[M - 1 > 0 ? M - 1 : 1];
    uint32_t has_mask = 0, neg_mask = 0;
#pragma unroll
    for (int b = 0; b < R; ++b) {
#pragma unroll
        for (int ml = 0; ml < (1 << b); ++ml) {
            const int idx = (1 << b) - 1 + ml, hs = hmin + b;
            bool neg;
            const bool has_w = ntt_stage_twiddle(p, t, hs, (base | (ml << hmin)) & ((1 << hs) - 1), col, w[idx], neg);
            has_mask |= (has_w ? 1u : 0u) << idx; neg_mask |= (neg ? 1u : 0u) << idx;
        }
    }
#pragma unroll
    for (int s = 0; s < R; ++s) {
        const int b = DIT ? s : R - 1 - s;            // DIF: largest half-size first; DIT: smallest first (compile-time: v[] stays in registers)
#pragma unroll
        for (int m = 0; m < M; ++m) {
            if ((m >> b) & 1) continue;
            const int idx = (1 << b) - 1 + (m & ((1 << b) - 1));
            ntt_butterfly(DIT, (has_mask >> idx) & 1u, (neg_mask >> idx) & 1u, w[idx], v[m], v[m | (1 << b)]);
        }
    }
#pragma unroll
    for (int m = 0; m < M; ++m) { const int e = ntt_swz(t.elem(base | (m << hmin), col), p); fr_to_planes(v[m], s_lo[e], s_hi[e]); }
}
// Rounds of a pass: the k stages in execution order, three at a time; the last round takes what is left.
// Round r covers execution-order stages [3r, min(3r+3, k)): DIF -> half-size bits hmin = k - 3r - R, DIT -> hmin = 3r.
B2_HD int ntt_round_count(const NttPass& p) { return (p.k + 2) / 3; }
B2_HD int ntt_round_size(const NttPass& p, int r) { const int left = p.k - 3 * r; return left < 3 ? left : 3; }
B2_HD int ntt_round_hmin(const NttPass& p, int r) { return p.dit ? 3 * r : p.k - 3 * r - ntt_round_size(p, r); }
B2_HD void ntt_phase_round_any(const NttPass& p, const NttTile& t, uint4* s_lo, uint4* s_hi, int r, int slot) {
    const int R = ntt_round_size(p, r), hmin = ntt_round_hmin(p, r);
    if (p.dit) {
        if (R == 3) ntt_phase_round<3, true>(p, t, s_lo, s_hi, hmin, slot);
        else if (R == 2) ntt_phase_round<2, true>(p, t, s_lo, s_hi, hmin, slot);
        else ntt_phase_round<1, true>(p, t, s_lo, s_hi, hmin, slot);
    } else {
        if (R == 3) ntt_phase_round<3, false>(p, t, s_lo, s_hi, hmin, slot);
        else if (R == 2) ntt_phase_round<2, false>(p, t, s_lo, s_hi, hmin, slot);
        else ntt_phase_round<1, false>(p, t, s_lo, s_hi, hmin, slot);
    }
}
// stage order: DIF runs half-sizes k-1 .. 0, DIT 0 .. k-1
B2_HD int ntt_stage_hs(const NttPass& p, int i) { return p.dit ? i : p.k - 1 - i; }
B2_HD void ntt_phase_store(const NttPass& p, const NttTile& t, Fr* data, const uint4* s_lo, const uint4* s_hi, int e) {
    int row = t.row_of(e), col = t.col_of(e);          // output row k1 (natural order)
    int src = ntt_swz(p.dit ? e : t.elem((int)bitrev_u32((uint32_t)row, p.k), col), p);
    Fr x = fr_from_planes(s_lo[src], s_hi[src]);
    uint64_t g = t.gpos(row, col);
    if (p.tw_mode == 1 && row != 0) x = x * ntt_twiddle(p, t.col0 + col, (uint32_t)row);
    if (p.scale_mode == 2) x = x * ntt_scale(p, g);
    if (p.out_canon) { Fr c; x.to_canonical(c.v); x = c; }
    fr_store(data + (p.perm_store ? ntt_digit_index(p.digits, g) : g), x);
}

#if defined(__CUDACC__)
// ---- TMA tile staging (cp.async.bulk.tensor + mbarrier) for plain loads of contiguous tiles ------------------
// A contiguous pass reads its 2048-element tile from 64 KB of consecutive global memory.  Seen as a rank-3 tensor of
// u64 -- (2 words, 2 halves, N elements), strides 8 / 16 / 32 B -- a box of (2, 1, 256) is 256 element *halves*, 16 B
// each, packed: exactly one eighth of a plane of the tile's shared-memory layout (the tensor map's swizzle modes fault
// with 16-byte box rows -- tools/tma_probe.cu -- so a TMA-filled tile keeps plain slots, NttPass::no_swz).  Sixteen boxes fill both planes; one elected thread
// issues them against an mbarrier and every thread waits on its phase -- no address arithmetic, no LDG / STS, no
// registers for data in flight.  Only loads with nothing fused into them qualify (ntt_pass_tma_ok).
B2_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
B2_D void mbar_init(uint64_t* b, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(count), "r"(smem_u32(b)) : "memory"); }
B2_D void mbar_expect_tx(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(bytes), "r"(smem_u32(b)) : "memory"); }
B2_D bool mbar_try_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
B2_D void tma_load_3d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
struct alignas(64) NttTmap { unsigned char bytes[128]; };      // a CUtensorMap, opaque here (cuda.h is included by ntt.cu only)
#endif
inline bool ntt_pass_tma_ok(const NttPass& p) {
    return p.contiguous && !p.dit && !p.in_canon && !p.check_canon && !p.fuse_abc && p.scale_mode != 1 && p.tw_mode != 2 && p.k + p.logc == NTT_TILE_LOG;
}
#if defined(__CUDACC__)
// THREADS x MINB: CTA size and CTAs per SM the kernel is compiled for; ROUNDS: three stages at a time on eight values in
// registers (ntt_phase_round) instead of one stage per barrier (ntt_phase_stage).  A 2048-element tile is 64 KB of shared
// memory, so at most three CTAs fit an SM.  B2G16_NTT_VARIANT: 0 = one stage per barrier, 256 x 3 (24 warps per SM);
// 1 / 2 / 3 = register rounds at 128 x 3 (168 registers), 256 x 2 (128 registers, spills), 256 x 1 (no spills, 8 warps).
template <int THREADS, int MINB, bool ROUNDS, bool TMA = false> static __global__ void __launch_bounds__(THREADS, MINB)
ntt_pass_kernel(const Fr* src, Fr* data, const NttPass p, const __grid_constant__ NttTmap tmap) {
    extern __shared__ __align__(1024) uint4 smem[];
    __shared__ __align__(8) uint64_t mbar;
    const int T = 1 << (p.k + p.logc);
    uint4* s_lo = smem;
    uint4* s_hi = smem + T;
    uint32_t phase = 0;
    if (TMA) {
        if (threadIdx.x == 0) { mbar_init(&mbar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncthreads();
    }
    for (uint64_t tile_id = blockIdx.x; tile_id < ((1ull << p.log_n) >> (p.k + p.logc)); tile_id += gridDim.x) {
        const NttTile t = ntt_make_tile(p, tile_id);
        if (TMA) {
            if (threadIdx.x == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the previous tile's generic-proxy accesses come first
                mbar_expect_tx(&mbar, (uint32_t)(2 * sizeof(uint4)) << NTT_TILE_LOG);
                const int e0 = (int)(tile_id << NTT_TILE_LOG);
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int q = 0; q < (1 << NTT_TILE_LOG) / 256; ++q) tma_load_3d((h ? s_hi : s_lo) + q * 256, &tmap, &mbar, 0, h, e0 + q * 256);
            }
            uint32_t spins = 0;
            while (!mbar_try_wait(&mbar, phase)) { if (++spins > (1u << 26)) __trap(); }   // a lost transaction must not hang the device
            phase ^= 1u;
        } else {
            for (int e = threadIdx.x; e < T; e += blockDim.x) ntt_phase_load(p, t, src, s_lo, s_hi, e);
            __syncthreads();
        }
        if (ROUNDS) {
            const int n_rounds = ntt_round_count(p);
            for (int r = 0; r < n_rounds; ++r) {
                const int slots = T >> ntt_round_size(p, r);
                for (int s = threadIdx.x; s < slots; s += blockDim.x) ntt_phase_round_any(p, t, s_lo, s_hi, r, s);
                __syncthreads();
            }
        } else {
            for (int i = 0; i < p.k; ++i) {
                const int hs = ntt_stage_hs(p, i);
                for (int b = threadIdx.x; b < T / 2; b += blockDim.x) ntt_phase_stage(p, t, s_lo, s_hi, hs, b);
                __syncthreads();
            }
        }
        for (int e = threadIdx.x; e < T; e += blockDim.x) ntt_phase_store(p, t, data, s_lo, s_hi, e);
        __syncthreads();
    }
}

// out[digit_index(pos)] = in[pos]  (digit-reversed -> natural), or the inverse gather
static __global__ void ntt_permute_kernel(const Fr* in, Fr* out, NttDigits d, int to_natural) {
    uint64_t n = 1ull << d.log_n;
    for (uint64_t pos = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; pos < n; pos += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t idx = ntt_digit_index(d, pos);
        if (to_natural) fr_store(out + idx, fr_load(in + pos));
        else fr_store(out + pos, fr_load(in + idx));
    }
}
#endif

}  // namespace b2
