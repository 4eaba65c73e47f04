// Fr evaluation-domain transforms on the device: domain tables, the natural-order
// fft / inverse_fft / coset API and the r1cs_to_qap witness-map tail (H polynomial).
// Replaces crypto3-math `evaluation_domain<Fr>` and zk `r1cs_to_qap::witness_map`
// (un-vendored); called on behalf of compound_proof::circuit_proofs
// (libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137).
#include <cstdlib>
#include <cuda.h>
#include "common.cuh"
#include "ntt_plan.cuh"

namespace b2 {

struct NttDomain {
    int log_n = 0;
    NttPlan plan;
    Fr *tlo_f = nullptr, *thi_f = nullptr, *tlo_i = nullptr, *thi_i = nullptr;   // w^e, w^-e
    Fr *glo_f = nullptr, *ghi_f = nullptr, *glo_i = nullptr, *ghi_i = nullptr;   // g^e, n^-1 g^-e
    Fr *glo_f_raw = nullptr, *glo_i_raw = nullptr;                                 // glo_* as plain residues
    Fr n_inv, zinv;                 // n^-1 ;  (g^n - 1)^-1
    Fr* tw[NTT_MAX_PASSES] = {nullptr, nullptr, nullptr, nullptr};   // per strided pass: w_{2^L}^x, x < 2^(L-1) (ntt_stage_twiddle)
};

// out[e] = pre * base^(e << shift)
__global__ void pow_table_kernel(Fr* out, Fr base, Fr pre, int shift, uint32_t count) {
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < count; e += gridDim.x * blockDim.x) {
        uint64_t x = (uint64_t)e << shift;
        Fr r = pre, b = base;
        while (x) { if (x & 1) r = r * b; b = b.sqr(); x >>= 1; }
        fr_store(out + e, r);
    }
}

static Fr host_fr_u64(uint64_t x) {
    uint32_t c[8] = {(uint32_t)x, (uint32_t)(x >> 32), 0, 0, 0, 0, 0, 0};
    return Fr::from_canonical(c);
}
static Fr host_pow_u64(Fr b, uint64_t e) {
    Fr r = Fr::one();
    while (e) { if (e & 1) r = r * b; b = b.sqr(); e >>= 1; }
    return r;
}
// 7^((r-1) >> log_n): primitive 2^log_n-th root of unity
Fr host_root_of_unity(int log_n) {
    uint32_t e[8];
    for (int i = 0; i < 8; ++i) e[i] = FrParams::P(i);
    e[0] -= 1;                                         // r - 1 (r is odd, no borrow)
    for (int s = 0; s < log_n; ++s)
        for (int i = 0; i < 8; ++i) e[i] = (e[i] >> 1) | (i < 7 ? e[i + 1] << 31 : 0);
    return host_fr_u64(7).pow_limbs(e, 8);
}

static Fr* make_table(Ctx* c, Fr base, Fr pre, int shift, uint32_t count) {
    Fr* d = nullptr;
    B2_CUDA(cudaMalloc((void**)&d, (size_t)count * sizeof(Fr)));
    pow_table_kernel<<<grid_for(count, 256), 256, 0, c->stream>>>(d, base, pre, shift, count);
    c->launches++;
    B2_CUDA(cudaGetLastError());
    return d;
}

void ntt_init_roots(Ctx* c) {
    if (c->roots_fwd) return;
    Fr w = host_root_of_unity(NTT_ROOT_LOG);
    c->roots_fwd = make_table(c, w, Fr::one(), 0, 1u << (NTT_ROOT_LOG - 1));
    c->roots_inv = make_table(c, w.inv(), Fr::one(), 0, 1u << (NTT_ROOT_LOG - 1));
    const int smem = (int)(2 * sizeof(uint4) << NTT_TILE_LOG);
    B2_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<256, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B2_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<256, 3, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B2_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<128, 3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B2_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<256, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B2_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<256, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
}

NttDomain* ntt_domain(Ctx* c, int log_n) {
    auto it = c->domains.find(log_n);
    if (it != c->domains.end()) return it->second;
    B2_REQUIRE(log_n >= 0 && log_n <= 30, B2G16_E_INVALID_ARG, "domain size must be 2^0..2^30");
    ntt_init_roots(c);
    NttDomain* d = new NttDomain();
    d->log_n = log_n;
    d->plan = ntt_make_plan(log_n);
    const int lo = d->plan.lo_bits, hi = log_n - lo;
    const uint32_t n_lo = 1u << lo, n_hi = 1u << (hi > 0 ? hi : 0);
    Fr w = host_root_of_unity(log_n), wi = w.inv();
    Fr g = host_fr_u64(7), gi = g.inv();
    d->n_inv = host_fr_u64(1ull << log_n).inv();
    d->zinv = (host_pow_u64(g, 1ull << log_n) - Fr::one()).inv();
    d->tlo_f = make_table(c, w, Fr::one(), 0, n_lo);  d->thi_f = make_table(c, w, Fr::one(), lo, n_hi);
    d->tlo_i = make_table(c, wi, Fr::one(), 0, n_lo); d->thi_i = make_table(c, wi, Fr::one(), lo, n_hi);
    d->glo_f = make_table(c, g, Fr::one(), 0, n_lo);  d->ghi_f = make_table(c, g, Fr::one(), lo, n_hi);
    d->glo_i = make_table(c, gi, d->n_inv, 0, n_lo);  d->ghi_i = make_table(c, gi, Fr::one(), lo, n_hi);
    // plain-residue copies: pre * base^e with pre = (Montgomery form of 1/R) = the integer 1
    Fr raw_one = Fr::zero(); raw_one.v[0] = 1;
    d->glo_f_raw = make_table(c, g, raw_one, 0, n_lo);
    d->glo_i_raw = make_table(c, gi, d->n_inv * raw_one, 0, n_lo);
    // Butterfly twiddle tables of the strided passes: 2^(L-1) entries each, i.e. n/2 for the first pass (2 GB at 2^27)
    // and n/2^(k_0+1), .. for the later ones.  HBM streams them at a fraction of the time the multiplier would need to
    // rebuild the twiddles (the transform is bound by the integer pipe, not by memory); when the allocation does not fit
    // the pass keeps the two-level tlo * thi reconstruction.
    if (!getenv("B2G16_NTT_NO_TW_TABLES")) {
        for (int p = 0; p < d->plan.digits.n_pass; ++p) {
            if (d->plan.contiguous[p]) continue;
            const int L = d->plan.L[p];
            const uint32_t count = 1u << (L - 1);
            Fr* t = nullptr;
            if (cudaMalloc((void**)&t, (size_t)count * sizeof(Fr)) != cudaSuccess) { cudaGetLastError(); continue; }
            pow_table_kernel<<<grid_for(count, 256), 256, 0, c->stream>>>(t, host_root_of_unity(L), Fr::one(), 0, count);
            c->launches++;
            B2_CUDA(cudaGetLastError());
            d->tw[p] = t;
        }
    }
    c->domains[log_n] = d;
    return d;
}

void ntt_free_domains(Ctx* c) {
    for (auto& kv : c->domains) {
        NttDomain* d = kv.second;
        Fr* t[10] = {d->tlo_f, d->thi_f, d->tlo_i, d->thi_i, d->glo_f, d->ghi_f, d->glo_i, d->ghi_i, d->glo_f_raw, d->glo_i_raw};
        for (Fr* p : t) cudaFree(p);
        for (Fr* p : d->tw) cudaFree(p);
        delete d;
    }
    c->domains.clear();
    cudaFree(c->roots_fwd); cudaFree(c->roots_inv);
    c->roots_fwd = c->roots_inv = nullptr;
}

static NttTables tables(Ctx* c, const NttDomain* d, bool inverse) {
    NttTables t;
    t.roots = inverse ? c->roots_inv : c->roots_fwd;
    t.tlo = inverse ? d->tlo_i : d->tlo_f; t.thi = inverse ? d->thi_i : d->thi_f;
    t.glo = inverse ? d->glo_i : d->glo_f; t.ghi = inverse ? d->ghi_i : d->ghi_f;
    t.glo_raw = inverse ? d->glo_i_raw : d->glo_f_raw;
    for (int p = 0; p < NTT_MAX_PASSES; ++p) t.tw[p] = d->tw[p];
    t.inverse = inverse ? 1 : 0;
    return t;
}

// Tensor map of a vector of n Fr elements for the plane-wise tile load (ntt_kernels.cuh "TMA tile staging").  The encoder
// is a driver entry point, fetched through the runtime so that the library does not link libcuda.
static bool make_ntt_tmap(const Fr* src, uint64_t n, NttTmap* out) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { cudaGetLastError(); fn = nullptr; }
        return reinterpret_cast<EncodeFn>(fn);
    }();
    if (!encode) return false;
    static_assert(sizeof(CUtensorMap) == sizeof(NttTmap), "tensor map size");
    const cuuint64_t gdim[3] = {2, 2, n};                      // u64 words per half, halves per element, elements
    const cuuint64_t gstride[2] = {16, 32};                    // bytes between halves, between elements
    const cuuint32_t box[3] = {2, 1, 256}, estride[3] = {1, 1, 1};
    return encode(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<Fr*>(src), gdim, gstride, box, estride,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static void launch_pass(Ctx* c, const Fr* src, Fr* dst, const NttPass& p) {
    const uint64_t tiles = ntt_pass_tiles(p);
    const int T = 1 << (p.k + p.logc);
    const size_t smem = 2 * sizeof(uint4) * (size_t)T;
    const int max_grid = 148 * 3 * 4;                      // 3 resident CTAs per SM, a few waves
    int grid = (int)(tiles < (uint64_t)max_grid ? tiles : (uint64_t)max_grid);
    static const int variant = getenv("B2G16_NTT_VARIANT") ? atoi(getenv("B2G16_NTT_VARIANT")) : 0;
    static const int use_tma = getenv("B2G16_NTT_TMA") ? atoi(getenv("B2G16_NTT_TMA")) : 1;   // on by default; 0 = per-thread loads
    NttTmap tmap{};
    if (variant == 0 && use_tma && ntt_pass_tma_ok(p) && make_ntt_tmap(src, 1ull << p.log_n, &tmap)) {
        NttPass q = p;
        q.no_swz = 1;
        ntt_pass_kernel<256, 3, false, true><<<grid, 256, smem, c->stream>>>(src, dst, q, tmap);
    } else if (variant == 0) {
        const int block = T / 2 < 256 ? (T / 2 < 32 ? 32 : T / 2) : 256;
        ntt_pass_kernel<256, 3, false><<<grid, block, smem, c->stream>>>(src, dst, p, tmap);
    } else {
        const int cap = variant == 1 ? 128 : 256;
        const int block = T / 8 < cap ? (T / 8 < 32 ? 32 : T / 8) : cap;     // one register-round slot = 8 elements
        if (variant == 1) ntt_pass_kernel<128, 3, true><<<grid, block, smem, c->stream>>>(src, dst, p, tmap);
        else if (variant == 2) ntt_pass_kernel<256, 2, true><<<grid, block, smem, c->stream>>>(src, dst, p, tmap);
        else ntt_pass_kernel<256, 1, true><<<grid, block, smem, c->stream>>>(src, dst, p, tmap);
    }
    c->launches++;
    B2_CUDA(cudaGetLastError());
}

void ntt_natural_ex(Ctx* c, const Fr* src, Fr* dst, Fr* tmp, int log_n, bool inverse, bool coset,
                    bool in_canon, bool out_canon, int* d_bad) {
    NttDomain* d = ntt_domain(c, log_n);
    NttConsts k{d->n_inv, d->zinv};
    for (auto& s : ntt_steps_natural(d->plan, tables(c, d, inverse), k, src, dst, tmp, inverse, coset, in_canon, out_canon, d_bad))
        launch_pass(c, s.src, s.dst, s.pass);
}

bool ntt_needs_tmp(int log_n) { return ntt_make_plan(log_n).digits.n_pass > 1; }

void ntt_natural(Ctx* c, Fr* d_data, int log_n, bool inverse, bool coset) {
    DevBuf<Fr> tmp;
    if (ntt_needs_tmp(log_n)) tmp.alloc((size_t)1 << log_n, c->stream);
    ntt_natural_ex(c, d_data, d_data, tmp.get(), log_n, inverse, coset, false, false, nullptr);
}

void h_from_abc_dev(Ctx* c, Fr* d_a, Fr* d_b, Fr* d_c, int log_m, uint32_t* d_h_canonical, bool in_canon, int* d_bad) {
    NttDomain* d = ntt_domain(c, log_m);
    NttConsts k{d->n_inv, d->zinv};
    for (auto& s : ntt_steps_h(d->plan, tables(c, d, false), tables(c, d, true), k, d_a, d_b, d_c, reinterpret_cast<Fr*>(d_h_canonical), in_canon, d_bad))
        launch_pass(c, s.src, s.dst, s.pass);
}

// ------------------------------------------------------------------ distributed transform
// One transform of size N = 2^log_n over G = 2^log_g ranks, rank r holding x[r M .. (r+1) M),
// M = N / G.  With n = r M + j and k = k1 + G k2:
//     X[k1 + G k2] = sum_j w_M^(j k2) [ w_N^(j k1) sum_r w_G^(r k1) x[r M + j] ]
// so the only cross-rank arithmetic is a G-point transform per column j followed by the
// twiddle w_N^(j k1); everything else is the ordinary local transform of size M.  The host
// side (python/b2g16/dist.py) moves the data with all-to-all exchanges so that a rank holds,
// for its chunk of M/G columns, the G values of every column: a G x (M/G) row-major matrix.
// This kernel does the per-column work on that matrix, one thread per column, operands in
// registers.  Forward: G-point DIF, then twiddles.  Inverse: twiddles (inverse roots), then
// the inverse G-point transform, times `scale`.  Data stays in canonical form throughout
// (the transform is linear and every constant is in Montgomery form; ntt_plan.cuh).
struct CrossRoots { Fr w[4]; };       // w_G^i (forward or inverse), i < G/2, G <= 8
// Row r of the input / row k of the output, each `cols` elements.  Local form: rows of one G x cols
// matrix.  Peer form: in[r] points into rank r's block and out[k] into rank k's buffer, both mapped
// into this process with CUDA IPC, so the two exchanges around the cross step become NVLink loads and
// stores issued by the kernel itself (compute and transfer overlap thread by thread; no staging pass).
struct CrossPtrs { const Fr* in[8]; Fr* out[8]; };

template <int LOG_G> __global__ void __launch_bounds__(128)
ntt_cross_kernel(const CrossPtrs ptrs, uint64_t cols, uint64_t col0, const Fr* __restrict__ tlo,
                 const Fr* __restrict__ thi, int lo_bits, CrossRoots roots, Fr scale, int inverse, int use_scale, int* bad) {
    constexpr int G = 1 << LOG_G;
    for (uint64_t jj = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; jj < cols; jj += (uint64_t)gridDim.x * blockDim.x) {
        Fr v[G];
#pragma unroll
        for (int r = 0; r < G; ++r) {
            v[r] = fr_load(ptrs.in[r] + jj);
            if (bad && Fr::geq_p(v[r].v)) atomicOr(bad, 1);
        }
        const uint64_t j = col0 + jj;                               // global column: exponent of the twiddle base
        const uint32_t lo = (uint32_t)(j & ((1ull << lo_bits) - 1)), hi = (uint32_t)(j >> lo_bits);
        Fr t = fr_load(tlo + lo);
        if (hi) t = t * fr_load(thi + hi);                          // w_N^(+-j)
        if (inverse) {                                              // v[k1] *= w_N^(-j k1), with the overall scale folded in
            Fr p = scale;
            if (use_scale) v[0] = v[0] * p;
#pragma unroll
            for (int k = 1; k < G; ++k) { p = (k == 1 && !use_scale) ? t : p * t; v[k] = v[k] * p; }
        }
        // G-point radix-2 DIF in registers: natural in, bit-reversed out
#pragma unroll
        for (int s = LOG_G - 1; s >= 0; --s) {
            const int h = 1 << s;
#pragma unroll
            for (int b = 0; b < G / 2; ++b) {
                const int i0 = ((b >> s) << (s + 1)) | (b & (h - 1)), i1 = i0 + h;
                const int e = (b & (h - 1)) << (LOG_G - 1 - s);    // w_G^e
                Fr u = v[i0], w = v[i1];
                v[i0] = u + w;
                Fr d = u - w;
                v[i1] = e ? d * roots.w[e] : d;
            }
        }
        if (!inverse) {                                             // y[k1] * w_N^(j k1)
            Fr p = t;
#pragma unroll
            for (int k = 1; k < G; ++k) {
                int src = 0;
#pragma unroll
                for (int bit = 0; bit < LOG_G; ++bit) src |= ((k >> bit) & 1) << (LOG_G - 1 - bit);
                v[src] = v[src] * p;
                if (k + 1 < G) p = p * t;
            }
        }
#pragma unroll
        for (int k = 0; k < G; ++k) {
            int src = 0;
#pragma unroll
            for (int bit = 0; bit < LOG_G; ++bit) src |= ((k >> bit) & 1) << (LOG_G - 1 - bit);
            fr_store(ptrs.out[k] + jj, v[src]);
        }
    }
}

// data[i] *= table power of g at global exponent offset + i (coset shift of a rank's block)
__global__ void __launch_bounds__(256) fr_coset_scale_kernel(Fr* data, uint64_t count, uint64_t offset, const Fr* __restrict__ glo,
                                                             const Fr* __restrict__ ghi, int lo_bits, int* bad) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t e = offset + i;
        const uint32_t lo = (uint32_t)(e & ((1ull << lo_bits) - 1)), hi = (uint32_t)(e >> lo_bits);
        Fr t = fr_load(glo + lo);
        if (hi) t = t * fr_load(ghi + hi);
        Fr x = fr_load(data + i);
        if (bad && Fr::geq_p(x.v)) atomicOr(bad, 1);
        fr_store(data + i, x * t);
    }
}
// out[c * rows + r] = in[r * cols + c]   (32-byte elements)
__global__ void __launch_bounds__(256) fr_transpose_kernel(const Fr* __restrict__ in, Fr* __restrict__ out, uint64_t rows, uint64_t cols) {
    const uint64_t n = rows * cols;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / cols, c = i % cols;
        fr_store(out + c * rows + r, fr_load(in + i));
    }
}

static void ntt_cross_launch(Ctx* c, const CrossPtrs& ptrs, int log_n, int log_g, uint64_t chunk, bool inverse, bool coset, int* d_bad);

void ntt_cross_dev(Ctx* c, const Fr* d_in, Fr* d_out, int log_n, int log_g, uint64_t chunk, bool inverse, bool coset, int* d_bad) {
    B2_REQUIRE(log_g >= 1 && log_g <= 3 && log_n >= 2 * log_g && log_n <= 30, B2G16_E_INVALID_ARG, "distributed ntt: 2, 4 or 8 ranks, 2^log_n >= ranks^2");
    const uint64_t cols = (1ull << log_n) >> (2 * log_g);
    CrossPtrs ptrs{};
    for (int r = 0; r < (1 << log_g); ++r) { ptrs.in[r] = d_in + (uint64_t)r * cols; ptrs.out[r] = d_out + (uint64_t)r * cols; }
    ntt_cross_launch(c, ptrs, log_n, log_g, chunk, inverse, coset, d_bad);
}
// peers_in[r] / peers_out[k]: base of rank r's (k's) M-element buffer as mapped in this process
void ntt_cross_p2p_dev(Ctx* c, const Fr* const* peers_in, Fr* const* peers_out, int log_n, int log_g, uint64_t rank, bool inverse, bool coset, int* d_bad) {
    B2_REQUIRE(log_g >= 1 && log_g <= 3 && log_n >= 2 * log_g && log_n <= 30, B2G16_E_INVALID_ARG, "distributed ntt: 2, 4 or 8 ranks, 2^log_n >= ranks^2");
    const uint64_t cols = (1ull << log_n) >> (2 * log_g);
    CrossPtrs ptrs{};
    for (int r = 0; r < (1 << log_g); ++r) {
        B2_REQUIRE(peers_in[r] && peers_out[r], B2G16_E_INVALID_ARG, "distributed ntt: missing peer buffer");
        ptrs.in[r] = peers_in[r] + rank * cols; ptrs.out[r] = peers_out[r] + rank * cols;
    }
    ntt_cross_launch(c, ptrs, log_n, log_g, rank, inverse, coset, d_bad);
}
static void ntt_cross_launch(Ctx* c, const CrossPtrs& ptrs, int log_n, int log_g, uint64_t chunk, bool inverse, bool coset, int* d_bad) {
    B2_REQUIRE(log_g >= 1 && log_g <= 3, B2G16_E_INVALID_ARG, "distributed ntt: 2, 4 or 8 ranks");
    B2_REQUIRE(log_n >= 2 * log_g && log_n <= 30, B2G16_E_INVALID_ARG, "distributed ntt: need 2^log_n >= ranks^2");
    B2_REQUIRE(chunk < (1ull << log_g), B2G16_E_INVALID_ARG, "distributed ntt: chunk index");
    NttDomain* d = ntt_domain(c, log_n);
    const int G = 1 << log_g;
    const uint64_t cols = (1ull << log_n) >> (2 * log_g);
    Fr wg = host_root_of_unity(log_g);
    if (inverse) wg = wg.inv();
    CrossRoots roots;
    roots.w[0] = Fr::one();
    for (int i = 1; i < 4; ++i) roots.w[i] = roots.w[i - 1] * wg;
    // inverse: the local inverse transform already scaled by 1/M; plain needs 1/G more, the coset
    // variant's g^-e table carries 1/N, so it needs M to end at 1/N overall
    Fr scale = inverse ? (coset ? host_fr_u64(1ull << (log_n - log_g)) : host_fr_u64((uint64_t)G).inv()) : Fr::one();
    const Fr* tlo = inverse ? d->tlo_i : d->tlo_f; const Fr* thi = inverse ? d->thi_i : d->thi_f;
    const int grid = grid_for(cols, 128, 148 * 8);
#define B2_CROSS(LG) ntt_cross_kernel<LG><<<grid, 128, 0, c->stream>>>(ptrs, cols, chunk * cols, tlo, thi, d->plan.lo_bits, roots, scale, inverse ? 1 : 0, inverse ? 1 : 0, d_bad)
    if (log_g == 1) B2_CROSS(1); else if (log_g == 2) B2_CROSS(2); else B2_CROSS(3);
#undef B2_CROSS
    c->launches++;
    B2_CUDA(cudaGetLastError());
}
void fr_coset_scale_dev(Ctx* c, Fr* d_data, uint64_t count, int log_n, uint64_t offset, bool inverse, int* d_bad) {
    B2_REQUIRE(offset + count <= (1ull << log_n), B2G16_E_INVALID_ARG, "coset scale: range outside the domain");
    NttDomain* d = ntt_domain(c, log_n);
    fr_coset_scale_kernel<<<grid_for(count, 256), 256, 0, c->stream>>>(d_data, count, offset, inverse ? d->glo_i : d->glo_f, inverse ? d->ghi_i : d->ghi_f, d->plan.lo_bits, d_bad);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}
void fr_transpose_dev(Ctx* c, const Fr* d_in, Fr* d_out, uint64_t rows, uint64_t cols) {
    fr_transpose_kernel<<<grid_for(rows * cols, 256), 256, 0, c->stream>>>(d_in, d_out, rows, cols);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}

}  // namespace b2
