// Shared host-side plumbing for libb2g16: context, error reporting, device buffers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <map>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b2g16.h"
#include "ec.cuh"

namespace b2 {

struct NttDomain;

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;      // main stream (caller-visible ordering)
    bool own_stream = false;
    cudaStream_t aux[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[4] = {nullptr, nullptr, nullptr, nullptr};
    std::string last_error;
    // Queue of the context: one call in flight at a time; a waiting priority call (the `priority` flag of
    // compound_proof::circuit_proofs, compound_proof.hpp:134 -- PoSt proofs jump ahead of sealing proofs) is admitted
    // before any waiting ordinary call (ContextGate below).
    std::mutex mu;
    std::condition_variable cv;
    bool busy = false;
    int hi_waiting = 0;
    Fr* roots_fwd = nullptr;            // w_{2^11}^j, j < 1024
    Fr* roots_inv = nullptr;
    std::map<int, NttDomain*> domains;
    uint64_t launches = 0;              // kernels launched by this library (bench "gpu_launches")
    float msm_ms[5] = {0, 0, 0, 0, 0};  // last timed MSM: digits+hist, scan, scatter, accumulate, reduce
    bool time_msm = false;
    int msm_window_override = 0;        // 0 = pick by size
    int serial_above_log = 24;          // proofs with a larger domain run their five multiexps one after the other (B2G16_SERIAL_LOG)
    int red_log = 3;                    // log2 of the items per thread and level in the bucket reduction (B2G16_RED_LOG)
    int acc_variant = 102;              // bucket-accumulation kernel variant (B2G16_ACC_VARIANT)
    int red_bits = 1;                   // finish the bucket reduction by bit decomposition once <= 16384 items per window are left (B2G16_RED_BITS)
    int tree_levels = -1;               // batched-affine tree levels before the XYZZ accumulate: -1 = by bucket occupancy, 0 = off (B2G16_TREE_LEVELS)
    int tree_k = 16;                    // outputs per thread in a tree level (B2G16_TREE_K)
    int tree_k2 = 32;                   // values per thread in the upper levels of a tree level's product tree (B2G16_TREE_K2)
    int key_chunk_mib = 32;             // pinned ring slot of the streamed parameter-file loader, MiB ("key_chunk_mib")
    int tree_wgroup = 0;                // windows per tree pass; 0 = as many as fit the free memory (B2G16_TREE_WGROUP)
    void* fixed_base[2] = {nullptr, nullptr};   // generator tables (G1, G2), built on first use
    cudaEvent_t ev_stage[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // prover stage marks (created once, b2g16_init)
    // Launch log of the bucket-accumulation kernels (b2g16_set_kernel_log): one entry per multiexp, events recorded on
    // the launching stream right around the kernel, read back after the work has been synchronised.  This is how
    // bench.py takes the dominant kernel's duration from the proof's own launches instead of a separate run.
    struct KernelLogEntry { int group, c; uint64_t n; cudaEvent_t e0, e1; };
    bool klog_on = false;
    std::vector<KernelLogEntry> klog;
    uint32_t* d_klog_entries = nullptr;                   // sorted (point, window) entries of each logged launch, 64 slots
};

// RAII admission to a context: blocks while another call runs; ordinary calls also yield to waiting priority calls.
struct ContextGate {
    Ctx* c;
    ContextGate(Ctx* ctx, bool priority) : c(ctx) {
        std::unique_lock<std::mutex> l(c->mu);
        if (priority) ++c->hi_waiting;
        c->cv.wait(l, [&] { return !c->busy && (priority || c->hi_waiting == 0); });
        if (priority) --c->hi_waiting;
        c->busy = true;
    }
    ~ContextGate() {
        { std::lock_guard<std::mutex> l(c->mu); c->busy = false; }
        c->cv.notify_all();
    }
    ContextGate(const ContextGate&) = delete;
    ContextGate& operator=(const ContextGate&) = delete;
};

struct CudaError { cudaError_t e; const char* what; const char* file; int line; };

#define B2_CUDA(expr)                                                                         \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) throw ::b2::CudaError{_e, #expr, __FILE__, __LINE__};           \
    } while (0)

struct ArgError { int code; std::string msg; };
#define B2_REQUIRE(cond, code, msg)                                   \
    do { if (!(cond)) throw ::b2::ArgError{(code), (msg)}; } while (0)

// RAII stream-ordered device buffer
template <class T> struct DevBuf {
    T* p = nullptr; size_t n = 0; cudaStream_t s = nullptr;
    DevBuf() {}
    DevBuf(size_t count, cudaStream_t st) { alloc(count, st); }
    void alloc(size_t count, cudaStream_t st) {
        free(); n = count; s = st;
        if (count) B2_CUDA(cudaMallocAsync((void**)&p, count * sizeof(T), st));
    }
    void free() { if (p) { cudaFreeAsync(p, s); p = nullptr; n = 0; } }
    ~DevBuf() { free(); }
    DevBuf(const DevBuf&) = delete; DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n), s(o.s) { o.p = nullptr; }
    DevBuf& operator=(DevBuf&& o) noexcept { free(); p = o.p; n = o.n; s = o.s; o.p = nullptr; return *this; }
    T* get() const { return p; }
};

inline int grid_for(size_t n, int block, int max_blocks = 148 * 16) {
    size_t g = (n + block - 1) / block;
    if (g > (size_t)max_blocks) g = max_blocks;
    return g ? (int)g : 1;
}

// ---- entry points implemented across the .cu files (all operate on ctx->stream) ----
// ingest.cu : ABI byte formats <-> Montgomery limbs.  d_bad (optional) is OR-ed with 1 when a
// value is not canonical (>= modulus), with 2 when a point is not on the curve.
void fr_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, Fr* d_out, size_t n, int* d_bad);
void fr_to_bytes_dev(Ctx* c, const Fr* d_in, uint8_t* d_bytes, size_t n);
void g1_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, G1Affine* d_out, size_t n, int* d_bad, bool check_curve);
void g2_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, G2Affine* d_out, size_t n, int* d_bad, bool check_curve);
void g1_to_bytes_dev(Ctx* c, const G1Affine* d_in, uint8_t* d_bytes, size_t n, cudaStream_t st);
void g2_to_bytes_dev(Ctx* c, const G2Affine* d_in, uint8_t* d_bytes, size_t n, cudaStream_t st);
void scalars_check_dev(Ctx* c, const uint8_t* d_bytes, size_t n, int* d_bad);
void field_op_dev(Ctx* c, int field, int op, const uint8_t* d_a, const uint8_t* d_b, size_t n, uint8_t* d_out, int* d_bad);
void fr_dot_dev(Ctx* c, const uint8_t* d_a, const uint8_t* d_b, size_t n, uint8_t* d_out32, int* d_bad);   // sum a_i b_i mod r, canonical in/out
double mad32_peak_dev(Ctx* c, double* elapsed_ms);
// ntt.cu
NttDomain* ntt_domain(Ctx* c, int log_n);
void ntt_free_domains(Ctx* c);
void ntt_init_roots(Ctx* c);
// natural-order in/out transform (the evaluation_domain API); src may equal dst, tmp needs
// 2^log_n elements when log_n > 11; in_canon/out_canon fuse the canonical<->Montgomery steps.
void ntt_natural_ex(Ctx* c, const Fr* src, Fr* dst, Fr* tmp, int log_n, bool inverse, bool coset,
                    bool in_canon, bool out_canon, int* d_bad);
void ntt_natural(Ctx* c, Fr* d_data, int log_n, bool inverse, bool coset);
bool ntt_needs_tmp(int log_n);
// building blocks of the transform sharded over 2^log_g ranks (see ntt.cu "distributed transform")
void ntt_cross_dev(Ctx* c, const Fr* d_in, Fr* d_out, int log_n, int log_g, uint64_t chunk, bool inverse, bool coset, int* d_bad);
void ntt_cross_p2p_dev(Ctx* c, const Fr* const* peers_in, Fr* const* peers_out, int log_n, int log_g, uint64_t rank, bool inverse, bool coset, int* d_bad);
void fr_coset_scale_dev(Ctx* c, Fr* d_data, uint64_t count, int log_n, uint64_t offset, bool inverse, int* d_bad);
void fr_transpose_dev(Ctx* c, const Fr* d_in, Fr* d_out, uint64_t rows, uint64_t cols);
Fr host_root_of_unity(int log_n);
// witness-map chain: a,b,c evaluations (natural order; destroyed) -> H coefficients written
// to d_h_canonical as canonical 32-byte scalars in natural order.
void h_from_abc_dev(Ctx* c, Fr* d_a, Fr* d_b, Fr* d_c, int log_m, uint32_t* d_h_canonical, bool in_canon, int* d_bad);
// msm_g1.cu / msm_g2.cu : sum_i s_i P_i over resident Montgomery-affine bases and canonical
// device scalars; the affine result is written to d_out_bytes in the ABI encoding (96 / 192 B).
void msm_g1_dev(Ctx* c, const G1Affine* d_bases, const uint32_t* d_scalars, size_t n, uint8_t* d_out_bytes, cudaStream_t st, int window_bits);
void msm_g2_dev(Ctx* c, const G2Affine* d_bases, const uint32_t* d_scalars, size_t n, uint8_t* d_out_bytes, cudaStream_t st, int window_bits);
// same reduction, but the (at most 32) window sums are left in d_win for the caller to combine: result =
// sum_w 2^(c_bits w) d_win[w]; n_windows = 0 for the empty sum
void msm_g1_windows_dev(Ctx* c, const G1Affine* d_bases, const uint32_t* d_scalars, size_t n, G1Xyzz* d_win, int* c_bits, int* n_windows, cudaStream_t st);
void msm_g2_windows_dev(Ctx* c, const G2Affine* d_bases, const uint32_t* d_scalars, size_t n, G2Xyzz* d_win, int* c_bits, int* n_windows, cudaStream_t st);
void gen_mul_g1_dev(Ctx* c, const uint32_t* d_scalars, size_t n, G1Affine* d_out);
void gen_mul_g2_dev(Ctx* c, const uint32_t* d_scalars, size_t n, G2Affine* d_out);
void g1_subgroup_check_dev(Ctx* c, const G1Affine* d_pts, size_t n, int* d_bad);   // r * P == O, else *d_bad |= 4
void sum_g1_dev(Ctx* c, const G1Affine* d_pts, size_t n, uint8_t* d_out_bytes);
void sum_g2_dev(Ctx* c, const G2Affine* d_pts, size_t n, uint8_t* d_out_bytes);

}  // namespace b2
