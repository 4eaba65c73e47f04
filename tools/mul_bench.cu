// Candidate Fq multipliers on sm_100a, throughput + correctness against the production Fq:
//   rr28 : 14 x 28-bit limbs, 64-bit column accumulators, IMAD.WIDE without carry (full rate)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o tools/mul_bench tools/mul_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../crypto3-fil-proofs_b200/csrc/fp.cuh"
using namespace b2;

namespace rr {
constexpr int N = 14, W = 28;
constexpr uint32_t MASK = (1u << W) - 1;
__device__ __constant__ uint32_t kP[N] = {0xfffaaabu, 0xfefffffu, 0x3ffffb9u, 0xfffeb15u, 0x6241eabu, 0xa0f6b0fu, 0xf6730d2u,
                                          0xf38512bu, 0x4774b84u, 0x4bacd76u, 0xba7b643u, 0xe69a4b1u, 0x1ea397fu, 0x001a011u};
__device__ __constant__ uint32_t kR2[N] = {0x10370edu, 0x6d1c345u, 0xe243d62u, 0xec45c53u, 0x3b1d65au, 0x093317du, 0xb4f36a0u,
                                           0x5d74088u, 0xc10ea72u, 0x865d118u, 0x7320a75u, 0xfd5cd50u, 0xcc8a759u, 0x000c8d4u};
__device__ __constant__ uint32_t kInv = 0xffcfffdu;

struct F { uint32_t v[N]; };

__device__ __forceinline__ uint64_t madw(uint32_t a, uint32_t b, uint64_t c) {
    uint64_t r; asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c)); return r;
}
// r = a*b / 2^392 mod p, result limbs < 2^28 (+1 carry into the top limb), value < 2p for inputs < 32p
template <bool CONSTP> __device__ __forceinline__ F mul(const F& a, const F& b) {
    uint64_t t[N + 1];
    uint32_t p[N];
#pragma unroll
    for (int j = 0; j < N; ++j) { t[j] = 0; p[j] = kP[j]; }
    t[N] = 0;
    const uint32_t inv = kInv;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const uint32_t bi = b.v[i];
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = madw(a.v[j], bi, t[j]);
        const uint32_t m = ((uint32_t)t[0] * inv) & MASK;
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = madw(m, p[j], t[j]);
        t[1] += t[0] >> W;
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = t[j + 1];
        t[N] = 0;
    }
    F r;
#pragma unroll
    for (int j = 0; j < N - 1; ++j) { r.v[j] = (uint32_t)t[j] & MASK; t[j + 1] += t[j] >> W; }
    r.v[N - 1] = (uint32_t)t[N - 1];
    return r;
}
__device__ F from_fq_canonical(const uint32_t* c /*12 x 32*/) {
    F r;
    for (int j = 0; j < N; ++j) {
        int bit = j * W, i = bit >> 5, sh = bit & 31;
        uint64_t x = c[i];
        if (i + 1 < 12) x |= (uint64_t)c[i + 1] << 32;
        r.v[j] = (uint32_t)(x >> sh) & MASK;
    }
    return r;
}
__device__ void to_canonical32(const F& a, uint32_t* c /*12*/) {
    // a < 2p, limbs normalised: subtract p if >= p, then repack
    uint32_t d[N]; int64_t borrow = 0;
    for (int j = 0; j < N; ++j) { int64_t x = (int64_t)a.v[j] - kP[j] + borrow; d[j] = (uint32_t)x & MASK; borrow = x >> W; }
    const uint32_t* s = borrow < 0 ? a.v : d;
    for (int i = 0; i < 12; ++i) c[i] = 0;
    for (int j = 0; j < N; ++j) {
        int bit = j * W, i = bit >> 5, sh = bit & 31;
        uint64_t x = (uint64_t)s[j] << sh;
        c[i] |= (uint32_t)x;
        if (i + 1 < 12) c[i + 1] |= (uint32_t)(x >> 32);
    }
}
}  // namespace rr

__global__ void check_kernel(int* bad, int n) {
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    uint32_t x[12], y[12];
    uint32_t s = id * 2654435761u + 12345u;
    for (int i = 0; i < 12; ++i) { s = s * 1664525u + 1013904223u; x[i] = s; s = s * 1664525u + 1013904223u; y[i] = s; }
    x[11] &= 0x0fffffffu; y[11] &= 0x0fffffffu;          // < 2^380 < p
    if (id == 0) { for (int i = 0; i < 12; ++i) x[i] = FqParams::P(i); x[0] -= 1; }   // p - 1
    Fq fx = Fq::from_canonical(x), fy = Fq::from_canonical(y);
    uint32_t want[12]; (fx * fy).to_canonical(want);
    rr::F a = rr::from_fq_canonical(x), b = rr::from_fq_canonical(y), r2;
    for (int j = 0; j < rr::N; ++j) r2.v[j] = rr::kR2[j];
    rr::F am = rr::mul<true>(a, r2);
    rr::F r = rr::mul<true>(am, b);
    uint32_t got[12]; rr::to_canonical32(r, got);
    for (int i = 0; i < 12; ++i) if (got[i] != want[i]) { atomicAdd(bad, 1); break; }
}

template <int CH> __global__ void __launch_bounds__(256) k_rr(uint32_t* sink, int iters) {
    rr::F x[CH], y;
    for (int i = 0; i < rr::N; ++i) { y.v[i] = (threadIdx.x * 2654435761u + i) & rr::MASK; for (int c = 0; c < CH; ++c) x[c].v[i] = (i * 7 + c) & rr::MASK; }
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int c = 0; c < CH; ++c) x[c] = rr::mul<true>(x[c], y);
    uint32_t s = 0; for (int c = 0; c < CH; ++c) for (int i = 0; i < rr::N; ++i) s ^= x[c].v[i];
    if (s == 0x1234567u) sink[0] = s;
}
template <int CH> __global__ void __launch_bounds__(256) k_fq(uint32_t* sink, int iters) {
    Fq x[CH], y;
    for (int i = 0; i < 12; ++i) { y.v[i] = threadIdx.x * 2654435761u + i; for (int c = 0; c < CH; ++c) x[c].v[i] = i * 7 + c; }
    y.v[11] &= 0x0fffffffu;
    for (int c = 0; c < CH; ++c) x[c].v[11] &= 0x0fffffffu;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int c = 0; c < CH; ++c) x[c] = x[c] * y;
    uint32_t s = 0; for (int c = 0; c < CH; ++c) for (int i = 0; i < 12; ++i) s ^= x[c].v[i];
    if (s == 0x1234567u) sink[0] = s;
}


// ---- fd24: Fq product on the FP64 pipe.  16 limbs of 24 bits; every DFMA below is exact (column sums
// stay below 2^53: at most 32 products of < 2^48 plus a carry), Montgomery radix 2^384 like the
// 32-bit-limb integer path, so both paths share one Montgomery representation.
namespace fd {
constexpr int N = 16;
__device__ __constant__ double kP[N] = {16755371.0, 16777215.0, 16759294.0, 11621375.0, 11272190.0, 16131102.0, 10548912.0, 6762706.0,
                                        8721087.0, 4949235.0, 14115959.0, 4410284.0, 1812406.0, 15112779.0, 15350143.0, 1704209.0};
constexpr uint32_t INV24 = 0xfcfffdu, MASK = 0xffffffu;
constexpr double TWO52 = 4503599627370496.0, INV_2_24 = 1.0 / 16777216.0, TWO24 = 16777216.0;
struct F { uint32_t v[N]; };

__device__ __forceinline__ double u2d(uint32_t x) { return __hiloint2double(0x43300000, (int)x) - TWO52; }
__device__ __forceinline__ uint32_t d2u(double x) { return (uint32_t)__double2loint(x + TWO52); }      // 0 <= x < 2^32, integer valued
__device__ __forceinline__ double floor24(double t) { return __fma_rz(t, INV_2_24, TWO52) - TWO52; }     // floor(t / 2^24), 0 <= t < 2^53

// r = a * b / 2^384 mod p; limbs of a, b < 2^24, values < 2p  ->  limbs < 2^24, value < 1.5 p
__device__ __forceinline__ F mul(const F& a, const F& b) {
    double ad[N], t[N + 1];
#pragma unroll
    for (int j = 0; j < N; ++j) { ad[j] = u2d(a.v[j]); t[j] = 0.0; }
    t[N] = 0.0;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const double bi = u2d(b.v[i]);
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = fma(ad[j], bi, t[j]);
        const double q = floor24(t[0]);
        const double lo = fma(q, -TWO24, t[0]);
        const uint32_t m = (d2u(lo) * INV24) & MASK;
        const double md = u2d(m);
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = fma(md, kP[j], t[j]);
        t[1] = fma(t[0], INV_2_24, t[1]);                  // t[0] is now a multiple of 2^24: exact
#pragma unroll
        for (int j = 0; j < N; ++j) t[j] = t[j + 1];
        t[N] = 0.0;
    }
    F r;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        const double q = floor24(t[j]);
        r.v[j] = d2u(fma(q, -TWO24, t[j]));
        if (j + 1 < N) t[j + 1] += q;
    }
    return r;
}
__device__ F from_fq_mont(const Fq& x) {                  // same value, 24-bit limbs
    F r;
    for (int j = 0; j < N; ++j) {
        int bit = 24 * j, i = bit >> 5, sh = bit & 31;
        uint64_t w = x.v[i];
        if (i + 1 < 12) w |= (uint64_t)x.v[i + 1] << 32;
        r.v[j] = (uint32_t)(w >> sh) & MASK;
    }
    return r;
}
__device__ Fq to_fq_mont(const F& a) {                    // a < 2p -> canonical Montgomery limbs
    uint32_t c[12];
    for (int i = 0; i < 12; ++i) c[i] = 0;
    for (int j = 0; j < N; ++j) {
        int bit = 24 * j, i = bit >> 5, sh = bit & 31;
        uint64_t w = (uint64_t)a.v[j] << sh;
        c[i] |= (uint32_t)w;
        if (i + 1 < 12) c[i + 1] |= (uint32_t)(w >> 32);
    }
    Fq r; Fq::final_sub(r.v, c, 0);
    return r;
}
}  // namespace fd

__global__ void check_fd_kernel(int* bad, int n) {
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    uint32_t x[12], y[12];
    uint32_t s = id * 2654435761u + 777u;
    for (int i = 0; i < 12; ++i) { s = s * 1664525u + 1013904223u; x[i] = s; s = s * 1664525u + 1013904223u; y[i] = s; }
    x[11] &= 0x0fffffffu; y[11] &= 0x0fffffffu;
    if (id == 0) { for (int i = 0; i < 12; ++i) x[i] = y[i] = FqParams::P(i); x[0] -= 1; y[0] -= 1; }
    Fq fx = Fq::from_canonical(x), fy = Fq::from_canonical(y);
    Fq want = fx * fy * fy;
    fd::F a = fd::from_fq_mont(fx), b = fd::from_fq_mont(fy);
    Fq got = fd::to_fq_mont(fd::mul(fd::mul(a, b), b));
    if (got != want) atomicAdd(bad, 1);
}
template <int CH> __global__ void __launch_bounds__(256) k_fd(uint32_t* sink, int iters) {
    fd::F x[CH], y;
    for (int i = 0; i < fd::N; ++i) { y.v[i] = (threadIdx.x * 2654435761u + i) & fd::MASK; for (int c = 0; c < CH; ++c) x[c].v[i] = (i * 7 + c) & fd::MASK; }
    y.v[15] &= 0xfffff; for (int c = 0; c < CH; ++c) x[c].v[15] &= 0xfffff;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int c = 0; c < CH; ++c) x[c] = fd::mul(x[c], y);
    uint32_t s = 0; for (int c = 0; c < CH; ++c) for (int i = 0; i < fd::N; ++i) s ^= x[c].v[i];
    if (s == 0x1234567u) sink[0] = s;
}
// warp-specialised: even warps multiply on the integer pipe, odd warps on the FP64 pipe
__global__ void __launch_bounds__(256) k_both(uint32_t* sink, int it_int, int it_fd) {
    const int warp = threadIdx.x >> 5;
    uint32_t s = 0;
    if (warp & 1) {
        fd::F x, y;
        for (int i = 0; i < fd::N; ++i) { y.v[i] = (threadIdx.x * 2654435761u + i) & fd::MASK; x.v[i] = (i * 7) & fd::MASK; }
        y.v[15] &= 0xfffff; x.v[15] &= 0xfffff;
        for (int it = 0; it < it_fd; ++it) x = fd::mul(x, y);
        for (int i = 0; i < fd::N; ++i) s ^= x.v[i];
    } else {
        Fq x, y;
        for (int i = 0; i < 12; ++i) { y.v[i] = threadIdx.x * 2654435761u + i; x.v[i] = i * 7; }
        y.v[11] &= 0x0fffffffu; x.v[11] &= 0x0fffffffu;
        for (int it = 0; it < it_int; ++it) x = x * y;
        for (int i = 0; i < 12; ++i) s ^= x.v[i];
    }
    if (s == 0x1234567u) sink[0] = s;
}

// ---- raw pipe rates with operands that cannot be hoisted (each product reads a live accumulator)
__global__ void __launch_bounds__(256) k_wide_live(uint64_t* sink, int iters) {
    uint32_t a = threadIdx.x * 2654435761u + 1;
    uint64_t acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"((uint32_t)acc[(k + 3) & 7]));
    uint64_t s = 0; for (int k = 0; k < 8; ++k) s ^= acc[k];
    if (s == 0x1234567u) sink[0] = s;
}
__global__ void __launch_bounds__(256) k_wide_mul_only(uint64_t* sink, int iters) {     // IMAD.WIDE with RZ addend, xor-folded
    uint32_t a = threadIdx.x * 2654435761u + 1;
    uint64_t acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc[k]) : "r"((uint32_t)(acc[k] >> 32)), "r"((uint32_t)acc[(k + 3) & 7]));
    uint64_t s = 0; for (int k = 0; k < 8; ++k) s ^= acc[k];
    if (s == 0x1234567u) sink[0] = s;
}
__global__ void __launch_bounds__(256) k_imad32_live(uint32_t* sink, int iters) {
    uint32_t a = threadIdx.x * 2654435761u + 1;
    uint32_t acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(acc[k]) : "r"(a), "r"(acc[(k + 3) & 7]));
    uint32_t s = 0; for (int k = 0; k < 8; ++k) s ^= acc[k];
    if (s == 0x1234567u) sink[0] = s;
}
__global__ void __launch_bounds__(256) k_dfma_live(double* sink, int iters) {
    double a = 1.0 + threadIdx.x * 1e-9;
    double acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = __fma_rz(a, acc[(k + 3) & 7], acc[k]);
    double s = 0; for (int k = 0; k < 8; ++k) s += acc[k];
    if (s == 0.1234567) sink[0] = s;
}
// IMAD.WIDE and DFMA chains in the same warp (do the two pipes overlap?)
__global__ void __launch_bounds__(256) k_mix_live(double* sink, int iters) {
    uint32_t a = threadIdx.x * 2654435761u + 1;
    double fa = 1.0 + threadIdx.x * 1e-9;
    uint64_t acc[8]; double facc[8];
    for (int k = 0; k < 8; ++k) { acc[k] = a + k; facc[k] = fa + k; }
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc[k]) : "r"((uint32_t)(acc[k] >> 32)), "r"((uint32_t)acc[(k + 3) & 7]));
                facc[k] = __fma_rz(fa, facc[(k + 3) & 7], facc[k]);
            }
    double s = 0; for (int k = 0; k < 8; ++k) s += facc[k] + (double)acc[k];
    if (s == 0.1234567) sink[0] = s;
}
template <class K> double run(K launch, double ops) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ops / (ms * 1e-3);
}
int main() {
    void* sink; cudaMalloc(&sink, 64);
    int* bad; cudaMalloc(&bad, 4); cudaMemset(bad, 0, 4);
    check_kernel<<<64, 256>>>(bad, 64 * 256);
    int hb = -1; cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
    printf("{\"rr28_mismatches\": %d}\n", hb);
    cudaMemset(bad, 0, 4);
    check_fd_kernel<<<64, 256>>>(bad, 64 * 256);
    cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
    printf("{\"fd24_mismatches\": %d}\n", hb);
    for (int bps : {1, 2, 3}) {
        const int blocks = 148 * bps, threads = 256; const double T = (double)blocks * threads;
        double f1 = run([&] { k_fd<1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double q1 = run([&] { k_fq<1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        printf("{\"warps_per_smsp\": %d, \"fd24_mul_per_s\": %.4g, \"fq_mul_per_s\": %.4g", bps * 2, f1, q1);
        for (int fdit : {256, 384, 512}) {
            double b = run([&] { k_both<<<blocks, threads>>>((uint32_t*)sink, 512, fdit); }, T / 2 * (512 + fdit));
            printf(", \"both_512_%d\": %.4g", fdit, b);
        }
        printf("}\n");
    }
    for (int bps : {1, 2, 4}) {
        const int blocks = 148 * bps, threads = 256;
        const double T = (double)blocks * threads;
        double fq1 = run([&] { k_fq<1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double rr1 = run([&] { k_rr<1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double rr2 = run([&] { k_rr<2><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512 * 2);
        printf("{\"warps_per_smsp\": %d, \"fq_mul1\": %.4g, \"rr28_mul1\": %.4g, \"rr28_mul2\": %.4g}\n", bps * 2, fq1, rr1, rr2);
    }
    {
        const int blocks = 148 * 8, threads = 256; const double T = (double)blocks * threads * 2048 * 64;
        double w = run([&] { k_wide_live<<<blocks, threads>>>((uint64_t*)sink, 2048); }, T);
        double wm = run([&] { k_wide_mul_only<<<blocks, threads>>>((uint64_t*)sink, 2048); }, T);
        double i32 = run([&] { k_imad32_live<<<blocks, threads>>>((uint32_t*)sink, 2048); }, T);
        double df = run([&] { k_dfma_live<<<blocks, threads>>>((double*)sink, 2048); }, T);
        double mx = run([&] { k_mix_live<<<blocks, threads>>>((double*)sink, 2048); }, T);
        printf("{\"mad_wide_live_per_s\": %.4g, \"mul_wide_only_per_s\": %.4g, \"imad32_per_s\": %.4g, \"dfma_per_s\": %.4g, \"mix_pairs_per_s\": %.4g}\n", w, wm, i32, df, mx);
    }
    printf("cuda status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
