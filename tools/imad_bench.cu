// Micro-benchmarks of the integer multiply-add pipe on sm_100a (MSM / NTT roofline inputs):
// IMAD.WIDE without carry, IMAD.WIDE.X carry chains, and whole Fr / Fq Montgomery products.
#include <cstdio>
#include <cuda_runtime.h>
#include "../crypto3-fil-proofs_b200/csrc/fp.cuh"
using namespace b2;

__global__ void __launch_bounds__(256) k_wide(uint64_t* sink, int iters) {
    uint32_t a = threadIdx.x * 2654435761u + 1, b = a * 747796405u + 2891336453u;
    uint64_t acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
    uint64_t s = 0; for (int k = 0; k < 8; ++k) s ^= acc[k];
    if (s == 0x1234567u) sink[0] = s;
}
// rows of 6 carry-chained wide MADs on a 12-limb accumulator, CH independent accumulators
template <int CH> __global__ void __launch_bounds__(256) k_carry(uint32_t* sink, int iters) {
    uint32_t a[12], acc[CH][12];
    for (int i = 0; i < 12; ++i) { a[i] = threadIdx.x * 2654435761u + i; for (int c = 0; c < CH; ++c) acc[c][i] = i + c; }
    uint32_t b = a[3] * 747796405u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                acc[c][0] = ptx::mad_lo_cc(a[0], b, acc[c][0]); acc[c][1] = ptx::madc_hi_cc(a[0], b, acc[c][1]);
#pragma unroll
                for (int j = 2; j < 12; j += 2) { acc[c][j] = ptx::madc_lo_cc(a[j], b, acc[c][j]); acc[c][j + 1] = ptx::madc_hi_cc(a[j], b, acc[c][j + 1]); }
            }
    }
    uint32_t s = 0; for (int c = 0; c < CH; ++c) for (int i = 0; i < 12; ++i) s ^= acc[c][i];
    if (s == 0x1234567u) sink[0] = s;
}
template <class F, int CH> __global__ void __launch_bounds__(256) k_mul(uint32_t* sink, int iters) {
    F x[CH], y;
    for (int i = 0; i < F::N; ++i) { y.v[i] = threadIdx.x * 2654435761u + i; for (int c = 0; c < CH; ++c) x[c].v[i] = i * 7 + c; }
    y.v[F::N - 1] &= 0x0fffffffu;
    for (int c = 0; c < CH; ++c) x[c].v[F::N - 1] &= 0x0fffffffu;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int c = 0; c < CH; ++c) x[c] = x[c] * y;
    uint32_t s = 0; for (int c = 0; c < CH; ++c) for (int i = 0; i < F::N; ++i) s ^= x[c].v[i];
    if (s == 0x1234567u) sink[0] = s;
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
    for (int bps : {1, 2, 4, 8}) {
        const int blocks = 148 * bps, threads = 256;
        const double T = (double)blocks * threads;
        double w = run([&] { k_wide<<<blocks, threads>>>((uint64_t*)sink, 2048); }, T * 2048 * 64);
        double c1 = run([&] { k_carry<1><<<blocks, threads>>>((uint32_t*)sink, 2048); }, T * 2048 * 4 * 6);
        double c2 = run([&] { k_carry<2><<<blocks, threads>>>((uint32_t*)sink, 2048); }, T * 2048 * 4 * 6 * 2);
        double c4 = run([&] { k_carry<4><<<blocks, threads>>>((uint32_t*)sink, 2048); }, T * 2048 * 4 * 6 * 4);
        double fq1 = run([&] { k_mul<Fq, 1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double fq2 = run([&] { k_mul<Fq, 2><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512 * 2);
        double fr1 = run([&] { k_mul<Fr, 1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double fr2 = run([&] { k_mul<Fr, 2><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512 * 2);
        printf("{\"warps_per_smsp\": %d, \"wide_nocarry_per_s\": %.4g, \"wide_carry1\": %.4g, \"wide_carry2\": %.4g, \"wide_carry4\": %.4g, "
               "\"fq_mul1\": %.4g, \"fq_mul2\": %.4g, \"fr_mul1\": %.4g, \"fr_mul2\": %.4g}\n", bps * 2, w, c1, c2, c4, fq1, fq2, fr1, fr2);
    }
    printf("cuda status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
