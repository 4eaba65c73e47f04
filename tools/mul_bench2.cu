// Throughput + correctness of the split / Karatsuba / lazy-reduction multipliers of csrc/fp_wide.cuh against the
// production CIOS product (fp.cuh) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o tools/mul_bench2 tools/mul_bench2.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../crypto3-fil-proofs_b200/csrc/fp_wide.cuh"
using namespace b2;

__global__ void check_kernel(int* bad, int n) {
    int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    Fq a, b, c, d;
    uint32_t s = id * 2654435761u + 12345u;
    for (int i = 0; i < 12; ++i) {
        s = s * 1664525u + 1013904223u; a.v[i] = s; s = s * 1664525u + 1013904223u; b.v[i] = s;
        s = s * 1664525u + 1013904223u; c.v[i] = s; s = s * 1664525u + 1013904223u; d.v[i] = s;
    }
    a.v[11] &= 0x0fffffffu; b.v[11] &= 0x0fffffffu; c.v[11] &= 0x0fffffffu; d.v[11] &= 0x0fffffffu;
    if (id == 0) { for (int i = 0; i < 12; ++i) a.v[i] = b.v[i] = c.v[i] = d.v[i] = FqParams::P(i); a.v[0] -= 1; b.v[0] -= 1; c.v[0] -= 1; d.v[0] -= 2; }
    const Fq want = a * b;
    if (wide::fq_mul<false>(a, b) != want) atomicAdd(bad, 1);
    if (wide::fq_mul<true>(a, b) != want) atomicAdd(bad + 1, 1);
    const Fq2 x{a, c}, y{b, d};
    const Fq2 w2 = x.mul_inline(y);
    if (wide::fq2_mul_lazy<false>(x, y) != w2) atomicAdd(bad + 2, 1);
    if (wide::fq2_mul_lazy<true>(x, y) != w2) atomicAdd(bad + 3, 1);
}

template <int MODE> __global__ void __launch_bounds__(256) k_fq(uint32_t* sink, int iters) {
    Fq x, y;
    for (int i = 0; i < 12; ++i) { y.v[i] = threadIdx.x * 2654435761u + i; x.v[i] = i * 7 + blockIdx.x; }
    y.v[11] &= 0x0fffffffu; x.v[11] &= 0x0fffffffu;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) x = x * y;
        else if (MODE == 1) x = wide::fq_mul<false>(x, y);
        else x = wide::fq_mul<true>(x, y);
    }
    uint32_t s = 0; for (int i = 0; i < 12; ++i) s ^= x.v[i];
    if (s == 0x1234567u) sink[0] = s;
}
template <int MODE> __global__ void __launch_bounds__(128) k_fq2(uint32_t* sink, int iters) {
    Fq2 x, y;
    for (int i = 0; i < 12; ++i) { y.c0.v[i] = threadIdx.x * 2654435761u + i; y.c1.v[i] = threadIdx.x * 40503u + i; x.c0.v[i] = i * 7 + blockIdx.x; x.c1.v[i] = i * 13 + 5; }
    y.c0.v[11] &= 0x0fffffffu; y.c1.v[11] &= 0x0fffffffu; x.c0.v[11] &= 0x0fffffffu; x.c1.v[11] &= 0x0fffffffu;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) x = x.mul_inline(y);
        else if (MODE == 1) x = wide::fq2_mul_lazy<false>(x, y);
        else x = wide::fq2_mul_lazy<true>(x, y);
    }
    uint32_t s = 0; for (int i = 0; i < 12; ++i) s ^= x.c0.v[i] ^ x.c1.v[i];
    if (s == 0x1234567u) sink[0] = s;
}
template <class K> double run(K launch, double ops) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return ops / (best * 1e-3);
}
int main() {
    void* sink; cudaMalloc(&sink, 64);
    int* bad; cudaMalloc(&bad, 16); cudaMemset(bad, 0, 16);
    check_kernel<<<256, 256>>>(bad, 256 * 256);
    int hb[4] = {-1, -1, -1, -1}; cudaMemcpy(hb, bad, 16, cudaMemcpyDeviceToHost);
    printf("{\"mismatches_split\": %d, \"mismatches_karatsuba\": %d, \"mismatches_fq2_lazy\": %d, \"mismatches_fq2_lazy_karatsuba\": %d}\n", hb[0], hb[1], hb[2], hb[3]);
    for (int bps : {1, 2, 4}) {
        const int blocks = 148 * bps, threads = 256; const double T = (double)blocks * threads;
        double q0 = run([&] { k_fq<0><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double q1 = run([&] { k_fq<1><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        double q2 = run([&] { k_fq<2><<<blocks, threads>>>((uint32_t*)sink, 512); }, T * 512);
        printf("{\"warps_per_smsp\": %d, \"fq_cios_per_s\": %.4g, \"fq_split_per_s\": %.4g, \"fq_karatsuba_per_s\": %.4g}\n", bps * 2, q0, q1, q2);
    }
    for (int bps : {2, 4, 8}) {
        const int blocks = 148 * bps, threads = 128; const double T = (double)blocks * threads;
        double q0 = run([&] { k_fq2<0><<<blocks, threads>>>((uint32_t*)sink, 256); }, T * 256);
        double q1 = run([&] { k_fq2<1><<<blocks, threads>>>((uint32_t*)sink, 256); }, T * 256);
        double q2 = run([&] { k_fq2<2><<<blocks, threads>>>((uint32_t*)sink, 256); }, T * 256);
        printf("{\"warps_per_smsp\": %d, \"fq2_karatsuba3_per_s\": %.4g, \"fq2_lazy_per_s\": %.4g, \"fq2_lazy_karatsuba_per_s\": %.4g}\n", bps, q0, q1, q2);
    }
    printf("cuda status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
