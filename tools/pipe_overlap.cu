// Do the integer-multiply pipe and the FP64 pipe overlap across warps on sm_100a?
// Even warps run integer chains (IMAD.WIDE, or IMAD.WIDE.X carry chains), odd warps DFMA chains.
#include <cstdio>
#include <cuda_runtime.h>
#include "../crypto3-fil-proofs_b200/csrc/fp.cuh"
using namespace b2;

template <int MODE> __device__ __forceinline__ uint32_t int_work(int iters) {
    uint32_t s = 0;
    if (MODE == 0) {            // IMAD.WIDE, live operands
        uint32_t a = threadIdx.x * 2654435761u + 1;
        uint64_t acc[8];
        for (int k = 0; k < 8; ++k) acc[k] = a + k;
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int k = 0; k < 8; ++k) asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc[k]) : "r"((uint32_t)(acc[k] >> 32)), "r"((uint32_t)acc[(k + 3) & 7]));
        for (int k = 0; k < 8; ++k) s ^= (uint32_t)acc[k] ^ (uint32_t)(acc[k] >> 32);
    } else {                    // IMAD.WIDE.X carry chains (rows of 6 on 12 limbs, 64 products per iteration..)
        uint32_t a[12], acc[12];
        for (int i = 0; i < 12; ++i) { a[i] = threadIdx.x * 2654435761u + i; acc[i] = i; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int r = 0; r < 11; ++r) {          // 66 products ~ 64
                uint32_t b = acc[r];
                acc[0] = ptx::mad_lo_cc(a[0], b, acc[0]); acc[1] = ptx::madc_hi_cc(a[0], b, acc[1]);
#pragma unroll
                for (int j = 2; j < 12; j += 2) { acc[j] = ptx::madc_lo_cc(a[j], b, acc[j]); acc[j + 1] = ptx::madc_hi_cc(a[j], b, acc[j + 1]); }
            }
        }
        for (int i = 0; i < 12; ++i) s ^= acc[i];
    }
    return s;
}
__device__ __forceinline__ double fp_work(int iters) {
    double a = 1.0 + threadIdx.x * 1e-9, acc[8];
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = __fma_rz(a, acc[(k + 3) & 7], acc[k]);
    double s = 0; for (int k = 0; k < 8; ++k) s += acc[k];
    return s;
}
// which: 1 = int warps only work, 2 = fp warps only, 3 = both
template <int MODE> __global__ void __launch_bounds__(256) k(double* sink, int which, int it_int, int it_fp) {
    const int warp = threadIdx.x >> 5;
    if (warp & 1) { if (which & 2) { double s = fp_work(it_fp); if (s == 0.1234567) sink[0] = s; } }
    else { if (which & 1) { uint32_t s = int_work<MODE>(it_int); if (s == 0x1234567u) sink[1] = s; } }
}
template <class K> float run(K launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
    double* sink; cudaMalloc(&sink, 64);
    for (int bps : {2, 4}) {
        const int blocks = 148 * bps, threads = 256, it = 4096;
        float i0 = run([&] { k<0><<<blocks, threads>>>(sink, 1, it, it); });
        float i1 = run([&] { k<1><<<blocks, threads>>>(sink, 1, it, it); });
        float f = run([&] { k<0><<<blocks, threads>>>(sink, 2, it, it); });
        float b0 = run([&] { k<0><<<blocks, threads>>>(sink, 3, it, it); });
        float b1 = run([&] { k<1><<<blocks, threads>>>(sink, 3, it, it); });
        printf("{\"warps_per_smsp_each\": %d, \"ms_imad_wide_only\": %.3f, \"ms_imad_wide_x_only\": %.3f, \"ms_dfma_only\": %.3f, \"ms_wide_plus_dfma\": %.3f, \"ms_wide_x_plus_dfma\": %.3f}\n",
               bps, i0, i1, f, b0, b1);
    }
    printf("cuda status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
