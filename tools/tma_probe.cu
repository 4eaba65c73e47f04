// Isolated check of the tensor-map tile load used by ntt_pass_kernel<..., TMA>: a (2, 2, N) u64 view of N 32-byte
// elements, boxes of (2, 1, 256) = 256 element halves, into two 32 KB planes.  argv[1]: swizzle (0 none, 3 = 128B).
// Prints whether plane slot swz(e) holds half h of element e.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o tools/tma_probe tools/tma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

struct alignas(64) Tmap { unsigned char b[128]; };
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256) probe(const __grid_constant__ Tmap tmap, uint4* out, int tile) {
    extern __shared__ __align__(1024) uint4 smem[];
    __shared__ __align__(8) uint64_t mbar;
    uint4* s_lo = smem; uint4* s_hi = smem + 2048;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(1), "r"(s32(&mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(65536), "r"(s32(&mbar)) : "memory");
        for (int h = 0; h < 2; ++h)
            for (int q = 0; q < 8; ++q)
                asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                             ::"r"(s32((h ? s_hi : s_lo) + q * 256)), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(s32(&mbar)), "r"(0), "r"(h), "r"(tile * 2048 + q * 256) : "memory");
    }
    uint32_t ok = 0, spins = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(ok) : "r"(s32(&mbar)), "r"(0) : "memory");
        if (++spins > (1u << 24)) __trap();
    }
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) out[i] = smem[i];
}

int main(int argc, char** argv) {
    const int swz = argc > 1 ? atoi(argv[1]) : 3;
    const uint64_t N = 1 << 14;
    std::vector<uint32_t> h(N * 8);
    for (uint64_t e = 0; e < N; ++e) for (int w = 0; w < 8; ++w) h[e * 8 + w] = (uint32_t)(e * 16 + w);
    uint32_t* d; cudaMalloc(&d, N * 32); cudaMemcpy(d, h.data(), N * 32, cudaMemcpyHostToDevice);
    uint4* dout; cudaMalloc(&dout, 65536);
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    Tmap tm{};
    const cuuint64_t gdim[3] = {2, 2, N}, gstride[2] = {16, 32};
    const cuuint32_t box[3] = {2, 1, 256}, es[3] = {1, 1, 1};
    CUresult r = ((EncodeFn)fn)((CUtensorMap*)&tm, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, d, gdim, gstride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                (CUtensorMapSwizzle)swz, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d swizzle=%d\n", (int)r, swz);
    if (r != CUDA_SUCCESS) return 2;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    const int tile = 3;
    probe<<<1, 256, 65536>>>(tm, dout, tile);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 3;
    std::vector<uint32_t> o(16384);
    cudaMemcpy(o.data(), dout, 65536, cudaMemcpyDeviceToHost);
    int bad_plain = 0, bad_swz = 0;
    for (int hh = 0; hh < 2; ++hh) for (int e = 0; e < 2048; ++e) {
        const int es2 = e ^ ((e >> 3) & 7);
        const uint32_t want = (uint32_t)((tile * 2048 + e) * 16 + hh * 4);
        if (o[(hh * 2048 + e) * 4] != want) ++bad_plain;
        if (o[(hh * 2048 + es2) * 4] != want) ++bad_swz;
    }
    printf("mismatches: plain layout %d, swizzled layout (e ^ ((e >> 3) & 7)) %d; first words %u %u %u %u\n", bad_plain, bad_swz, o[0], o[4], o[8], o[32 * 4]);
    return 0;
}
