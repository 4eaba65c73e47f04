// An entry point that allocates nothing: b2g16_fr_dot_device_noalloc.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It is the
// no-allocation variant of b2g16_fr_dot_device (ingest.cu: out = sum_i a_i b_i mod r over canonical scalars -- the
// scalar side of the multiexp linearity identity), whose per-block partial sums and "bad input" flag normally live in
// stream-ordered allocations.  Here they live in ONE file-scope device array of fixed size:
//
//     header (256 B: the non-canonical-input flag) + B2G16_NOALLOC_MAX_BLOCKS partial sums of 32 B  =  131,328 bytes
//
// held by every GPU that loads the library until the process exits.  One block reduces 4096 elements, so the largest
// input is B2G16_NOALLOC_MAX_N = 4096 * 4096 = 2^24 scalars per operand (512 MiB each); longer inputs are refused.
// In host code the array's name is only a host-side stand-in for the symbol, not a device address: the address is
// taken once per device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.
// Every call on a device shares the array and finds the previous call's contents in it: the flag is cleared on the
// stream at the start of each call, the partial sums are fully overwritten before they are read.  Two calls that
// overlap on the GPU (different streams / contexts on one device) race on the array: serialise them.
// The call enqueues a memset and two kernels on the context's stream and returns; it calls no cudaMalloc / cudaFree /
// cudaMallocAsync and does not synchronise, so it can be captured into a CUDA graph.
#include <mutex>

#include "api_util.cuh"
#include "ntt_kernels.cuh"

namespace b2 {

constexpr int NOALLOC_HEADER_BYTES = 256;
constexpr int NOALLOC_PER_BLOCK = 4096;                       // elements one 256-thread block reduces (16 per thread)
constexpr int NOALLOC_MAX_BLOCKS = B2G16_NOALLOC_MAX_BLOCKS;  // 4096
static_assert((uint64_t)NOALLOC_MAX_BLOCKS * NOALLOC_PER_BLOCK == B2G16_NOALLOC_MAX_N, "header and kernel disagree on the maximum");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_HEADER_BYTES + NOALLOC_MAX_BLOCKS * sizeof(Fr)];

// scratch layout, from the device address of g_noalloc_scratch
B2_D int* noalloc_flag(unsigned char* s) { return reinterpret_cast<int*>(s); }
B2_D Fr* noalloc_partials(unsigned char* s) { return reinterpret_cast<Fr*>(s + NOALLOC_HEADER_BYTES); }

// block b: partial[b] = sum over its 4096 elements of a_i b_i / R (Montgomery product of canonical operands)
__global__ void __launch_bounds__(256) fr_dot_partial_noalloc_kernel(const Fr* __restrict__ a, const Fr* __restrict__ b, size_t n, unsigned char* scratch) {
    __shared__ Fr sh[256];
    Fr acc = Fr::zero();
    int bad = 0;
    const size_t base = (size_t)blockIdx.x * NOALLOC_PER_BLOCK;
    for (int k = 0; k < NOALLOC_PER_BLOCK / 256; ++k) {
        const size_t i = base + (size_t)k * 256 + threadIdx.x;
        if (i < n) {
            const Fr x = fr_load(a + i), y = fr_load(b + i);
            if (Fr::geq_p(x.v) || Fr::geq_p(y.v)) bad = 1;
            acc = acc + x * y;
        }
    }
    if (bad) atomicOr(noalloc_flag(scratch), 1);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = sh[threadIdx.x] + sh[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) fr_store(noalloc_partials(scratch) + blockIdx.x, sh[0]);
}
// one block folds the partial sums; thread 0 restores the plain residue (one product with R^2) and reports the flag
__global__ void __launch_bounds__(256) fr_dot_final_noalloc_kernel(unsigned char* scratch, int count, Fr* out, int* status_out) {
    __shared__ Fr sh[256];
    Fr acc = Fr::zero();
    for (int i = threadIdx.x; i < count; i += 256) acc = acc + fr_load(noalloc_partials(scratch) + i);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = sh[threadIdx.x] + sh[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        fr_store(out, sh[0] * Fr::r2());
        if (status_out) *status_out = *noalloc_flag(scratch) ? B2G16_E_NOT_CANONICAL : B2G16_OK;
    }
}

namespace {
// device address of the scratch array, per device ordinal; taken once, with that device current
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {};

unsigned char* noalloc_scratch_for(int device) {
    B2_REQUIRE(device >= 0 && device < 64, B2G16_E_INVALID_ARG, "no-alloc entry point: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        B2_CUDA(cudaSetDevice(device));
        unsigned char* p = nullptr;
        B2_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));     // NOT &g_noalloc_scratch: that is a host address
        B2_REQUIRE(p != nullptr, B2G16_E_CUDA, "no-alloc entry point: cudaGetSymbolAddress returned null");
        g_noalloc_ptr[device] = p;
    }
    return g_noalloc_ptr[device];
}
}  // namespace
}  // namespace b2

using namespace b2;

extern "C" int b2g16_fr_dot_device_noalloc(b2g16_ctx* h, const uint8_t* d_a32, const uint8_t* d_b32, size_t n, uint8_t* d_out32, int* d_status) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_out32 || (n && (!d_a32 || !d_b32))) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        B2_REQUIRE(n <= (size_t)B2G16_NOALLOC_MAX_N, B2G16_E_INVALID_ARG,
                   "b2g16_fr_dot_device_noalloc: n exceeds B2G16_NOALLOC_MAX_N (2^24 scalars); use b2g16_fr_dot_device for longer inputs");
        unsigned char* scratch = noalloc_scratch_for(c->device);
        const int blocks = (int)((n + NOALLOC_PER_BLOCK - 1) / NOALLOC_PER_BLOCK);
        B2_CUDA(cudaMemsetAsync(scratch, 0, NOALLOC_HEADER_BYTES, c->stream));              // the flag must start at zero
        if (blocks) {
            fr_dot_partial_noalloc_kernel<<<blocks, 256, 0, c->stream>>>(reinterpret_cast<const Fr*>(d_a32), reinterpret_cast<const Fr*>(d_b32), n, scratch);
            c->launches++;
        }
        fr_dot_final_noalloc_kernel<<<1, 256, 0, c->stream>>>(scratch, blocks, reinterpret_cast<Fr*>(d_out32), d_status);
        c->launches++;
        B2_CUDA(cudaGetLastError());
    });
}
