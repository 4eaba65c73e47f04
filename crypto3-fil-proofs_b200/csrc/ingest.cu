// Ingest / egress kernels: ABI byte encodings (include/b2g16.h) <-> Montgomery limbs, the
// fr32 canonical-scalar contract of the reference
// (libs/storage/include/nil/filecoin/storage/proofs/core/fr32.hpp:36-52: 32 LE bytes that
// must be < r), elementwise field ops for parity tests, and the MAD32 peak probe that is
// the MSM roofline denominator.  All HBM-streaming, one value per thread.
#include "common.cuh"
#include "ntt_kernels.cuh"

namespace b2 {

template <class F> B2_D F load_limbs(const uint8_t* p) {
    F r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
#pragma unroll
    for (int i = 0; i < F::N / 4; ++i) { uint4 t = q[i]; r.v[4 * i] = t.x; r.v[4 * i + 1] = t.y; r.v[4 * i + 2] = t.z; r.v[4 * i + 3] = t.w; }
    return r;
}
template <class F> B2_D void store_limbs(uint8_t* p, const F& f) {
    uint4* q = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < F::N / 4; ++i) q[i] = make_uint4(f.v[4 * i], f.v[4 * i + 1], f.v[4 * i + 2], f.v[4 * i + 3]);
}
// canonical bytes -> Montgomery; flags non-canonical input
template <class F> B2_D F ingest(const uint8_t* p, int& bad) {
    F c = load_limbs<F>(p);
    if (F::geq_p(c.v)) bad |= 1;
    return c * F::r2();
}
template <class F> B2_D void egress(uint8_t* p, const F& m) {
    F c; m.to_canonical(c.v);
    store_limbs<F>(p, c);
}

__global__ void fr_from_bytes_kernel(const uint8_t* in, Fr* out, size_t n, int* d_bad) {
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        fr_store(out + i, ingest<Fr>(in + 32 * i, bad));
    if (bad && d_bad) atomicOr(d_bad, bad);
}
__global__ void fr_to_bytes_kernel(const Fr* in, uint8_t* out, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        egress<Fr>(out + 32 * i, fr_load(in + i));
}
__global__ void scalars_check_kernel(const uint8_t* in, size_t n, int* d_bad) {
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        Fr c = load_limbs<Fr>(in + 32 * i);
        if (Fr::geq_p(c.v)) bad = 1;
    }
    if (bad) atomicOr(d_bad, 1);
}

B2_D Fq fq_small(uint32_t k) { Fq r = Fq::zero(); Fq one = Fq::one(); for (uint32_t i = 0; i < k; ++i) r = r + one; return r; }
B2_D bool on_curve(const G1Affine& p) { return p.is_identity() || p.y.sqr() == p.x.sqr() * p.x + fq_small(4); }
B2_D bool on_curve(const G2Affine& p) {
    Fq four = fq_small(4);
    Fq2 b{four, four};                                  // 4 (1 + u)
    return p.is_identity() || p.y.sqr() == p.x.sqr() * p.x + b;
}
B2_D void ingest_point(const uint8_t* p, G1Affine& o, int& bad) { o.x = ingest<Fq>(p, bad); o.y = ingest<Fq>(p + 48, bad); }
B2_D void ingest_point(const uint8_t* p, G2Affine& o, int& bad) {
    o.x.c0 = ingest<Fq>(p, bad); o.x.c1 = ingest<Fq>(p + 48, bad); o.y.c0 = ingest<Fq>(p + 96, bad); o.y.c1 = ingest<Fq>(p + 144, bad);
}
B2_D void egress_point(uint8_t* p, const G1Affine& a) { egress<Fq>(p, a.x); egress<Fq>(p + 48, a.y); }
B2_D void egress_point(uint8_t* p, const G2Affine& a) {
    egress<Fq>(p, a.x.c0); egress<Fq>(p + 48, a.x.c1); egress<Fq>(p + 96, a.y.c0); egress<Fq>(p + 144, a.y.c1);
}

template <class A, int BYTES> __global__ void __launch_bounds__(128) points_from_bytes_kernel(const uint8_t* in, A* out, size_t n, int* d_bad, int check_curve) {
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        A p; ingest_point(in + (size_t)BYTES * i, p, bad);
        if (check_curve && !on_curve(p)) bad |= 2;
        out[i] = p;
    }
    if (bad && d_bad) atomicOr(d_bad, bad);
}
template <class A, int BYTES> __global__ void __launch_bounds__(128) points_to_bytes_kernel(const A* in, uint8_t* out, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        egress_point(out + (size_t)BYTES * i, in[i]);
}

template <class F> __global__ void __launch_bounds__(128) field_op_kernel(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out, int* d_bad) {
    constexpr int B = 4 * F::N;
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        F x = ingest<F>(a + B * i, bad), y = ingest<F>(b + B * i, bad), r;
        switch (op) {
            case 0: r = x + y; break;
            case 1: r = x - y; break;
            case 2: r = x * y; break;
            case 3: r = x.inv(); break;
            case 4: r = x.sqr(); break;
            default: r = x.neg();
        }
        egress<F>(out + B * i, r);
    }
    if (bad && d_bad) atomicOr(d_bad, bad);
}

// out = sum_i a_i b_i mod r over canonical scalars (measurement aid: the scalar side of the multiexp linearity identity
// sum_i s_i (k_i G) = (sum_i s_i k_i) G that bench.py checks at 2^26 points).  A Montgomery product of two canonical
// operands is a b / R; the partial sums stay in that form and one product with R^2 at the end restores the plain residue.
__global__ void __launch_bounds__(256) fr_dot_partial_kernel(const uint8_t* a, const uint8_t* b, size_t n, Fr* partial, int* d_bad) {
    __shared__ Fr sh[256];
    Fr acc = Fr::zero();
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        Fr x = load_limbs<Fr>(a + 32 * i), y = load_limbs<Fr>(b + 32 * i);
        if (Fr::geq_p(x.v) || Fr::geq_p(y.v)) bad = 1;
        acc = acc + x * y;
    }
    if (bad && d_bad) atomicOr(d_bad, 1);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = sh[threadIdx.x] + sh[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) fr_store(partial + blockIdx.x, sh[0]);
}
__global__ void __launch_bounds__(32) fr_dot_final_kernel(const Fr* partial, int count, uint8_t* out32) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    Fr acc = Fr::zero();
    for (int i = 0; i < count; ++i) acc = acc + fr_load(partial + i);
    store_limbs<Fr>(out32, acc * Fr::r2());
}

// Sustained 32x32 -> 64 multiply rate of the integer pipe: 8 chains of IMAD.WIDE per thread whose
// multiplicands are live accumulator halves, so ptxas cannot hoist the product out of the loop
// (a loop-invariant a*b is folded into IADD3 pairs and would time the ALU pipe instead).
// Measured on B200: IMAD.WIDE (with or without .X carry) issues at 32 lanes/clk/SM -- half the
// 32-bit IMAD rate -- i.e. ~9.2e12 full products/s; this is the MSM/NTT roofline denominator.
__global__ void __launch_bounds__(256) mad32_probe_kernel(uint64_t* sink, uint32_t seed, int iters) {
    uint32_t a = seed ^ (threadIdx.x * 2654435761u);
    uint64_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = a + k;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc[k]) : "r"((uint32_t)(acc[k] >> 32)), "r"((uint32_t)acc[(k + 3) & 7]));
        }
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s ^= acc[k];
    if (s == 0x1234567u) sink[0] = s;                      // keep the chains alive
}

static void check_launch(Ctx* c) { c->launches++; B2_CUDA(cudaGetLastError()); }

void fr_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, Fr* d_out, size_t n, int* d_bad) {
    if (!n) return;
    fr_from_bytes_kernel<<<grid_for(n, 256), 256, 0, c->stream>>>(d_bytes, d_out, n, d_bad); check_launch(c);
}
void fr_to_bytes_dev(Ctx* c, const Fr* d_in, uint8_t* d_bytes, size_t n) {
    if (!n) return;
    fr_to_bytes_kernel<<<grid_for(n, 256), 256, 0, c->stream>>>(d_in, d_bytes, n); check_launch(c);
}
void scalars_check_dev(Ctx* c, const uint8_t* d_bytes, size_t n, int* d_bad) {
    if (!n) return;
    scalars_check_kernel<<<grid_for(n, 256), 256, 0, c->stream>>>(d_bytes, n, d_bad); check_launch(c);
}
void g1_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, G1Affine* d_out, size_t n, int* d_bad, bool check_curve) {
    if (!n) return;
    points_from_bytes_kernel<G1Affine, 96><<<grid_for(n, 128), 128, 0, c->stream>>>(d_bytes, d_out, n, d_bad, check_curve); check_launch(c);
}
void g2_from_bytes_dev(Ctx* c, const uint8_t* d_bytes, G2Affine* d_out, size_t n, int* d_bad, bool check_curve) {
    if (!n) return;
    points_from_bytes_kernel<G2Affine, 192><<<grid_for(n, 128), 128, 0, c->stream>>>(d_bytes, d_out, n, d_bad, check_curve); check_launch(c);
}
void g1_to_bytes_dev(Ctx* c, const G1Affine* d_in, uint8_t* d_bytes, size_t n, cudaStream_t st) {
    if (!n) return;
    points_to_bytes_kernel<G1Affine, 96><<<grid_for(n, 128), 128, 0, st>>>(d_in, d_bytes, n); check_launch(c);
}
void g2_to_bytes_dev(Ctx* c, const G2Affine* d_in, uint8_t* d_bytes, size_t n, cudaStream_t st) {
    if (!n) return;
    points_to_bytes_kernel<G2Affine, 192><<<grid_for(n, 128), 128, 0, st>>>(d_in, d_bytes, n); check_launch(c);
}
void field_op_dev(Ctx* c, int field, int op, const uint8_t* d_a, const uint8_t* d_b, size_t n, uint8_t* d_out, int* d_bad) {
    if (!n) return;
    if (field == 0) field_op_kernel<Fr><<<grid_for(n, 128), 128, 0, c->stream>>>(op, d_a, d_b, n, d_out, d_bad);
    else field_op_kernel<Fq><<<grid_for(n, 128), 128, 0, c->stream>>>(op, d_a, d_b, n, d_out, d_bad);
    check_launch(c);
}

void fr_dot_dev(Ctx* c, const uint8_t* d_a, const uint8_t* d_b, size_t n, uint8_t* d_out32, int* d_bad) {
    const int blocks = grid_for(n, 256, 148 * 8);
    DevBuf<Fr> partial(blocks, c->stream);
    fr_dot_partial_kernel<<<blocks, 256, 0, c->stream>>>(d_a, d_b, n, partial.get(), d_bad); check_launch(c);
    fr_dot_final_kernel<<<1, 32, 0, c->stream>>>(partial.get(), blocks, d_out32); check_launch(c);
}

double mad32_peak_dev(Ctx* c, double* elapsed_ms) {
    DevBuf<uint64_t> sink(1, c->stream);
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0)); B2_CUDA(cudaEventCreate(&e1));
    mad32_probe_kernel<<<blocks, threads, 0, c->stream>>>(sink.get(), 1, 64); check_launch(c);      // warm-up
    B2_CUDA(cudaEventRecord(e0, c->stream));
    mad32_probe_kernel<<<blocks, threads, 0, c->stream>>>(sink.get(), 2, iters); check_launch(c);
    B2_CUDA(cudaEventRecord(e1, c->stream));
    B2_CUDA(cudaEventSynchronize(e1));
    float ms = 0; B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (elapsed_ms) *elapsed_ms = ms;
    return (double)blocks * threads * iters * 64.0 / (ms * 1e-3);
}

}  // namespace b2
