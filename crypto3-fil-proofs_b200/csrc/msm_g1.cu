// G1 MSM instantiation, generator multiples and point sums (see msm_impl.cuh).
#include "msm_impl.cuh"

namespace b2 {

// standard BLS12-381 G1 generator, canonical limbs (on-curve asserted by the oracle tests)
static const uint32_t G1_GEN_X[12] = {0xdb22c6bbu, 0xfb3af00au, 0xf97a1aefu, 0x6c55e83fu, 0x171bac58u, 0xa14e3a3fu, 0x9774b905u, 0xc3688c4fu, 0x4fa9ac0fu, 0x2695638cu, 0x3197d794u, 0x17f1d3a7u};
static const uint32_t G1_GEN_Y[12] = {0x46c5e7e1u, 0x0caa2329u, 0xa2888ae4u, 0xd03cc744u, 0x2c04b3edu, 0x00db18cbu, 0xd5d00af6u, 0xfcf5e095u, 0x741d8ae4u, 0xa09e30edu, 0xe3aaa0f1u, 0x08b3f481u};

void msm_g1_dev(Ctx* c, const G1Affine* d_bases, const uint32_t* d_scalars, size_t n, uint8_t* d_out_bytes, cudaStream_t st, int window_bits) {
    msm_run<Fq>(c, d_bases, d_scalars, n, d_out_bytes, st, window_bits);
}
void msm_g1_windows_dev(Ctx* c, const G1Affine* d_bases, const uint32_t* d_scalars, size_t n, G1Xyzz* d_win /*32*/, int* c_bits, int* n_windows, cudaStream_t st) {
    MsmPlan plan{0, 0};
    msm_run<Fq>(c, d_bases, d_scalars, n, nullptr, st, 0, d_win, &plan);
    *c_bits = plan.c; *n_windows = plan.W;
}
void gen_mul_g1_dev(Ctx* c, const uint32_t* d_scalars, size_t n, G1Affine* d_out) {
    if (!n) return;
    G1Affine gen{Fq::from_canonical(G1_GEN_X), Fq::from_canonical(G1_GEN_Y)};
    G1Affine* table = fixed_base_table<Fq>(c, 0, gen);
    gen_mul_kernel<Fq><<<grid_for(n, 128), 128, 0, c->stream>>>(table, d_scalars, n, d_out);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}
void g1_subgroup_check_dev(Ctx* c, const G1Affine* d_pts, size_t n, int* d_bad) {
    if (!n) return;
    subgroup_check_kernel<Fq><<<grid_for(n, 64), 64, 0, c->stream>>>(d_pts, n, d_bad);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}
void sum_g1_dev(Ctx* c, const G1Affine* d_pts, size_t n, uint8_t* d_out_bytes) {
    sum_points_kernel<Fq><<<1, 32, 0, c->stream>>>(d_pts, n, d_out_bytes);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}

}  // namespace b2
