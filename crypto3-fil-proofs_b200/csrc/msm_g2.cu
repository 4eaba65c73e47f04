// G2 MSM instantiation, generator multiples and point sums (see msm_impl.cuh).
#include "msm_impl.cuh"

namespace b2 {

// standard BLS12-381 G2 generator, canonical limbs: x.c0, x.c1, y.c0, y.c1
static const uint32_t G2_GEN[4][12] = {
    {0xc121bdb8u, 0xd48056c8u, 0xa805bbefu, 0x0bac0326u, 0x7ae3d177u, 0xb4510b64u, 0xfa403b02u, 0xc6e47ad4u, 0x2dc51051u, 0x26080527u, 0xf08f0a91u, 0x024aa2b2u},
    {0x5d042b7eu, 0xe5ac7d05u, 0x13945d57u, 0x334cf112u, 0xdc7f5049u, 0xb5da61bbu, 0x9920b61au, 0x596bd0d0u, 0x88274f65u, 0x7dacd3a0u, 0x52719f60u, 0x13e02b60u},
    {0x08b82801u, 0xe1935486u, 0x3baca289u, 0x923ac9ccu, 0x5160d12cu, 0x6d429a69u, 0x8cbdd3a7u, 0xadfd9baau, 0xda2e351au, 0x8cc9cdc6u, 0x727d6e11u, 0x0ce5d527u},
    {0xf05f79beu, 0xaaa9075fu, 0x5cec1da1u, 0x3f370d27u, 0x572e99abu, 0x267492abu, 0x85a763afu, 0xcb3e287eu, 0x2bc28b99u, 0x32acd2b0u, 0x2ea734ccu, 0x0606c4a0u}};

void msm_g2_dev(Ctx* c, const G2Affine* d_bases, const uint32_t* d_scalars, size_t n, uint8_t* d_out_bytes, cudaStream_t st, int window_bits) {
    msm_run<Fq2>(c, d_bases, d_scalars, n, d_out_bytes, st, window_bits);
}
void msm_g2_windows_dev(Ctx* c, const G2Affine* d_bases, const uint32_t* d_scalars, size_t n, G2Xyzz* d_win /*32*/, int* c_bits, int* n_windows, cudaStream_t st) {
    MsmPlan plan{0, 0};
    msm_run<Fq2>(c, d_bases, d_scalars, n, nullptr, st, 0, d_win, &plan);
    *c_bits = plan.c; *n_windows = plan.W;
}
void gen_mul_g2_dev(Ctx* c, const uint32_t* d_scalars, size_t n, G2Affine* d_out) {
    if (!n) return;
    G2Affine gen{{Fq::from_canonical(G2_GEN[0]), Fq::from_canonical(G2_GEN[1])}, {Fq::from_canonical(G2_GEN[2]), Fq::from_canonical(G2_GEN[3])}};
    G2Affine* table = fixed_base_table<Fq2>(c, 1, gen);
    gen_mul_kernel<Fq2><<<grid_for(n, 128), 128, 0, c->stream>>>(table, d_scalars, n, d_out);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}
void sum_g2_dev(Ctx* c, const G2Affine* d_pts, size_t n, uint8_t* d_out_bytes) {
    sum_points_kernel<Fq2><<<1, 32, 0, c->stream>>>(d_pts, n, d_out_bytes);
    c->launches++;
    B2_CUDA(cudaGetLastError());
}

}  // namespace b2
