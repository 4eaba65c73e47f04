// The device-resident proving key (scheme_params<r1cs_gg_ppzksnark<bls12<381>>>,
// libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:38-67, dense layout) and the
// constraint matrices it is proved against, shared by prove.cu (prover, loaders from host memory) and
// pk_stream.cu (loader that streams a parameter file).
#pragma once
#include "common.cuh"

namespace b2 {

struct DevCsr { uint64_t n_rows = 0, nnz = 0; uint32_t* row_ptr = nullptr; uint32_t* col = nullptr; Fr* val = nullptr; };

}  // namespace b2

struct b2g16_pk {
    uint32_t log_m = 0;
    uint64_t n_cons = 0, n_in = 0, n_aux = 0, n_vars = 0, m = 0;
    b2::DevCsr A, B, C;
    b2::G1Affine *a = nullptr, *b_g1 = nullptr, *l = nullptr, *h = nullptr;
    b2::G2Affine* b_g2 = nullptr;
    b2::G1Affine alpha_g1, beta_g1, delta_g1;     // host copies, Montgomery
    b2::G2Affine beta_g2, delta_g2;
};

namespace b2 {
// prove.cu
b2g16_pk* new_pk_with_cs(Ctx* c, const b2g16_r1cs_view* cs, int* d_bad);     // shape + constraint matrices on the device
void free_pk(b2g16_pk* pk);
// the five single key elements from ABI bytes: validated on the device (range, curve), host copies stored in pk
void pk_set_single_points(Ctx* c, b2g16_pk* pk, const uint8_t* alpha_g1, const uint8_t* beta_g1, const uint8_t* delta_g1, const uint8_t* beta_g2,
                          const uint8_t* delta_g2, int* d_bad);
}  // namespace b2
