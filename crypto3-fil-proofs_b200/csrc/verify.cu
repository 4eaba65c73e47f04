// Groth16 verifier and the 192-byte proof codec behind the C ABI.
//
// Replaces what the reference reaches through verify_seal / verify_batch_seal
// (libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:339-395, :409-485) ->
// compound_proof::verify (libs/storage/.../core/proof/compound_proof.hpp:97-102, a stub) ->
// crypto3 zk `verify<r1cs_gg_ppzksnark>` with the processed verification key that
// mapped_scheme_params keeps next to the proving key (.../core/crypto/mapped_scheme_params.hpp:61).
// seal_commit_phase2 refuses to return a proof that does not verify (seal.hpp:310-313), so the
// prover's caller needs this on the same boundary.
//
// Split of work: the public-input combination  IC_0 + sum_j x_j IC_j  is a multiexp and runs on
// the GPU over resident IC bases (msm_g1_dev); the three/four pairings of one check are O(1)
// and run on the host (host_bls.hpp).
#include <memory>
#include <random>

#include "api_util.cuh"
#include "host_bls.hpp"

using namespace b2;

struct b2g16_pvk {
    size_t n_ic = 0;
    b2h::G1 alpha_g1;
    b2h::G2 beta_g2, gamma_g2, delta_g2;
    b2h::G2 neg_gamma_g2, neg_delta_g2;
    b2h::Fq12 miller_alpha_beta;           // f_{|x|,beta}(alpha), not yet exponentiated
    G1Affine* d_ic = nullptr;              // resident Montgomery-affine IC bases
};

namespace {

struct DecodedProof { b2h::G1 a, c; b2h::G2 b; };

// decompress + on-curve (by construction) + prime-order subgroup membership, as bellman's
// Proof::read does for every proof it deserialises; like Proof::read, the point at infinity is refused for A, B
// and C (miller_loop would silently skip such a term)
bool decode_proof(const uint8_t* p, DecodedProof* o) {
    if (!b2h::g1_decompress(p, &o->a) || !b2h::g2_decompress(p + 48, &o->b) || !b2h::g1_decompress(p + 144, &o->c)) return false;
    if (o->a.inf || o->b.inf || o->c.inf) return false;
    return b2h::in_subgroup(o->a) && b2h::in_subgroup(o->b) && b2h::in_subgroup(o->c);
}

// X = IC_0 + sum_j x_j IC_j on the device; inputs32 = n_in canonical scalars (without the leading ONE)
b2h::G1 input_combination(Ctx* c, const b2g16_pvk* vk, const uint8_t* inputs32, size_t n_in) {
    const size_t n = n_in + 1;
    std::vector<uint8_t> s(n * 32, 0);
    s[0] = 1;
    if (n_in) memcpy(s.data() + 32, inputs32, n_in * 32);
    DevBuf<uint8_t> d_s(n * 32, c->stream), d_out(96, c->stream);
    DevFlag bad(c);
    B2_CUDA(cudaMemcpyAsync(d_s.get(), s.data(), n * 32, cudaMemcpyHostToDevice, c->stream));
    scalars_check_dev(c, d_s.get(), n, bad.ptr());
    msm_g1_dev(c, vk->d_ic, reinterpret_cast<const uint32_t*>(d_s.get()), n, d_out.get(), c->stream, 0);
    uint8_t out[96];
    B2_CUDA(cudaMemcpyAsync(out, d_out.get(), 96, cudaMemcpyDeviceToHost, c->stream));
    bad.check("public inputs");
    b2h::G1 x;
    B2_REQUIRE(b2h::g1_from_abi(out, &x), B2G16_E_INTERNAL, "verify: input combination not canonical");
    return x;
}

}  // namespace

extern "C" {

int b2g16_vk_prepare(b2g16_ctx* h, const uint8_t* vk, size_t n_ic, b2g16_pvk** out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !vk || !out || n_ic == 0) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        std::unique_ptr<b2g16_pvk> p(new b2g16_pvk());
        p->n_ic = n_ic;
        B2_REQUIRE(b2h::g1_from_abi(vk, &p->alpha_g1) && b2h::g2_from_abi(vk + 96, &p->beta_g2) && b2h::g2_from_abi(vk + 288, &p->gamma_g2) &&
                   b2h::g2_from_abi(vk + 480, &p->delta_g2), B2G16_E_NOT_CANONICAL, "vk: coordinate not canonical");
        B2_REQUIRE(b2h::on_curve(p->alpha_g1) && b2h::on_curve(p->beta_g2) && b2h::on_curve(p->gamma_g2) && b2h::on_curve(p->delta_g2),
                   B2G16_E_NOT_ON_CURVE, "vk: point not on curve");
        B2_REQUIRE(b2h::in_subgroup(p->alpha_g1) && b2h::in_subgroup(p->beta_g2) && b2h::in_subgroup(p->gamma_g2) && b2h::in_subgroup(p->delta_g2),
                   B2G16_E_NOT_ON_CURVE, "vk: point not in the prime-order subgroup");
        p->neg_gamma_g2 = p->gamma_g2.neg();
        p->neg_delta_g2 = p->delta_g2.neg();
        B2_REQUIRE(b2h::miller_loop({{p->alpha_g1, p->beta_g2}}, &p->miller_alpha_beta), B2G16_E_INVALID_ARG, "vk: degenerate alpha/beta");
        DevFlag bad(c);
        B2_CUDA(cudaMalloc((void**)&p->d_ic, n_ic * sizeof(G1Affine)));
        B2_CUDA(cudaMemcpyAsync(p->d_ic, vk + 672, n_ic * 96, cudaMemcpyHostToDevice, c->stream));
        g1_from_bytes_dev(c, reinterpret_cast<const uint8_t*>(p->d_ic), p->d_ic, n_ic, bad.ptr(), true);
        g1_subgroup_check_dev(c, p->d_ic, n_ic, bad.ptr());          // bellman's VerifyingKey::read checks the IC points too
        try { bad.check("vk ic"); } catch (...) { cudaFree(p->d_ic); throw; }
        *out = p.release();
    });
}

int b2g16_pvk_free(b2g16_ctx* h, b2g16_pvk* vk) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c) return B2G16_E_INVALID_ARG;
    if (!vk) return B2G16_OK;
    return guard(c, [&] { B2_CUDA(cudaStreamSynchronize(c->stream)); cudaFree(vk->d_ic); delete vk; });
}

// e(A, B) == e(alpha, beta) . e(X, gamma) . e(C, delta)   (SURVEY App. B step 7), checked as
// e(-A, B) . e(alpha, beta) . e(X, gamma) . e(C, delta) == 1 with the alpha/beta Miller value cached.
int b2g16_verify(b2g16_ctx* h, const b2g16_pvk* vk, const uint8_t* inputs32, size_t n_in, const uint8_t proof[192], int* ok) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !vk || !proof || !ok || (n_in && !inputs32)) return B2G16_E_INVALID_ARG;
    *ok = 0;
    return guard(c, [&] {
        B2_REQUIRE(n_in + 1 == vk->n_ic, B2G16_E_INVALID_ARG, "verify: public input count does not match the key");
        DecodedProof p;
        if (!decode_proof(proof, &p)) return;                         // malformed proof: ok = 0, not an error
        const b2h::G1 x = input_combination(c, vk, inputs32, n_in);
        b2h::Fq12 f;
        if (!b2h::miller_loop({{p.a.neg(), p.b}, {x, vk->gamma_g2}, {p.c, vk->delta_g2}}, &f)) return;
        *ok = b2h::final_exponentiation(f * vk->miller_alpha_beta) == b2h::Fq12::one() ? 1 : 0;
    });
}

// Random linear combination of n_proofs checks into one product of n_proofs + 3 Miller loops and a
// single final exponentiation (verify_batch_seal, seal.hpp:409-485).  rand32: n_proofs canonical
// scalars chosen by the caller, or NULL to draw 128-bit ones from std::random_device.
int b2g16_verify_batch(b2g16_ctx* h, const b2g16_pvk* vk, const uint8_t* inputs32, size_t n_in, const uint8_t* proofs, size_t n_proofs,
                       const uint8_t* rand32, int* ok) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !vk || !ok || (n_proofs && !proofs) || (n_proofs && n_in && !inputs32)) return B2G16_E_INVALID_ARG;
    *ok = 0;
    return guard(c, [&] {
        B2_REQUIRE(n_in + 1 == vk->n_ic, B2G16_E_INVALID_ARG, "verify: public input count does not match the key");
        if (n_proofs == 0) { *ok = 1; return; }
        std::vector<uint8_t> rho(n_proofs * 32, 0);
        if (rand32) {
            for (size_t i = 0; i < n_proofs; ++i) B2_REQUIRE(b2h::fr_canonical(rand32 + 32 * i), B2G16_E_NOT_CANONICAL, "verify: random scalar not canonical");
            memcpy(rho.data(), rand32, rho.size());
        } else {
            std::random_device rd;
            for (size_t i = 0; i < n_proofs; ++i) for (int k = 0; k < 4; ++k) { uint32_t w = rd(); memcpy(&rho[32 * i + 4 * k], &w, 4); }
        }
        std::vector<b2h::PairTerm> terms;
        b2h::Pt<b2h::Fq> sum_x = b2h::Pt<b2h::Fq>::identity(), sum_c = sum_x, sum_alpha = sum_x;
        for (size_t i = 0; i < n_proofs; ++i) {
            DecodedProof p;
            if (!decode_proof(proofs + 192 * i, &p)) return;
            const uint8_t* r = &rho[32 * i];
            const b2h::G1 x = input_combination(c, vk, n_in ? inputs32 + 32 * n_in * i : nullptr, n_in);
            terms.push_back({b2h::scalar_mul(p.a, r).affine(), p.b});
            sum_x = sum_x.add(b2h::scalar_mul(x, r));
            sum_c = sum_c.add(b2h::scalar_mul(p.c, r));
            sum_alpha = sum_alpha.add(b2h::scalar_mul(vk->alpha_g1, r));
        }
        terms.push_back({sum_alpha.affine().neg(), vk->beta_g2});
        terms.push_back({sum_x.affine(), vk->neg_gamma_g2});
        terms.push_back({sum_c.affine(), vk->neg_delta_g2});
        *ok = b2h::pairing_product_is_one(terms) ? 1 : 0;
    });
}

// 192-byte proof <-> ABI affine points.  decode validates encoding, curve and subgroup.
int b2g16_proof_decode(const uint8_t proof[192], uint8_t a96[96], uint8_t b192[192], uint8_t c96[96]) {
    if (!proof || !a96 || !b192 || !c96) return B2G16_E_INVALID_ARG;
    DecodedProof p;
    if (!decode_proof(proof, &p)) return B2G16_E_NOT_ON_CURVE;
    b2h::g1_to_abi(p.a, a96); b2h::g2_to_abi(p.b, b192); b2h::g1_to_abi(p.c, c96);
    return B2G16_OK;
}
int b2g16_proof_encode(const uint8_t a96[96], const uint8_t b192[192], const uint8_t c96[96], uint8_t proof[192]) {
    if (!proof || !a96 || !b192 || !c96) return B2G16_E_INVALID_ARG;
    DecodedProof p;
    if (!b2h::g1_from_abi(a96, &p.a) || !b2h::g2_from_abi(b192, &p.b) || !b2h::g1_from_abi(c96, &p.c)) return B2G16_E_NOT_CANONICAL;
    if (!b2h::on_curve(p.a) || !b2h::on_curve(p.b) || !b2h::on_curve(p.c)) return B2G16_E_NOT_ON_CURVE;
    b2h::g1_compress(p.a, proof); b2h::g2_compress(p.b, proof + 48); b2h::g1_compress(p.c, proof + 144);
    return B2G16_OK;
}

}  // extern "C"
