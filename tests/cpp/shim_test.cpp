// Exercises the crypto3-shaped C++ shim (include/nil/crypto3/b2g16_shim.hpp) the way
// compound_proof::circuit_proofs would: multiexp over G1/G2, evaluation_domain transforms and
// prove<r1cs_gg_ppzksnark<bls12<381>>>.  Inputs and the oracle's expected outputs are raw .bin files
// written by tests/test_cpp_shim.py into the directory given as argv[1].
//   shim_test <dir>      -> runs on the GPU, exit 0 iff every result is byte-identical
//   shim_test --no-gpu   -> checks that, without a device, the shim throws (no CPU fallback)
#include <cstdio>
#include <fstream>
#include <iostream>
#include <iterator>
#include "nil/crypto3/b2g16_shim.hpp"

using namespace nil::crypto3;
typedef algebra::curves::bls12<381> curve;
typedef zk::snark::r1cs_gg_ppzksnark<curve> scheme;
typedef scheme::fr_value fr;
typedef scheme::g1_value g1;
typedef scheme::g2_value g2;

static std::string dir;
template <class T> std::vector<T> load(const char* name) {
    std::ifstream f(dir + "/" + name + ".bin", std::ios::binary);
    if (!f) { std::cerr << "missing " << name << "\n"; exit(2); }
    std::vector<char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    std::vector<T> v(raw.size() / sizeof(T));
    if (!v.empty()) memcpy((void*)v.data(), raw.data(), v.size() * sizeof(T));
    return v;
}
static int fails = 0;
static void expect(bool ok, const char* what) { printf("%s %s\n", ok ? "ok  " : "FAIL", what); if (!ok) ++fails; }

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    if (std::string(argv[1]) == "--no-gpu") {
        try { std::vector<g1> b(1); std::vector<fr> s(1); algebra::multiexp(b.begin(), b.end(), s.begin(), s.end(), 1); }
        catch (const b2g16::error& e) { printf("ok   no device -> %s\n", e.what()); return e.code == B2G16_E_CUDA ? 0 : 1; }
        return 1;
    }
    dir = argv[1];
    static_assert(sizeof(fr) == 32 && sizeof(g1) == 96 && sizeof(g2) == 192, "value types are the ABI encodings");
    {   // multiexp (a14)
        auto b1 = load<g1>("msm_g1_bases"); auto b2 = load<g2>("msm_g2_bases"); auto s = load<fr>("msm_scalars");
        g1 r1 = algebra::multiexp<algebra::policies::multiexp_method_BDLO12>(b1.begin(), b1.end(), s.begin(), s.end(), 8);
        expect(r1 == load<g1>("msm_g1_out")[0], "multiexp<BDLO12> over G1");
        g2 r2 = algebra::multiexp_with_mixed_addition(b2.begin(), b2.end(), s.begin(), s.begin() + b2.size(), 8);
        expect(r2 == load<g2>("msm_g2_out")[0], "multiexp_with_mixed_addition over G2");
        bool threw = false;
        try { algebra::multiexp(b1.begin(), b1.end(), s.begin(), s.begin() + 1, 1); } catch (const b2g16::error&) { threw = true; }
        expect(threw, "length mismatch is rejected");
        // value arithmetic on the point / scalar types (group law through the library): P + Q, k * P, -P, scalar field ops
        fr one{}; one.data[0] = 1;
        fr two = one + one, three = two + one;
        std::vector<g1> pq = {b1[0], b1[1]}; std::vector<fr> k11 = {one, one}, k23 = {two, three};
        expect(b1[0] + b1[1] == algebra::multiexp(pq.begin(), pq.end(), k11.begin(), k11.end()), "P + Q on g1 values");
        expect(two * b1[0] + three * b1[1] == algebra::multiexp(pq.begin(), pq.end(), k23.begin(), k23.end()), "k * P on g1 values");
        expect((b1[0] + (-b1[0])).is_zero() && b1[0] - b1[1] + b1[1] == b1[0], "-P and P - Q on g1 values");
        expect(b2[0] + b2[0] == two * b2[0] && (b2[0] - b2[0]).is_zero(), "P + P, k * P, P - P on g2 values");
        expect(three * inversed(three) == one && two - three + one == fr{} && -one + one == fr{}, "scalar field operators");
    }
    {   // evaluation_domain (a13)
        auto x = load<fr>("ntt_in");
        auto dom = math::make_evaluation_domain<curve::scalar_field_type>(x.size());
        auto y = x; dom->fft(y);
        expect(y == load<fr>("ntt_fwd"), "evaluation_domain::fft");
        dom->inverse_fft(y);
        expect(y == x, "inverse_fft(fft(x)) == x");
        y = x; dom->icosetFFT(y);
        expect(y == load<fr>("ntt_icoset"), "icosetFFT");
        bool threw = false;
        try { math::evaluation_domain<curve::scalar_field_type> bad(12); } catch (const b2g16::error&) { threw = true; }
        expect(threw, "non power-of-two domain is rejected");
    }
    {   // generate + prove (a11), explicit r, s
        auto shape = load<std::uint64_t>("cs_shape");            // log_m, n_cons, n_in, n_aux
        scheme::constraint_system_type cs;
        cs.log_m = (std::uint32_t)shape[0]; cs.n_cons = shape[1]; cs.n_in = shape[2]; cs.n_aux = shape[3];
        const char* names[3][3] = {{"A_rp", "A_col", "A_val"}, {"B_rp", "B_col", "B_val"}, {"C_rp", "C_col", "C_val"}};
        scheme::constraint_system_type::matrix* m[3] = {&cs.A, &cs.B, &cs.C};
        for (int i = 0; i < 3; ++i) { m[i]->row_ptr = load<std::uint32_t>(names[i][0]); m[i]->col = load<std::uint32_t>(names[i][1]); m[i]->val = load<fr>(names[i][2]); }
        auto tox = load<fr>("toxic"), z = load<fr>("z"), rs = load<fr>("rs");
        std::array<fr, 5> toxic; std::copy(tox.begin(), tox.begin() + 5, toxic.begin());
        auto kp = zk::snark::generate<scheme>(cs, toxic);
        expect(kp.second.gamma_ABC_g1 == load<g1>("ic"), "generate: verification key IC");
        scheme::primary_input_type primary(z.begin() + 1, z.begin() + 1 + cs.n_in);
        scheme::auxiliary_input_type aux(z.begin() + 1 + cs.n_in, z.end());
        auto proof = zk::snark::prove<scheme>(kp.first, primary, aux, rs[0], rs[1]);
        auto want = load<std::uint8_t>("proof");
        expect(want.size() == 192 && memcmp(proof.bytes().data(), want.data(), 192) == 0, "prove: 192 proof bytes");
        expect(proof.g_A == load<g1>("proof_A")[0] && proof.g_B == load<g2>("proof_B")[0] && proof.g_C == load<g1>("proof_C")[0], "prove: g_A, g_B, g_C");
        auto p2 = zk::snark::prove<scheme>(kp.first, primary, aux);                          // random r, s: valid, different bytes
        expect(memcmp(p2.bytes().data(), want.data(), 192) != 0, "prove with drawn r, s differs");
        // the seal_commit_phase2 tail (F/api/seal.hpp:304-313): write the multi-proof bytes, read them back, verify
        scheme::processed_verification_key_type pvk(kp.second);
        expect(zk::snark::verify<scheme>(pvk, primary, proof), "verify: oracle-identical proof accepted");
        expect(zk::snark::verify<scheme>(pvk, primary, p2), "verify: proof with drawn r, s accepted");
        expect(zk::snark::verify<scheme>(kp.second, primary, proof), "verify: from the plain verification key");
        std::uint8_t buf[384];
        proof.write(buf); p2.write(buf + 192);
        auto q = scheme::proof_type::read(buf);
        expect(q.g_A == proof.g_A && q.g_B == proof.g_B && q.g_C == proof.g_C, "proof read back: same points");
        expect(zk::snark::verify_batch<scheme>(pvk, {primary, primary}, {q, scheme::proof_type::read(buf + 192)}), "verify_batch: two partitions");
        auto bad_in = primary; bad_in[0].data[0] ^= 1;
        expect(!zk::snark::verify<scheme>(pvk, bad_in, proof), "verify: perturbed public input rejected");
        expect(!zk::snark::verify_batch<scheme>(pvk, {primary, bad_in}, {proof, p2}), "verify_batch: one bad input sinks the batch");
        buf[0] &= 0x7f;
        bool threw = false;
        try { scheme::proof_type::read(buf); } catch (const b2g16::error&) { threw = true; }
        expect(threw, "proof read: malformed encoding throws");
        // the parameter cache (parameter_cache.hpp:125-153): key -> .params file (bellman-style filtered, and dense) -> key
        for (int filtered = 1; filtered >= 0; --filtered) {
            std::string path = dir + (filtered ? "/v28-shim-filtered.params" : "/v28-shim-dense.params");
            kp.first.write_cached(path, kp.second, filtered != 0);
            scheme::verification_key_type vk2;
            auto pk2 = scheme::proving_key_type::read_cached(cs, path, &vk2);
            auto p3 = zk::snark::prove<scheme>(pk2, primary, aux, rs[0], rs[1]);
            expect(memcmp(p3.bytes().data(), want.data(), 192) == 0, filtered ? "key from a density-filtered .params file: same proof" : "key from a dense .params file: same proof");
            expect(vk2.gamma_ABC_g1 == kp.second.gamma_ABC_g1 && vk2.gamma_g2 == kp.second.gamma_g2 && vk2.alpha_g1 == kp.second.alpha_g1, "vk carried by the .params file");
        }
        {
            auto other = cs; other.n_aux -= 1;
            threw = false;
            try { scheme::proving_key_type::read_cached(other, dir + "/v28-shim-filtered.params"); } catch (const b2g16::error&) { threw = true; }
            expect(threw, "parameters of another circuit are refused");
            FILE* a = fopen((dir + "/v28-shim-filtered.params").c_str(), "rb"); FILE* b = fopen((dir + "/v28-shim-dense.params").c_str(), "rb");
            fseek(a, 0, SEEK_END); fseek(b, 0, SEEK_END);
            expect(ftell(a) < ftell(b), "filtered file is smaller than the dense one");
            fclose(a); fclose(b);
        }
    }
    printf("%s\n", fails ? "SHIM TEST FAILED" : "shim test passed");
    return fails ? 1 : 0;
}
