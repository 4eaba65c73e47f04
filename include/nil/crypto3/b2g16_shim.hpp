// Header-only C++ shim over the C ABI of libb2g16 (include/b2g16.h) under the crypto3 names the
// reference's callers use, so compound_proof::circuit_proofs
// (libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137) can be
// written once against either backend.  Names mirrored (the originals live in un-vendored
// submodules; the reference names them at the cited lines):
//   nil::crypto3::algebra::curves::bls12<381>, g1_type<>::value_type, scalar_field_type
//       (.../core/crypto/scheme_params.hpp:40-41, .../core/proof/compound_proof.hpp:130)
//   nil::crypto3::algebra::multiexp / multiexp_with_mixed_addition            
//   nil::crypto3::math::evaluation_domain / make_evaluation_domain            
//   nil::crypto3::zk::snark::r1cs_gg_ppzksnark<Curve>::{proof_type, proving_key_type, ...},
//   prove<Scheme>(pk, primary_input, auxiliary_input)
//       (.../core/proof/multi_proof.hpp:39-43, libs/filecoin/include/nil/filecoin/proofs/caches.hpp:41-43)
// Values are the ABI byte encodings (Fr 32 B LE canonical, affine points 96/192 B); every
// operation runs on the GPU -- there is no CPU fallback, a missing device throws b2g16::error.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../../b2g16.h"

namespace nil {
namespace crypto3 {
namespace b2g16 {

struct error : std::runtime_error {
    int code;
    error(int c, const std::string& m) : std::runtime_error("b2g16 error " + std::to_string(c) + ": " + m), code(c) {}
};

// One context per (process, device); created on first use.
class context {
    b2g16_ctx* h_ = nullptr;
public:
    explicit context(int device = 0) {
        int rc = b2g16_init(device, nullptr, &h_);
        if (rc != B2G16_OK) throw error(rc, "b2g16_init failed (no CUDA device / bad ordinal)");
    }
    ~context() { if (h_) b2g16_shutdown(h_); }
    context(const context&) = delete;
    context& operator=(const context&) = delete;
    b2g16_ctx* get() const { return h_; }
    void check(int rc) const { if (rc != B2G16_OK) throw error(rc, b2g16_last_error(h_)); }
    static context& instance(int device = 0) { static context c(device); return c; }
};

template <std::size_t N> struct bytes_value {
    std::array<std::uint8_t, N> data{};
    static constexpr std::size_t size = N;
    bool is_zero() const { for (auto b : data) if (b) return false; return true; }
    bool operator==(const bytes_value& o) const { return data == o.data; }
    bool operator!=(const bytes_value& o) const { return !(*this == o); }
    static bytes_value zero() { return bytes_value(); }
};


// Value arithmetic on the byte-encoded types, so that caller code written against crypto3's value types (P + Q, -P,
// k * P on g1_type<>::value_type / g2_type<>::value_type; a + b, a * b, a.inversed() on scalar values) compiles and
// gives canonical results.  Each operator is one library call on the device: meant for the O(1) glue arithmetic
// around a proof (multi_proof bookkeeping, test assertions), not for loops -- those are what multiexp is for.
namespace detail {
inline bytes_value<32> fr_op(int op, const bytes_value<32>& a, const bytes_value<32>* b) {
    bytes_value<32> r;
    context& c = context::instance();
    c.check(b2g16_field_op(c.get(), 0, op, a.data.data(), b ? b->data.data() : nullptr, 1, r.data.data()));
    return r;
}
// p - y on 48 little-endian bytes (y != 0, y < p)
inline void fq_negate(std::uint8_t* y) {
    static const std::uint8_t P[48] = {0xab, 0xaa, 0xff, 0xff, 0xff, 0xff, 0xfe, 0xb9, 0xff, 0xff, 0x53, 0xb1, 0xfe, 0xff, 0xab, 0x1e,
                                       0x24, 0xf6, 0xb0, 0xf6, 0xa0, 0xd2, 0x30, 0x67, 0xbf, 0x12, 0x85, 0xf3, 0x84, 0x4b, 0x77, 0x64,
                                       0xd7, 0xac, 0x4b, 0x43, 0xb6, 0xa7, 0x1b, 0x4b, 0x9a, 0xe6, 0x7f, 0x39, 0xea, 0x11, 0x01, 0x1a};
    bool zero = true;
    for (int i = 0; i < 48; ++i) zero = zero && y[i] == 0;
    if (zero) return;
    int borrow = 0;
    for (int i = 0; i < 48; ++i) { int v = (int)P[i] - (int)y[i] - borrow; borrow = v < 0; y[i] = (std::uint8_t)(v + (borrow << 8)); }
}
}  // namespace detail
inline bytes_value<32> operator+(const bytes_value<32>& a, const bytes_value<32>& b) { return detail::fr_op(0, a, &b); }
inline bytes_value<32> operator-(const bytes_value<32>& a, const bytes_value<32>& b) { return detail::fr_op(1, a, &b); }
inline bytes_value<32> operator*(const bytes_value<32>& a, const bytes_value<32>& b) { return detail::fr_op(2, a, &b); }
inline bytes_value<32> operator-(const bytes_value<32>& a) { return detail::fr_op(5, a, &a); }
inline bytes_value<32> inversed(const bytes_value<32>& a) { return detail::fr_op(3, a, &a); }
template <std::size_t N> typename std::enable_if<N == 96 || N == 192, bytes_value<N>>::type operator+(const bytes_value<N>& a, const bytes_value<N>& b) {
    std::uint8_t two[2 * N];
    std::memcpy(two, a.data.data(), N); std::memcpy(two + N, b.data.data(), N);
    bytes_value<N> r;
    context& c = context::instance();
    c.check(N == 96 ? b2g16_g1_sum(c.get(), two, 2, r.data.data()) : b2g16_g2_sum(c.get(), two, 2, r.data.data()));
    return r;
}
template <std::size_t N> typename std::enable_if<N == 96 || N == 192, bytes_value<N>>::type operator-(const bytes_value<N>& a) {
    bytes_value<N> r = a;                                   // (x, -y); the identity (all zero) stays the identity
    for (std::size_t o = N / 2; o < N; o += 48) detail::fq_negate(r.data.data() + o);
    return r;
}
template <std::size_t N> typename std::enable_if<N == 96 || N == 192, bytes_value<N>>::type operator-(const bytes_value<N>& a, const bytes_value<N>& b) { return a + (-b); }
template <std::size_t N> typename std::enable_if<N == 96 || N == 192, bytes_value<N>>::type operator*(const bytes_value<32>& k, const bytes_value<N>& p) {
    bytes_value<N> r;
    context& c = context::instance();
    c.check(N == 96 ? b2g16_msm_g1(c.get(), p.data.data(), k.data.data(), 1, r.data.data()) : b2g16_msm_g2(c.get(), p.data.data(), k.data.data(), 1, r.data.data()));
    return r;
}

// Cache entry names of S/core/parameter_cache.hpp:50-94,166-173 (v28-<prefix>-<sha256>.{params,vk,meta} under
// $FIL_PROOFS_PARAMETER_CACHE) and the metadata file (:108-123).
inline std::string cache_string(int (*fn)(const char*, int, char*, std::size_t), const char* a, int k) {
    char buf[4096];
    int rc = fn(a, k, buf, sizeof buf);
    if (rc != B2G16_OK) throw error(rc, "cache path does not fit");
    return buf;
}
inline std::string parameter_cache_identifier(const std::string& cache_prefix, const std::string& parameter_set_identifier) {
    char buf[512];
    int rc = b2g16_cache_identifier(cache_prefix.c_str(), parameter_set_identifier.c_str(), buf, sizeof buf);
    if (rc != B2G16_OK) throw error(rc, "cache identifier does not fit");
    return buf;
}
inline std::string parameter_cache_params_path(const std::string& id) { return cache_string(b2g16_cache_path, id.c_str(), 0); }
inline std::string parameter_cache_verifying_key_path(const std::string& id) { return cache_string(b2g16_cache_path, id.c_str(), 1); }
inline std::string parameter_cache_metadata_path(const std::string& id) { return cache_string(b2g16_cache_path, id.c_str(), 2); }
struct cache_entry_metadata { std::uint64_t sector_size; };
inline cache_entry_metadata read_cached_metadata(const std::string& path) {
    cache_entry_metadata m{};
    int rc = b2g16_meta_read(path.c_str(), &m.sector_size);
    if (rc != B2G16_OK) throw error(rc, "cannot read " + path);
    return m;
}
inline cache_entry_metadata write_cached_metadata(const std::string& path, cache_entry_metadata m) {
    int rc = b2g16_meta_write(path.c_str(), m.sector_size);
    if (rc != B2G16_OK) throw error(rc, "cannot write " + path);
    return m;
}

}  // namespace b2g16

namespace algebra {
namespace fields {
struct bls12_scalar_field_381 {
    typedef b2g16::bytes_value<32> value_type;                 // 32-byte LE canonical (fr32.hpp:36-52)
    static constexpr std::size_t modulus_bits = 255;
};
}  // namespace fields
namespace curves {
template <std::size_t Bits> struct bls12;
template <> struct bls12<381> {
    typedef fields::bls12_scalar_field_381 scalar_field_type;
    template <typename = void> struct g1_type { typedef b2g16::bytes_value<96> value_type; static constexpr int group = 1; };
    template <typename = void> struct g2_type { typedef b2g16::bytes_value<192> value_type; static constexpr int group = 2; };
};
}  // namespace curves
namespace policies {
struct multiexp_method_naive_plain {};
struct multiexp_method_bos_coster {};
struct multiexp_method_BDLO12 {};
}  // namespace policies

namespace detail {
template <class It> const std::uint8_t* flat(It first, It last, std::vector<std::uint8_t>& tmp) {
    typedef typename std::iterator_traits<It>::value_type V;
    const std::size_t n = (std::size_t)std::distance(first, last);
    tmp.resize(n * V::size);
    std::size_t k = 0;
    for (It it = first; it != last; ++it, ++k) std::memcpy(tmp.data() + k * V::size, it->data.data(), V::size);
    return tmp.data();
}
}  // namespace detail

// sum_i scalars[i] * bases[i]; `chunks` (the CPU thread count upstream) is accepted and ignored.
// The method tag is accepted for source compatibility: every method yields the same group element.
template <class Method = policies::multiexp_method_BDLO12, class BaseIt, class ScalarIt>
typename std::iterator_traits<BaseIt>::value_type multiexp(BaseIt bases_begin, BaseIt bases_end, ScalarIt scalars_begin,
                                                            ScalarIt scalars_end, std::size_t chunks = 1) {
    (void)chunks;
    typedef typename std::iterator_traits<BaseIt>::value_type P;
    static_assert(P::size == 96 || P::size == 192, "bases must be G1 (96 B) or G2 (192 B) affine values");
    const std::size_t n = (std::size_t)std::distance(bases_begin, bases_end);
    if ((std::size_t)std::distance(scalars_begin, scalars_end) != n) throw b2g16::error(B2G16_E_INVALID_ARG, "multiexp: bases and scalars differ in length");
    std::vector<std::uint8_t> b, s;
    const std::uint8_t* pb = detail::flat(bases_begin, bases_end, b);
    const std::uint8_t* ps = detail::flat(scalars_begin, scalars_end, s);
    b2g16::context& c = b2g16::context::instance();
    P out;
    c.check(P::size == 96 ? b2g16_msm_g1(c.get(), pb, ps, n, out.data.data()) : b2g16_msm_g2(c.get(), pb, ps, n, out.data.data()));
    return out;
}
template <class Method = policies::multiexp_method_BDLO12, class BaseIt, class ScalarIt>
typename std::iterator_traits<BaseIt>::value_type multiexp_with_mixed_addition(BaseIt bb, BaseIt be, ScalarIt sb, ScalarIt se, std::size_t chunks = 1) {
    return multiexp<Method>(bb, be, sb, se, chunks);          // 0/1 scalars need no special path on the device
}
}  // namespace algebra

namespace math {
// Radix-2 domain of size exactly m = 2^k over Fr (generator 7).  Natural order in and
// out, in place.  m must be a power of two: the "domain trap" of SURVEY 8(c) -- the library refuses
// other sizes rather than silently switching to an extended/step domain.
template <class FieldType> class evaluation_domain {
public:
    typedef typename FieldType::value_type value_type;
    std::size_t m;
    unsigned log_m;
    explicit evaluation_domain(std::size_t size) : m(size), log_m(0) {
        while (((std::size_t)1 << log_m) < m) ++log_m;
        if (((std::size_t)1 << log_m) != m) throw b2g16::error(B2G16_E_INVALID_ARG, "evaluation_domain: size must be a power of two");
    }
    void fft(std::vector<value_type>& a) const { run(a, 0, 0); }
    void inverse_fft(std::vector<value_type>& a) const { run(a, 1, 0); }
    // evaluate on / interpolate from the coset g*D, g = multiplicative generator 7
    void coset_fft(std::vector<value_type>& a) const { run(a, 0, 1); }
    void inverse_coset_fft(std::vector<value_type>& a) const { run(a, 1, 1); }
    // older libsnark-style spellings
    void FFT(std::vector<value_type>& a) const { fft(a); }
    void iFFT(std::vector<value_type>& a) const { inverse_fft(a); }
    void cosetFFT(std::vector<value_type>& a) const { coset_fft(a); }
    void icosetFFT(std::vector<value_type>& a) const { inverse_coset_fft(a); }
private:
    void run(std::vector<value_type>& a, int inverse, int coset) const {
        if (a.size() != m) throw b2g16::error(B2G16_E_INVALID_ARG, "evaluation_domain: vector length differs from the domain size");
        static_assert(sizeof(value_type) == 32, "Fr values are 32 contiguous bytes");
        b2g16::context& c = b2g16::context::instance();
        c.check(b2g16_ntt(c.get(), reinterpret_cast<std::uint8_t*>(a.data()), log_m, inverse, coset));
    }
};
template <class FieldType> std::shared_ptr<evaluation_domain<FieldType>> make_evaluation_domain(std::size_t m) {
    std::size_t p = 1;
    while (p < m) p <<= 1;
    return std::make_shared<evaluation_domain<FieldType>>(p);
}
}  // namespace math

namespace zk {
namespace snark {

// Constraint system in CSR form, z = (1, primary, auxiliary); domain 2^log_m >= n_cons + n_in + 1.
template <class FieldType> struct r1cs_constraint_system {
    typedef typename FieldType::value_type value_type;
    struct matrix { std::vector<std::uint32_t> row_ptr, col; std::vector<value_type> val; };
    std::uint32_t log_m = 0;
    std::size_t n_cons = 0, n_in = 0, n_aux = 0;
    matrix A, B, C;
    std::size_t num_inputs() const { return n_in; }
    std::size_t num_variables() const { return n_in + n_aux; }
    std::size_t num_constraints() const { return n_cons; }
    b2g16_r1cs_view view() const {
        b2g16_r1cs_view v;
        v.log_m = log_m; v.n_cons = n_cons; v.n_in = n_in; v.n_aux = n_aux;
        const matrix* m[3] = {&A, &B, &C};
        b2g16_csr* o[3] = {&v.A, &v.B, &v.C};
        for (int i = 0; i < 3; ++i) {
            o[i]->n_rows = n_cons; o[i]->nnz = m[i]->col.size();
            o[i]->row_ptr = m[i]->row_ptr.data(); o[i]->col = m[i]->col.data();
            o[i]->val = reinterpret_cast<const std::uint8_t*>(m[i]->val.data());
        }
        return v;
    }
};

template <class CurveType> struct r1cs_gg_ppzksnark {
    typedef CurveType curve_type;
    typedef typename CurveType::scalar_field_type scalar_field_type;
    typedef typename scalar_field_type::value_type fr_value;
    typedef typename CurveType::template g1_type<>::value_type g1_value;
    typedef typename CurveType::template g2_type<>::value_type g2_value;
    typedef r1cs_constraint_system<scalar_field_type> constraint_system_type;
    typedef std::vector<fr_value> primary_input_type;
    typedef std::vector<fr_value> auxiliary_input_type;

    // (A in G1, B in G2, C in G1); bytes() = 192-byte compressed form (constants.hpp:93)
    struct proof_type {
        std::array<std::uint8_t, 192> compressed{};
        g1_value g_A, g_C;
        g2_value g_B;
        const std::array<std::uint8_t, 192>& bytes() const { return compressed; }
        // multi_proof write / new_from_reader (F/api/seal.hpp:304-308, 382-383): 192 bytes per proof
        void write(std::uint8_t* out) const { std::memcpy(out, compressed.data(), 192); }
        static proof_type read(const std::uint8_t* in) {                // validates encoding, curve, subgroup
            proof_type p;
            std::memcpy(p.compressed.data(), in, 192);
            int rc = b2g16_proof_decode(in, p.g_A.data.data(), p.g_B.data.data(), p.g_C.data.data());
            if (rc != B2G16_OK) throw b2g16::error(rc, "proof: malformed 192-byte encoding");
            return p;
        }
    };
    struct verification_key_type {
        g1_value alpha_g1; g2_value beta_g2, gamma_g2, delta_g2; std::vector<g1_value> gamma_ABC_g1;
        g1_value beta_g1, delta_g1;      // carried by the file formats only (the verifier does not use them)
        // write_cached_verifying_key / read_cached_verifying_key (parameter_cache.hpp:131-145): the `.vk` cache entry
        void write(const std::string& path) const {
            b2g16_params p;
            std::memset(&p, 0, sizeof p);
            std::memcpy(p.alpha_g1, alpha_g1.data.data(), 96); std::memcpy(p.beta_g1, beta_g1.data.data(), 96);
            std::memcpy(p.beta_g2, beta_g2.data.data(), 192); std::memcpy(p.gamma_g2, gamma_g2.data.data(), 192);
            std::memcpy(p.delta_g1, delta_g1.data.data(), 96); std::memcpy(p.delta_g2, delta_g2.data.data(), 192);
            p.n_ic = gamma_ABC_g1.size();
            p.ic = const_cast<std::uint8_t*>(reinterpret_cast<const std::uint8_t*>(gamma_ABC_g1.data()));
            int rc = b2g16_vk_file_write(path.c_str(), &p);
            if (rc != B2G16_OK) throw b2g16::error(rc, "cannot write " + path);
        }
        static verification_key_type read(const std::string& path) {
            b2g16_params* p = nullptr;
            int rc = b2g16_vk_file_read(path.c_str(), &p);
            if (rc != B2G16_OK) throw b2g16::error(rc, "cannot read " + path);
            struct guard { b2g16_params* p; ~guard() { b2g16_params_free(p); } } g{p};
            verification_key_type vk;
            vk.from_params(*p);
            return vk;
        }
        void from_params(const b2g16_params& p) {
            std::memcpy(alpha_g1.data.data(), p.alpha_g1, 96); std::memcpy(beta_g1.data.data(), p.beta_g1, 96);
            std::memcpy(beta_g2.data.data(), p.beta_g2, 192); std::memcpy(gamma_g2.data.data(), p.gamma_g2, 192);
            std::memcpy(delta_g1.data.data(), p.delta_g1, 96); std::memcpy(delta_g2.data.data(), p.delta_g2, 192);
            gamma_ABC_g1.resize(p.n_ic);
            if (p.n_ic) std::memcpy(gamma_ABC_g1.data(), p.ic, 96 * p.n_ic);
        }
    };
    typedef verification_key_type verifying_key_type;            // both spellings occur (scheme_params.hpp:44, multi_proof.hpp:43)
    // the `pvk` kept next to the mapped proving key (mapped_scheme_params.hpp:61): alpha/beta Miller value
    // cached on the host, IC bases resident on the device
    class processed_verification_key_type {
        b2g16_pvk* h_ = nullptr;
        std::size_t n_in_ = 0;
    public:
        processed_verification_key_type() {}
        explicit processed_verification_key_type(const verification_key_type& vk) : n_in_(vk.gamma_ABC_g1.size() - 1) {
            std::vector<std::uint8_t> b(672 + 96 * vk.gamma_ABC_g1.size());
            std::memcpy(b.data(), vk.alpha_g1.data.data(), 96); std::memcpy(b.data() + 96, vk.beta_g2.data.data(), 192);
            std::memcpy(b.data() + 288, vk.gamma_g2.data.data(), 192); std::memcpy(b.data() + 480, vk.delta_g2.data.data(), 192);
            std::memcpy(b.data() + 672, vk.gamma_ABC_g1.data(), 96 * vk.gamma_ABC_g1.size());
            b2g16::context& c = b2g16::context::instance();
            c.check(b2g16_vk_prepare(c.get(), b.data(), vk.gamma_ABC_g1.size(), &h_));
        }
        processed_verification_key_type(processed_verification_key_type&& o) noexcept : h_(o.h_), n_in_(o.n_in_) { o.h_ = nullptr; }
        processed_verification_key_type& operator=(processed_verification_key_type&& o) noexcept { reset(); h_ = o.h_; n_in_ = o.n_in_; o.h_ = nullptr; return *this; }
        processed_verification_key_type(const processed_verification_key_type&) = delete;
        processed_verification_key_type& operator=(const processed_verification_key_type&) = delete;
        ~processed_verification_key_type() { reset(); }
        void reset() { if (h_) { b2g16_pvk_free(b2g16::context::instance().get(), h_); h_ = nullptr; } }
        const b2g16_pvk* handle() const { return h_; }
        std::size_t num_inputs() const { return n_in_; }
    };

    // Device-resident proving key (scheme_params.hpp:38-67 layout: vk, h, l, a, b_g1, b_g2), dense.
    class proving_key_type {
        b2g16_pk* h_ = nullptr;
        std::size_t n_vars_ = 0;
    public:
        proving_key_type() {}
        proving_key_type(b2g16_pk* h, std::size_t n_vars) : h_(h), n_vars_(n_vars) {}
        proving_key_type(proving_key_type&& o) noexcept : h_(o.h_), n_vars_(o.n_vars_) { o.h_ = nullptr; }
        proving_key_type& operator=(proving_key_type&& o) noexcept { reset(); h_ = o.h_; n_vars_ = o.n_vars_; o.h_ = nullptr; return *this; }
        proving_key_type(const proving_key_type&) = delete;
        proving_key_type& operator=(const proving_key_type&) = delete;
        ~proving_key_type() { reset(); }
        void reset() { if (h_) { b2g16_pk_free(b2g16::context::instance().get(), h_); h_ = nullptr; } }
        b2g16_pk* handle() const { return h_; }
        std::size_t num_variables_with_one() const { return n_vars_; }
        // upload a host key (the in-memory scheme_params vectors) -- caller keeps its buffers
        static proving_key_type load(const constraint_system_type& cs, const g1_value& alpha_g1, const g1_value& beta_g1, const g1_value& delta_g1,
                                     const g2_value& beta_g2, const g2_value& delta_g2, const std::vector<g1_value>& a, const std::vector<g1_value>& b_g1,
                                     const std::vector<g2_value>& b_g2, const std::vector<g1_value>& l, const std::vector<g1_value>& h) {
            b2g16_r1cs_view rv = cs.view();
            b2g16_pk_view v;
            v.alpha_g1 = alpha_g1.data.data(); v.beta_g1 = beta_g1.data.data(); v.delta_g1 = delta_g1.data.data();
            v.beta_g2 = beta_g2.data.data(); v.delta_g2 = delta_g2.data.data();
            v.a = reinterpret_cast<const std::uint8_t*>(a.data()); v.b_g1 = reinterpret_cast<const std::uint8_t*>(b_g1.data());
            v.b_g2 = reinterpret_cast<const std::uint8_t*>(b_g2.data()); v.l = reinterpret_cast<const std::uint8_t*>(l.data());
            v.h = reinterpret_cast<const std::uint8_t*>(h.data()); v.cs = &rv;
            b2g16::context& c = b2g16::context::instance();
            b2g16_pk* pk = nullptr;
            c.check(b2g16_pk_load(c.get(), &v, &pk));
            return proving_key_type(pk, 1 + cs.n_in + cs.n_aux);
        }
        // write_cached_params (parameter_cache.hpp:147-153): the resident key + vk as a bellperson-layout
        // `.params` file; `filtered` drops the points at infinity from a / b_g1 / b_g2 the way bellman's
        // generator does, so the file is what Filecoin's parameter files look like.
        void write_cached(const std::string& path, const verification_key_type& vk, bool filtered = true) const {
            b2g16::context& c = b2g16::context::instance();
            std::uint64_t shp[5];                                         // log_m, n_cons, n_in, n_aux, n_vars
            c.check(b2g16_pk_shape(h_, shp));
            const std::uint64_t nv = shp[4], counts[5] = {nv, nv, nv, shp[3], (std::uint64_t(1) << shp[0]) - 1};
            const std::size_t width[5] = {96, 96, 192, 96, 96};           // which = 0 a, 1 b_g1, 2 b_g2, 3 l, 4 h
            std::vector<std::uint8_t> q[5];
            b2g16_params p;
            std::memset(&p, 0, sizeof p);
            std::uint64_t kept[3] = {nv, nv, nv};
            for (int w = 0; w < 5; ++w) {
                q[w].resize(counts[w] * width[w]);
                c.check(b2g16_pk_export(c.get(), h_, w, q[w].data(), q[w].size()));
                if (filtered && w < 3) {
                    std::vector<std::uint8_t> f(q[w].size());
                    c.check(b2g16_query_filter(q[w].data(), nv, width[w], f.data(), nullptr, &kept[w]));
                    f.resize(kept[w] * width[w]);
                    q[w].swap(f);
                }
            }
            std::uint8_t* single[5] = {p.alpha_g1, p.beta_g1, p.delta_g1, p.beta_g2, p.delta_g2};   // which = 5..9
            const std::size_t sw[5] = {96, 96, 96, 192, 192};
            for (int w = 0; w < 5; ++w) c.check(b2g16_pk_export(c.get(), h_, 5 + w, single[w], sw[w]));
            std::memcpy(p.gamma_g2, vk.gamma_g2.data.data(), 192);
            p.n_ic = vk.gamma_ABC_g1.size(); p.ic = const_cast<std::uint8_t*>(reinterpret_cast<const std::uint8_t*>(vk.gamma_ABC_g1.data()));
            p.n_a = kept[0]; p.a = q[0].data(); p.n_b_g1 = kept[1]; p.b_g1 = q[1].data(); p.n_b_g2 = kept[2]; p.b_g2 = q[2].data();
            p.n_l = counts[3]; p.l = q[3].data(); p.n_h = counts[4]; p.h = q[4].data();
            int rc = b2g16_params_write(path.c_str(), &p);
            if (rc != B2G16_OK) throw b2g16::error(rc, "cannot write " + path);
        }
        // read_cached_params (parameter_cache.hpp:125-129): `.params` file -> resident key (+ the vk it
        // carries).  Dense files load as they are; density-filtered a / b queries are expanded with the
        // column occupancy of the constraint system's A / B -- plus ONE and every public input for a, whose
        // input-consistency rows make those polynomials non-zero (a file that does not fit is refused).
        static proving_key_type read_cached(const constraint_system_type& cs, const std::string& path, verification_key_type* vk_out = nullptr) {
            b2g16::context& c = b2g16::context::instance();
            b2g16_params* p = nullptr;
            int rc = b2g16_params_read(path.c_str(), &p);
            if (rc != B2G16_OK) throw b2g16::error(rc, "cannot read " + path);
            struct guard { b2g16_params* p; ~guard() { b2g16_params_free(p); } } g{p};
            const std::uint64_t nv = 1 + cs.n_in + cs.n_aux;
            if (p->n_l != cs.n_aux || p->n_h + 1 != (std::uint64_t(1) << cs.log_m) || p->n_ic != cs.n_in + 1)
                throw b2g16::error(B2G16_E_INVALID_ARG, path + ": parameters of another circuit (l / h / ic lengths)");
            b2g16_r1cs_view rv = cs.view();
            std::vector<std::uint8_t> store[3], dens(nv);
            auto dense = [&](const std::uint8_t* q, std::uint64_t n, const b2g16_csr& m, std::uint64_t n_always, std::size_t w, std::vector<std::uint8_t>& s) {
                if (n == nv) return q;
                if (b2g16_query_density(&m, nv, n_always, dens.data(), nullptr) != B2G16_OK) throw b2g16::error(B2G16_E_INVALID_ARG, "constraint system: column out of range");
                s.resize(nv * w);
                if (b2g16_query_expand(q, n, dens.data(), nv, w, s.data()) != B2G16_OK)
                    throw b2g16::error(B2G16_E_INVALID_ARG, path + ": filtered query does not match the constraint system's density");
                return static_cast<const std::uint8_t*>(s.data());
            };
            b2g16_pk_view v;
            v.alpha_g1 = p->alpha_g1; v.beta_g1 = p->beta_g1; v.delta_g1 = p->delta_g1; v.beta_g2 = p->beta_g2; v.delta_g2 = p->delta_g2;
            v.a = dense(p->a, p->n_a, rv.A, cs.n_in + 1, 96, store[0]);      // ONE and the inputs are always in the a query
            v.b_g1 = dense(p->b_g1, p->n_b_g1, rv.B, 0, 96, store[1]);
            v.b_g2 = dense(p->b_g2, p->n_b_g2, rv.B, 0, 192, store[2]);
            v.l = p->l; v.h = p->h; v.cs = &rv;
            b2g16_pk* pk = nullptr;
            c.check(b2g16_pk_load(c.get(), &v, &pk));
            if (vk_out) vk_out->from_params(*p);
            return proving_key_type(pk, nv);
        }
    };
    struct keypair_type { proving_key_type first; verification_key_type second; };
};

// generate<Scheme>(cs, toxic): throw-away parameters from explicit toxic waste (tau, alpha, beta, gamma,
// delta), computed on the device -- what parameter_cache.hpp:185-200 does through
// groth16::generate_random_parameters for cache misses.
template <class Scheme>
typename Scheme::keypair_type generate(const typename Scheme::constraint_system_type& cs, const std::array<typename Scheme::fr_value, 5>& toxic) {
    b2g16::context& c = b2g16::context::instance();
    b2g16_r1cs_view rv = cs.view();
    std::vector<std::uint8_t> vk(96 + 3 * 192 + 96 * (cs.n_in + 1));
    b2g16_pk* pk = nullptr;
    c.check(b2g16_pk_generate(c.get(), &rv, reinterpret_cast<const std::uint8_t*>(toxic.data()), &pk, vk.data()));
    typename Scheme::keypair_type kp;
    kp.first = typename Scheme::proving_key_type(pk, 1 + cs.n_in + cs.n_aux);
    std::memcpy(kp.second.alpha_g1.data.data(), vk.data(), 96);
    std::memcpy(kp.second.beta_g2.data.data(), vk.data() + 96, 192);
    std::memcpy(kp.second.gamma_g2.data.data(), vk.data() + 288, 192);
    std::memcpy(kp.second.delta_g2.data.data(), vk.data() + 480, 192);
    kp.second.gamma_ABC_g1.resize(cs.n_in + 1);
    std::memcpy(kp.second.gamma_ABC_g1.data(), vk.data() + 672, 96 * (cs.n_in + 1));
    c.check(b2g16_pk_export(c.get(), pk, 6, kp.second.beta_g1.data.data(), 96));     // for the .vk / .params files
    c.check(b2g16_pk_export(c.get(), pk, 7, kp.second.delta_g1.data.data(), 96));
    return kp;
}

// prove<Scheme>(pk, primary, auxiliary, r, s): explicit r, s (bit-exact parity).
template <class Scheme>
typename Scheme::proof_type prove(const typename Scheme::proving_key_type& pk, const typename Scheme::primary_input_type& primary,
                                  const typename Scheme::auxiliary_input_type& auxiliary, const typename Scheme::fr_value& r,
                                  const typename Scheme::fr_value& s, bool priority = false) {
    typedef typename Scheme::fr_value fr;
    if (1 + primary.size() + auxiliary.size() != pk.num_variables_with_one()) throw b2g16::error(B2G16_E_INVALID_ARG, "prove: assignment length differs from the key");
    std::vector<fr> z(1 + primary.size() + auxiliary.size());
    z[0].data[0] = 1;                                              // z_0 = ONE
    std::copy(primary.begin(), primary.end(), z.begin() + 1);
    std::copy(auxiliary.begin(), auxiliary.end(), z.begin() + 1 + primary.size());
    typename Scheme::proof_type proof;
    b2g16_trace tr;
    std::memset(&tr, 0, sizeof(tr));
    tr.A = proof.g_A.data.data(); tr.B = proof.g_B.data.data(); tr.C = proof.g_C.data.data();
    b2g16::context& c = b2g16::context::instance();
    c.check(b2g16_prove(c.get(), pk.handle(), reinterpret_cast<const std::uint8_t*>(z.data()), r.data.data(), s.data.data(), priority ? 1 : 0,
                        proof.compressed.data(), &tr));
    return proof;
}
// prove<Scheme>(pk, primary, auxiliary): r, s drawn host-side (upstream behaviour; not reproducible)
template <class Scheme>
typename Scheme::proof_type prove(const typename Scheme::proving_key_type& pk, const typename Scheme::primary_input_type& primary,
                                  const typename Scheme::auxiliary_input_type& auxiliary) {
    std::random_device rd;
    typename Scheme::fr_value r, s;
    for (auto* v : {&r, &s}) { for (auto& b : v->data) b = (std::uint8_t)rd(); v->data[31] &= 0x3f; }   // < 2^254 < r
    return prove<Scheme>(pk, primary, auxiliary, r, s);
}

// verify<Scheme>(pvk, primary_input, proof): the check behind compound_proof::verify
// (S/core/proof/compound_proof.hpp:97-102) that seal_commit_phase2 runs on its own output
// (F/api/seal.hpp:310-313).  A malformed proof is `false`, not an exception.
template <class Scheme>
bool verify(const typename Scheme::processed_verification_key_type& pvk, const typename Scheme::primary_input_type& primary,
            const typename Scheme::proof_type& proof) {
    b2g16::context& c = b2g16::context::instance();
    int ok = 0;
    c.check(b2g16_verify(c.get(), pvk.handle(), reinterpret_cast<const std::uint8_t*>(primary.data()), primary.size(), proof.compressed.data(), &ok));
    return ok != 0;
}
template <class Scheme>
bool verify(const typename Scheme::verification_key_type& vk, const typename Scheme::primary_input_type& primary, const typename Scheme::proof_type& proof) {
    typename Scheme::processed_verification_key_type pvk(vk);
    return verify<Scheme>(pvk, primary, proof);
}
// all partitions of a multi_proof in one product (verify_batch_seal, F/api/seal.hpp:409-485)
template <class Scheme>
bool verify_batch(const typename Scheme::processed_verification_key_type& pvk, const std::vector<typename Scheme::primary_input_type>& primaries,
                  const std::vector<typename Scheme::proof_type>& proofs) {
    if (primaries.size() != proofs.size()) throw b2g16::error(B2G16_E_INVALID_ARG, "verify_batch: one input vector per proof");
    std::vector<std::uint8_t> in, pr;
    for (const auto& v : primaries) {
        if (v.size() != pvk.num_inputs()) throw b2g16::error(B2G16_E_INVALID_ARG, "verify_batch: public input count differs from the key");
        in.insert(in.end(), reinterpret_cast<const std::uint8_t*>(v.data()), reinterpret_cast<const std::uint8_t*>(v.data()) + 32 * v.size());
    }
    for (const auto& p : proofs) pr.insert(pr.end(), p.compressed.begin(), p.compressed.end());
    b2g16::context& c = b2g16::context::instance();
    int ok = 0;
    c.check(b2g16_verify_batch(c.get(), pvk.handle(), in.data(), pvk.num_inputs(), pr.data(), proofs.size(), nullptr, &ok));
    return ok != 0;
}

}  // namespace snark
}  // namespace zk
}  // namespace crypto3
}  // namespace nil
