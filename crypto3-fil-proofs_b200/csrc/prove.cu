// Groth16 (r1cs_gg_ppzksnark, BLS12-381) prover on the device-resident proving key.
// Stands where compound_proof::circuit_proofs
// (libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137)
// calls crypto3-zk `prove<r1cs_gg_ppzksnark<bls12<381>>>(pk, primary, aux)` for
// seal_commit_phase2 (libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:298-308) and
// generate_{winning,window}_post (.../api/post.hpp:225,345).  Algorithm: SURVEY App. B.
//   device: ingest z -> a,b,c = A z, B z, C z (CSR SpMV, + input-consistency rows)
//           -> 7 NTTs (H coefficients) -> 5 MSMs (A, B1, B2, L over z; H over h)
//   host:   A = alpha + A' + r delta ; B = beta + B' + s delta ; C = H' + L' + sA + rB1 - rs delta
//           -> 192-byte compressed proof (F/constants.hpp:93 SINGLE_PARTITION_PROOF_LEN)
// The key layout mirrors scheme_params (S/core/crypto/scheme_params.hpp:38-67), dense.
#include <cstring>
#include <future>
#include <thread>
#include <vector>
#include "msm_impl.cuh"
#include "api_util.cuh"
#include "pk.cuh"
#include "ntt_kernels.cuh"
#include "host_bls.hpp"

namespace b2 {

// ------------------------------------------------------------------------------ kernels
// a,b,c evaluations on the domain; rows >= n_cons of A carry the inputs
__global__ void __launch_bounds__(256) abc_spmv_kernel(DevCsr A, DevCsr B, DevCsr C, const Fr* __restrict__ z, uint64_t n_cons, uint64_t n_in, uint64_t m,
                                                       Fr* __restrict__ a, Fr* __restrict__ b, Fr* __restrict__ c) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < m; i += (uint64_t)gridDim.x * blockDim.x) {
        Fr sa = Fr::zero(), sb = Fr::zero(), sc = Fr::zero();
        if (i < n_cons) {
            for (uint32_t k = A.row_ptr[i]; k < A.row_ptr[i + 1]; ++k) sa = sa + fr_load(A.val + k) * fr_load(z + A.col[k]);
            for (uint32_t k = B.row_ptr[i]; k < B.row_ptr[i + 1]; ++k) sb = sb + fr_load(B.val + k) * fr_load(z + B.col[k]);
            for (uint32_t k = C.row_ptr[i]; k < C.row_ptr[i + 1]; ++k) sc = sc + fr_load(C.val + k) * fr_load(z + C.col[k]);
        } else if (i - n_cons <= n_in) {
            sa = fr_load(z + (i - n_cons));
        }
        fr_store(a + i, sa); fr_store(b + i, sb); fr_store(c + i, sc);
    }
}
// The same evaluations for the rows row0 + stride * k, k < count, written as canonical scalars: a rank's share
// of a, b, c when the witness map of one proof is spread over several GPUs (cyclic row distribution, which is
// what the sharded inverse transform consumes; b2g16.dist.prove_sharded_full).
__global__ void __launch_bounds__(256) abc_spmv_strided_kernel(DevCsr A, DevCsr B, DevCsr C, const Fr* __restrict__ z, uint64_t n_cons, uint64_t n_in,
                                                               uint64_t row0, uint64_t stride, uint64_t count, Fr* __restrict__ a, Fr* __restrict__ b, Fr* __restrict__ c) {
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < count; k += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = row0 + stride * k;
        Fr sa = Fr::zero(), sb = Fr::zero(), sc = Fr::zero();
        if (i < n_cons) {
            for (uint32_t q = A.row_ptr[i]; q < A.row_ptr[i + 1]; ++q) sa = sa + fr_load(A.val + q) * fr_load(z + A.col[q]);
            for (uint32_t q = B.row_ptr[i]; q < B.row_ptr[i + 1]; ++q) sb = sb + fr_load(B.val + q) * fr_load(z + B.col[q]);
            for (uint32_t q = C.row_ptr[i]; q < C.row_ptr[i + 1]; ++q) sc = sc + fr_load(C.val + q) * fr_load(z + C.col[q]);
        } else if (i - n_cons <= n_in) {
            sa = fr_load(z + (i - n_cons));
        }
        Fr t;
        sa.to_canonical(t.v); fr_store(a + k, t);
        sb.to_canonical(t.v); fr_store(b + k, t);
        sc.to_canonical(t.v); fr_store(c + k, t);
    }
}
// out = (a * b - c) * zinv on canonical scalars (the quotient by Z on the coset, pointwise; any distribution)
__global__ void __launch_bounds__(256) h_pointwise_kernel(const Fr* __restrict__ a, const Fr* __restrict__ b, const Fr* __restrict__ c, uint64_t count, Fr zinv,
                                                          Fr* __restrict__ out) {
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < count; k += (uint64_t)gridDim.x * blockDim.x) {
        // canonical x is the Montgomery form of x/R: (a R^2 -> Montgomery a) * b -> canonical a b; a product with a
        // Montgomery-form constant keeps a canonical operand canonical
        const Fr ab = (fr_load(a + k) * Fr::r2()) * fr_load(b + k);
        fr_store(out + k, (ab - fr_load(c + k)) * zinv);
    }
}
// one layer of the forward-built synthetic instance: z[out0 + i] = (A_i . z)(B_i . z), rows [r0, r1)
__global__ void __launch_bounds__(256) synth_layer_kernel(DevCsr A, DevCsr B, Fr* __restrict__ z, uint64_t r0, uint64_t r1, uint64_t out0) {
    for (uint64_t i = r0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < r1; i += (uint64_t)gridDim.x * blockDim.x) {
        Fr sa = Fr::zero(), sb = Fr::zero();
        for (uint32_t k = A.row_ptr[i]; k < A.row_ptr[i + 1]; ++k) sa = sa + fr_load(A.val + k) * fr_load(z + A.col[k]);
        for (uint32_t k = B.row_ptr[i]; k < B.row_ptr[i + 1]; ++k) sb = sb + fr_load(B.val + k) * fr_load(z + B.col[k]);
        fr_store(z + out0 + i, sa * sb);
    }
}


// ------------------------------------------------- hashed synthetic instance (bench/test support)
// Same shape as python/b2g16/synth.py (3 non-zeros per row in A and B over variables of earlier
// layers, C_i = the fresh variable of row i) but every column and coefficient is a counter-based hash
// of (seed, matrix, row, slot), so the instance can be generated on the device at sizes whose host
// arrays would not fit (2^27 rows = 33 GB) and regenerated bit-identically in numpy at small sizes
// (synth.HashedR1cs) for parity against the oracle.
B2_HD uint64_t synth_mix(uint64_t x) {                       // splitmix64
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
B2_HD uint64_t synth_stream(uint64_t seed, uint64_t tag) { return synth_mix(seed ^ (tag * 0xD6E8FEB86659FD93ull)); }
constexpr int SYNTH_MAX_LAYERS = 96;
struct SynthLayers { int n; uint64_t first[SYNTH_MAX_LAYERS + 1]; };     // first row of each layer, first[n] = n_cons

__global__ void __launch_bounds__(256) synth_rows_kernel(uint64_t seed, uint64_t n_cons, uint64_t base_vars, SynthLayers L,
                                                         uint32_t* __restrict__ a_rp, uint32_t* __restrict__ a_col, Fr* __restrict__ a_val,
                                                         uint32_t* __restrict__ b_rp, uint32_t* __restrict__ b_col, Fr* __restrict__ b_val,
                                                         uint32_t* __restrict__ c_rp, uint32_t* __restrict__ c_col, Fr* __restrict__ c_val) {
    const uint64_t sa = synth_stream(seed, 1), sb = synth_stream(seed, 2), va = synth_stream(seed, 3), vb = synth_stream(seed, 4);
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i <= n_cons; i += (uint64_t)gridDim.x * blockDim.x) {
        a_rp[i] = (uint32_t)(3 * i); b_rp[i] = (uint32_t)(3 * i); c_rp[i] = (uint32_t)i;
        if (i == n_cons) break;
        int lo = 0, hi = L.n;                                // layer of row i: largest l with first[l] <= i
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (L.first[mid] <= i) lo = mid; else hi = mid; }
        const uint64_t visible = base_vars + L.first[lo];
        for (int j = 0; j < 3; ++j) {
            const uint64_t e = 3 * i + j;
            a_col[e] = (uint32_t)((synth_mix(sa + e) >> 11) % visible);
            b_col[e] = (uint32_t)((synth_mix(sb + e) >> 11) % visible);
            uint32_t w[8];
            for (int k = 0; k < 4; ++k) { uint64_t x = synth_mix(va + 4 * e + k); w[2 * k] = (uint32_t)x; w[2 * k + 1] = (uint32_t)(x >> 32); }
            w[7] &= 0x3fffffffu;                             // < 2^254 < r
            fr_store(a_val + e, Fr::from_canonical(w));
            for (int k = 0; k < 4; ++k) { uint64_t x = synth_mix(vb + 4 * e + k); w[2 * k] = (uint32_t)x; w[2 * k + 1] = (uint32_t)(x >> 32); }
            w[7] &= 0x3fffffffu;
            fr_store(b_val + e, Fr::from_canonical(w));
        }
        c_col[i] = (uint32_t)(base_vars + i);
        fr_store(c_val + i, Fr::one());
    }
}

// --------------------------------------------------------------- Groth16 setup kernels
// lag[i] = pref * w^i / (tau - w^i)     (Lagrange basis of the domain at tau)
__global__ void __launch_bounds__(128) lagrange_kernel(Fr* lag, Fr tau, Fr pref, Fr w, uint64_t m) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < m; i += (uint64_t)gridDim.x * blockDim.x) {
        Fr wi = Fr::one(), b = w;
        for (uint64_t e = i; e; e >>= 1) { if (e & 1) wi = wi * b; b = b.sqr(); }
        fr_store(lag + i, pref * wi * (tau - wi).inv());
    }
}
__global__ void csc_hist_kernel(const uint32_t* col, uint64_t nnz, uint32_t* hist) {
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < nnz; k += (uint64_t)gridDim.x * blockDim.x) atomicAdd(hist + col[k], 1u);
}
// entries grouped by column: slot order inside a column is irrelevant (field addition commutes)
__global__ void csc_scatter_kernel(const uint32_t* row_ptr, const uint32_t* col, uint64_t n_rows, uint32_t* cursor, uint32_t* ent_row, uint32_t* ent_k) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_rows; i += (uint64_t)gridDim.x * blockDim.x)
        for (uint32_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k) {
            uint32_t slot = atomicAdd(cursor + col[k], 1u);
            ent_row[slot] = (uint32_t)i; ent_k[slot] = k;
        }
}
// out[j] = sum over entries of column j of val * lag[row]   (+ lag[n_cons + j] for j <= n_in when add_inputs)
__global__ void __launch_bounds__(128) qap_column_kernel(const uint32_t* off, const uint32_t* ent_row, const uint32_t* ent_k, const Fr* val, const Fr* lag,
                                                         uint64_t n_vars, uint64_t n_cons, uint64_t n_in, int add_inputs, Fr* out) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_vars; j += (uint64_t)gridDim.x * blockDim.x) {
        Fr acc = Fr::zero();
        for (uint32_t e = off[j]; e < off[j + 1]; ++e) acc = acc + fr_load(val + ent_k[e]) * fr_load(lag + ent_row[e]);
        if (add_inputs && j <= n_in) acc = acc + fr_load(lag + n_cons + j);
        fr_store(out + j, acc);
    }
}
// k[j] = (beta u_j + alpha v_j + w_j) * (j <= n_in ? gamma^-1 : delta^-1), canonical out
__global__ void __launch_bounds__(128) setup_k_kernel(const Fr* u, const Fr* v, const Fr* w, Fr alpha, Fr beta, Fr ginv, Fr dinv, uint64_t n_vars, uint64_t n_in, uint32_t* out_canon) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_vars; j += (uint64_t)gridDim.x * blockDim.x) {
        Fr k = (beta * fr_load(u + j) + alpha * fr_load(v + j) + fr_load(w + j)) * (j <= n_in ? ginv : dinv);
        Fr c; k.to_canonical(c.v);
        fr_store(reinterpret_cast<Fr*>(out_canon) + j, c);
    }
}
// h scalars: tau^i * zt_over_delta, canonical
__global__ void __launch_bounds__(128) setup_h_kernel(Fr tau, Fr zd, uint64_t count, uint32_t* out_canon) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
        Fr t = zd, b = tau;
        for (uint64_t e = i; e; e >>= 1) { if (e & 1) t = t * b; b = b.sqr(); }
        Fr c; t.to_canonical(c.v);
        fr_store(reinterpret_cast<Fr*>(out_canon) + i, c);
    }
}

// ------------------------------------------------------------------------- host algebra
static Fr host_fr_from_bytes(const uint8_t* b32, bool* ok) {
    uint32_t c[8]; memcpy(c, b32, 32);
    if (Fr::geq_p(c)) { *ok = false; return Fr::zero(); }
    return Fr::from_canonical(c);
}
static Fq host_fq_from_bytes(const uint8_t* b48) { uint32_t c[12]; memcpy(c, b48, 48); return Fq::from_canonical(c); }
static void host_fq_to_bytes(const Fq& f, uint8_t* b48) { uint32_t c[12]; f.to_canonical(c); memcpy(b48, c, 48); }
static G1Affine host_g1_from_bytes(const uint8_t* b) { return {host_fq_from_bytes(b), host_fq_from_bytes(b + 48)}; }
static G2Affine host_g2_from_bytes(const uint8_t* b) {
    return {{host_fq_from_bytes(b), host_fq_from_bytes(b + 48)}, {host_fq_from_bytes(b + 96), host_fq_from_bytes(b + 144)}};
}
static void host_g1_to_bytes(const G1Affine& p, uint8_t* b) { host_fq_to_bytes(p.x, b); host_fq_to_bytes(p.y, b + 48); }
static void host_g2_to_bytes(const G2Affine& p, uint8_t* b) {
    host_fq_to_bytes(p.x.c0, b); host_fq_to_bytes(p.x.c1, b + 48); host_fq_to_bytes(p.y.c0, b + 96); host_fq_to_bytes(p.y.c1, b + 144);
}
// (point compression of the 192-byte proof: host_bls.hpp, shared with the verifier)

// ------------------------------------------------------------------------------ key load
static void free_csr(DevCsr& m) { cudaFree(m.row_ptr); cudaFree(m.col); cudaFree(m.val); m = DevCsr(); }
void free_pk(b2g16_pk* pk) {
    if (!pk) return;
    free_csr(pk->A); free_csr(pk->B); free_csr(pk->C);
    cudaFree(pk->a); cudaFree(pk->b_g1); cudaFree(pk->b_g2); cudaFree(pk->l); cudaFree(pk->h);
    delete pk;
}
static void load_csr(Ctx* c, const b2g16_csr& src, uint64_t n_cols, DevCsr& dst, int* d_bad) {
    B2_REQUIRE(src.row_ptr && (src.nnz == 0 || (src.col && src.val)), B2G16_E_INVALID_ARG, "constraint matrix pointers");
    B2_REQUIRE(src.nnz < (1ull << 32), B2G16_E_INVALID_ARG, "constraint matrix: nnz must fit 32 bits");
    B2_REQUIRE(src.row_ptr[0] == 0 && src.row_ptr[src.n_rows] == src.nnz, B2G16_E_INVALID_ARG, "constraint matrix: row_ptr inconsistent with nnz");
    for (uint64_t k = 0; k < src.nnz; ++k) B2_REQUIRE(src.col[k] < n_cols, B2G16_E_INVALID_ARG, "constraint matrix: column index out of range");
    dst.n_rows = src.n_rows; dst.nnz = src.nnz;
    B2_CUDA(cudaMalloc((void**)&dst.row_ptr, (src.n_rows + 1) * sizeof(uint32_t)));
    B2_CUDA(cudaMalloc((void**)&dst.col, (src.nnz + 1) * sizeof(uint32_t)));
    B2_CUDA(cudaMalloc((void**)&dst.val, (src.nnz + 1) * sizeof(Fr)));
    B2_CUDA(cudaMemcpyAsync(dst.row_ptr, src.row_ptr, (src.n_rows + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    if (src.nnz) {
        B2_CUDA(cudaMemcpyAsync(dst.col, src.col, src.nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
        B2_CUDA(cudaMemcpyAsync(dst.val, src.val, src.nnz * 32, cudaMemcpyHostToDevice, c->stream));
        fr_from_bytes_dev(c, reinterpret_cast<const uint8_t*>(dst.val), dst.val, src.nnz, d_bad);      // in place
    }
}
b2g16_pk* new_pk_with_cs(Ctx* c, const b2g16_r1cs_view* cs, int* d_bad) {
    B2_REQUIRE(cs && cs->log_m <= 30, B2G16_E_INVALID_ARG, "constraint system view");
    const uint64_t m = 1ull << cs->log_m;
    B2_REQUIRE(cs->n_cons + cs->n_in + 1 <= m, B2G16_E_INVALID_ARG, "domain too small: need n_cons + n_in + 1 <= 2^log_m");
    // shape-only view (no matrices): the caller will bring a, b, c itself (b2g16_prove_abc)
    const bool shape_only = !cs->A.row_ptr && !cs->B.row_ptr && !cs->C.row_ptr;
    B2_REQUIRE(shape_only || (cs->A.n_rows == cs->n_cons && cs->B.n_rows == cs->n_cons && cs->C.n_rows == cs->n_cons), B2G16_E_INVALID_ARG, "matrix row counts");
    b2g16_pk* pk = new b2g16_pk();
    pk->log_m = cs->log_m; pk->m = m; pk->n_cons = cs->n_cons; pk->n_in = cs->n_in; pk->n_aux = cs->n_aux;
    pk->n_vars = 1 + cs->n_in + cs->n_aux;
    if (shape_only) return pk;
    try {
        load_csr(c, cs->A, pk->n_vars, pk->A, d_bad);
        load_csr(c, cs->B, pk->n_vars, pk->B, d_bad);
        load_csr(c, cs->C, pk->n_vars, pk->C, d_bad);
    } catch (...) { free_pk(pk); throw; }
    return pk;
}
// single elements: validate through the same device path as the queries, keep host copies
void pk_set_single_points(Ctx* c, b2g16_pk* pk, const uint8_t* alpha_g1, const uint8_t* beta_g1, const uint8_t* delta_g1, const uint8_t* beta_g2,
                          const uint8_t* delta_g2, int* d_bad) {
    DevBuf<uint8_t> small(96 * 3 + 192 * 2, c->stream);
    uint8_t hb[96 * 3 + 192 * 2];
    memcpy(hb, alpha_g1, 96); memcpy(hb + 96, beta_g1, 96); memcpy(hb + 192, delta_g1, 96);
    memcpy(hb + 288, beta_g2, 192); memcpy(hb + 480, delta_g2, 192);
    B2_CUDA(cudaMemcpyAsync(small.get(), hb, sizeof(hb), cudaMemcpyHostToDevice, c->stream));
    g1_from_bytes_dev(c, small.get(), (G1Affine*)small.get(), 3, d_bad, true);
    g2_from_bytes_dev(c, small.get() + 288, (G2Affine*)(small.get() + 288), 2, d_bad, true);
    B2_CUDA(cudaStreamSynchronize(c->stream));       // hb is on this frame
    pk->alpha_g1 = host_g1_from_bytes(alpha_g1); pk->beta_g1 = host_g1_from_bytes(beta_g1); pk->delta_g1 = host_g1_from_bytes(delta_g1);
    pk->beta_g2 = host_g2_from_bytes(beta_g2); pk->delta_g2 = host_g2_from_bytes(delta_g2);
}
template <class A> static A* upload_points(Ctx* c, const uint8_t* bytes, size_t n, int* d_bad) {
    A* d = nullptr;
    B2_CUDA(cudaMalloc((void**)&d, (n + 1) * sizeof(A)));
    if (n) {
        // ABI bytes and Montgomery limbs have the same size: convert in place
        B2_CUDA(cudaMemcpyAsync(d, bytes, n * sizeof(A), cudaMemcpyHostToDevice, c->stream));
        if (sizeof(A) == sizeof(G1Affine)) g1_from_bytes_dev(c, (const uint8_t*)d, (G1Affine*)d, n, d_bad, true);
        else g2_from_bytes_dev(c, (const uint8_t*)d, (G2Affine*)d, n, d_bad, true);
    }
    return d;
}

// ---------------------------------------------------------------------------- the prover
struct MsmOut { uint8_t a[96], b1[96], l[96], h[96], b2[192]; };
// window sums of the five multiexps as the device leaves them (msm_*_windows_dev); combined on the host
struct WinBuf { G1Xyzz a[32], b1[32], l[32], h[32]; G2Xyzz b2[32]; };
// sum_w 2^(c w) win[w] by Horner in the host's 64-bit arithmetic.  Device XYZZ accumulators (12 x u32 Montgomery
// limbs per coordinate) and host_bls.hpp's Pt (6 x u64) are the same bytes.
template <class HF, class DX> static b2h::Aff<HF> host_horner(const DX* win, int c, int W) {
    static_assert(sizeof(DX) == sizeof(b2h::Pt<HF>), "device and host accumulator layouts differ");
    if (W == 0) return b2h::Aff<HF>::identity();
    b2h::Pt<HF> acc, t;
    memcpy(&acc, &win[W - 1], sizeof acc);
    for (int w = W - 2; w >= 0; --w) {
        for (int k = 0; k < c; ++k) acc = acc.dbl();
        memcpy(&t, &win[w], sizeof t);
        acc = acc.add(t);
    }
    return acc.affine();
}

// abc_host: optional {a32, b32, c32} host evaluations supplied by the caller's synthesis (b2g16_prove_abc);
// otherwise a, b, c come from the key's CSR matrices.
// the blinding terms of a proof depend on (key, r, s) only: r delta_1, s delta_1, r s delta_1, s delta_2 -- four of the six
// host scalar multiplications of the assembly, computed on a host thread while the device works
struct BlindTerms { b2h::Pt<b2h::Fq> r_d1, s_d1, rs_d1; b2h::Pt<b2h::Fq2> s_d2; };
static BlindTerms blind_terms(const b2g16_pk* pk, const uint8_t* r32, const uint8_t* s32);
static void assemble_proof(const b2g16_pk* pk, const MsmOut& out, const uint8_t* r32, const uint8_t* s32, uint8_t* proof_out, b2g16_trace* tr,
                           const BlindTerms* pre = nullptr);

// part / parts: this call reduces only the contiguous slice [part n / parts, (part+1) n / parts) of each of the
// five queries (one proof spread over several GPUs, b2g16_prove_partial); partial_out then receives the five
// partial sums and no proof is assembled.
static void prove_one(Ctx* c, const b2g16_pk* pk, const uint8_t* d_z_canon, const uint8_t* r32, const uint8_t* s32,
                      uint8_t* proof_out, b2g16_trace* tr, const uint8_t* const* abc_host = nullptr,
                      unsigned part = 0, unsigned parts = 1, MsmOut* partial_out = nullptr,
                      const uint8_t* d_h_slice = nullptr, uint64_t h_first = 0, uint64_t h_count = 0) {
    std::future<BlindTerms> blind;
    if (!partial_out) {
        bool ok = true;
        host_fr_from_bytes(r32, &ok); host_fr_from_bytes(s32, &ok);
        B2_REQUIRE(ok, B2G16_E_NOT_CANONICAL, "r / s not canonical");
        blind = std::async(std::launch::async, [pk, r32, s32] { return blind_terms(pk, r32, s32); });   // joined in assemble below (or by the future's destructor on unwind)
    }
    auto lo_of = [&](uint64_t n) { return (size_t)(n * part / parts); };
    auto cnt_of = [&](uint64_t n) { return (size_t)(n * (part + 1) / parts - n * part / parts); };
    const uint64_t m = pk->m, nv = pk->n_vars;
    cudaStream_t st = c->stream;
    cudaEvent_t* ev = c->ev_stage;                       // created once per context
    DevFlag bad(c);
    // d_h_slice: H coefficients [h_first, h_first + h_count) computed elsewhere (the sharded witness map): no
    // constraint evaluation and no transforms here, and this rank's share of the H query is exactly that range
    const bool ext_h = d_h_slice != nullptr;
    DevBuf<Fr> z(ext_h ? 0 : nv, st), a(ext_h ? 0 : m, st), b(ext_h ? 0 : m, st), cc(ext_h ? 0 : m, st);
    DevBuf<uint32_t> h(ext_h ? 0 : m * 8, st);
    DevBuf<uint8_t> d_win(sizeof(WinBuf), st);
    MsmOut out;
    static thread_local WinBuf win;                      // 36 KB: not on the stack
    int cb[5] = {0, 0, 0, 0, 0}, nw[5] = {0, 0, 0, 0, 0};
    B2_CUDA(cudaEventRecord(ev[0], st));
    if (ext_h) scalars_check_dev(c, d_z_canon, nv, bad.ptr());
    else fr_from_bytes_dev(c, d_z_canon, z.get(), nv, bad.ptr());
    B2_CUDA(cudaEventRecord(ev[1], st));
    if (ext_h) {
    } else if (abc_host) {
        Fr* dst[3] = {a.get(), b.get(), cc.get()};
        for (int i = 0; i < 3; ++i) {                                 // same size: upload, then convert in place
            B2_CUDA(cudaMemcpyAsync(dst[i], abc_host[i], m * 32, cudaMemcpyHostToDevice, st));
            fr_from_bytes_dev(c, reinterpret_cast<const uint8_t*>(dst[i]), dst[i], m, bad.ptr());
        }
    } else {
        B2_REQUIRE(pk->A.row_ptr && pk->B.row_ptr && pk->C.row_ptr, B2G16_E_INVALID_ARG, "prove: key was loaded without constraint matrices, use b2g16_prove_abc");
        abc_spmv_kernel<<<grid_for(m, 256), 256, 0, st>>>(pk->A, pk->B, pk->C, z.get(), pk->n_cons, pk->n_in, m, a.get(), b.get(), cc.get());
        c->launches++;
        B2_CUDA(cudaGetLastError());
    }
    if (tr && (tr->a || tr->b || tr->c)) {
        DevBuf<uint8_t> tmp(m * 32, st);
        uint8_t* dst[3] = {tr->a, tr->b, tr->c};
        Fr* src[3] = {a.get(), b.get(), cc.get()};
        for (int i = 0; i < 3; ++i) if (dst[i]) {
            fr_to_bytes_dev(c, src[i], tmp.get(), m);
            B2_CUDA(cudaMemcpyAsync(dst[i], tmp.get(), m * 32, cudaMemcpyDeviceToHost, st));
            B2_CUDA(cudaStreamSynchronize(st));
        }
    }
    B2_CUDA(cudaEventRecord(ev[2], st));
    // witness MSMs do not depend on H: fork them onto the aux streams while the NTT chain runs.  From here to the join,
    // work of this proof may be running on the aux streams against buffers that are freed stream-ordered on `st`
    // (d_win, z, the MSM scratch): if anything throws in between (e.g. an out-of-memory in a later multiexp's scratch),
    // wait for every stream before the buffers' destructors run, or a stale kernel could write into memory the pool has
    // already handed to the next proof.
    struct JoinOnUnwind {
        Ctx* c; bool armed = true;
        ~JoinOnUnwind() { if (armed) { for (auto s : c->aux) cudaStreamSynchronize(s); cudaStreamSynchronize(c->stream); } }
    } unwind{c};
    B2_CUDA(cudaEventRecord(c->ev_fork, st));
    for (int i = 0; i < 4; ++i) B2_CUDA(cudaStreamWaitEvent(c->aux[i], c->ev_fork, 0));
    const uint32_t* zc = reinterpret_cast<const uint32_t*>(d_z_canon);
    WinBuf* dw = reinterpret_cast<WinBuf*>(d_win.get());
    // above 2^24 the witness MSMs run one after the other on the main stream: each needs ~9 GB of sort
    // scratch at 2^27 and the integer pipe serialises them anyway; below, they overlap the NTT chain
    const bool serial = (int)pk->log_m > c->serial_above_log;
    cudaStream_t s_b2 = serial ? st : c->aux[0], s_a = serial ? st : c->aux[1], s_b1 = serial ? st : c->aux[2], s_l = serial ? st : c->aux[3];
    msm_g2_windows_dev(c, pk->b_g2 + lo_of(nv), zc + 8 * lo_of(nv), cnt_of(nv), dw->b2, &cb[4], &nw[4], s_b2);
    msm_g1_windows_dev(c, pk->a + lo_of(nv), zc + 8 * lo_of(nv), cnt_of(nv), dw->a, &cb[0], &nw[0], s_a);
    msm_g1_windows_dev(c, pk->b_g1 + lo_of(nv), zc + 8 * lo_of(nv), cnt_of(nv), dw->b1, &cb[1], &nw[1], s_b1);
    msm_g1_windows_dev(c, pk->l + lo_of(pk->n_aux), zc + 8 * (1 + pk->n_in + lo_of(pk->n_aux)), cnt_of(pk->n_aux), dw->l, &cb[2], &nw[2], s_l);
    if (!ext_h) h_from_abc_dev(c, a.get(), b.get(), cc.get(), (int)pk->log_m, h.get(), false, nullptr);
    B2_CUDA(cudaEventRecord(ev[3], st));
    if (ext_h) {
        const uint64_t hi = h_first + h_count < m - 1 ? h_first + h_count : m - 1;       // h_{m-1} = 0 has no base
        const uint64_t cnt = hi > h_first ? hi - h_first : 0;
        msm_g1_windows_dev(c, pk->h + h_first, reinterpret_cast<const uint32_t*>(d_h_slice), cnt, dw->h, &cb[3], &nw[3], st);
    } else
        msm_g1_windows_dev(c, pk->h + lo_of(m - 1), h.get() + 8 * lo_of(m - 1), cnt_of(m - 1), dw->h, &cb[3], &nw[3], st);
    for (int i = 0; i < 4; ++i) { B2_CUDA(cudaEventRecord(c->ev_join[i], c->aux[i])); B2_CUDA(cudaStreamWaitEvent(st, c->ev_join[i], 0)); }
    B2_CUDA(cudaMemcpyAsync(&win, dw, sizeof(WinBuf), cudaMemcpyDeviceToHost, st));
    if (tr && tr->h && !ext_h) B2_CUDA(cudaMemcpyAsync(tr->h, h.get(), m * 32, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaEventRecord(ev[4], st));
    bad.check("assignment");                              // synchronises the stream (which has joined the aux streams)
    unwind.armed = false;
    if (tr) {
        for (int i = 0; i < 8; ++i) tr->stage_ms[i] = 0.f;
        for (int i = 0; i < 4; ++i) cudaEventElapsedTime(&tr->stage_ms[i], ev[i], ev[i + 1]);   // ingest, abc, ntt/h, msm_h + join of the rest
    }
    // Horner over the windows of each multiexp, on the host (the stream is idle: bad.check synchronised it);
    // the five combinations are independent: one thread each, the G2 one (3x the work) on this thread
    {
        const WinBuf* w = &win;
        MsmOut* op = &out;
        std::thread th[4] = {
            std::thread([=] { b2h::g1_to_abi(host_horner<b2h::Fq>(w->a, cb[0], nw[0]), op->a); }),
            std::thread([=] { b2h::g1_to_abi(host_horner<b2h::Fq>(w->b1, cb[1], nw[1]), op->b1); }),
            std::thread([=] { b2h::g1_to_abi(host_horner<b2h::Fq>(w->l, cb[2], nw[2]), op->l); }),
            std::thread([=] { b2h::g1_to_abi(host_horner<b2h::Fq>(w->h, cb[3], nw[3]), op->h); })};
        b2h::g2_to_abi(host_horner<b2h::Fq2>(w->b2, cb[4], nw[4]), op->b2);
        for (auto& t : th) t.join();
    }
    if (partial_out) { *partial_out = out; return; }
    const BlindTerms pre = blind.get();
    assemble_proof(pk, out, r32, s32, proof_out, tr, &pre);
}

// host assembly from the five multiexp results
// Montgomery limbs of the device types (12 x u32, R = 2^384) and of host_bls.hpp (6 x u64) are the same bytes
static b2h::G1 to_host(const G1Affine& p) { b2h::G1 o; o.inf = p.is_identity(); memcpy(o.x.v, p.x.v, 48); memcpy(o.y.v, p.y.v, 48); return o; }
static b2h::G2 to_host(const G2Affine& p) {
    b2h::G2 o; o.inf = p.is_identity();
    memcpy(o.x.c0.v, p.x.c0.v, 48); memcpy(o.x.c1.v, p.x.c1.v, 48); memcpy(o.y.c0.v, p.y.c0.v, 48); memcpy(o.y.c1.v, p.y.c1.v, 48);
    return o;
}
// The ten scalar multiplications and three inversions of the assembly run in the 64-bit host arithmetic of
// host_bls.hpp (~0.1 us per Fq product; the limb-by-limb emulation of the device multiplier that the unit tests
// use costs 0.6 us, which made this step 18 ms of every proof).
static BlindTerms blind_terms(const b2g16_pk* pk, const uint8_t* r32, const uint8_t* s32) {
    bool ok = true;
    const Fr r = host_fr_from_bytes(r32, &ok), s = host_fr_from_bytes(s32, &ok);
    B2_REQUIRE(ok, B2G16_E_NOT_CANONICAL, "r / s not canonical");
    uint8_t rs32[32];
    { uint32_t c[8]; (r * s).to_canonical(c); memcpy(rs32, c, 32); }
    const b2h::G1 d1 = to_host(pk->delta_g1);
    BlindTerms t;
    t.r_d1 = b2h::scalar_mul(d1, r32); t.s_d1 = b2h::scalar_mul(d1, s32); t.rs_d1 = b2h::scalar_mul(d1, rs32);
    t.s_d2 = b2h::scalar_mul(to_host(pk->delta_g2), s32);
    return t;
}
static void assemble_proof(const b2g16_pk* pk, const MsmOut& out, const uint8_t* r32, const uint8_t* s32, uint8_t* proof_out, b2g16_trace* tr,
                           const BlindTerms* pre) {
    BlindTerms own;
    if (!pre) { own = blind_terms(pk, r32, s32); pre = &own; }
    typedef b2h::Pt<b2h::Fq> P1;
    typedef b2h::Pt<b2h::Fq2> P2;
    b2h::G1 ha, hb1, hl, hh;
    b2h::G2 hb2;
    B2_REQUIRE(b2h::g1_from_abi(out.a, &ha) && b2h::g1_from_abi(out.b1, &hb1) && b2h::g1_from_abi(out.l, &hl) && b2h::g1_from_abi(out.h, &hh) &&
               b2h::g2_from_abi(out.b2, &hb2), B2G16_E_INTERNAL, "prove: multiexp result not canonical");
    const P1 A = P1::from(ha).add(P1::from(to_host(pk->alpha_g1))).add(pre->r_d1);
    const P2 Bq = P2::from(hb2).add(P2::from(to_host(pk->beta_g2))).add(pre->s_d2);
    const P1 B1 = P1::from(hb1).add(P1::from(to_host(pk->beta_g1))).add(pre->s_d1);
    const b2h::G1 Aa = A.affine(), B1a = B1.affine();
    const P1 Cc = P1::from(hh).add(P1::from(hl)).add(b2h::scalar_mul(Aa, s32)).add(b2h::scalar_mul(B1a, r32)).add(P1::from(pre->rs_d1.affine().neg()));
    const b2h::G1 Ca = Cc.affine();
    const b2h::G2 Ba = Bq.affine();
    b2h::g1_compress(Aa, proof_out); b2h::g2_compress(Ba, proof_out + 48); b2h::g1_compress(Ca, proof_out + 144);
    if (tr) {
        if (tr->msm_a) memcpy(tr->msm_a, out.a, 96);
        if (tr->msm_b1) memcpy(tr->msm_b1, out.b1, 96);
        if (tr->msm_l) memcpy(tr->msm_l, out.l, 96);
        if (tr->msm_h) memcpy(tr->msm_h, out.h, 96);
        if (tr->msm_b2) memcpy(tr->msm_b2, out.b2, 192);
        if (tr->A) b2h::g1_to_abi(Aa, tr->A);
        if (tr->B) b2h::g2_to_abi(Ba, tr->B);
        if (tr->C) b2h::g1_to_abi(Ca, tr->C);
    }
}

}  // namespace b2

namespace b2 {
// Groth16 setup on the device for a key whose constraint system is already resident (pk->A/B/C).
static void setup_on_device(Ctx* c, b2g16_pk* pk, const uint8_t* toxic, uint8_t* vk_out, DevFlag& bad) {
        bool ok = true;
        Fr tx[5];
        for (int i = 0; i < 5; ++i) tx[i] = host_fr_from_bytes(toxic + 32 * i, &ok);
        B2_REQUIRE(ok, B2G16_E_NOT_CANONICAL, "toxic waste not canonical");
        const Fr tau = tx[0], alpha = tx[1], beta = tx[2], gamma = tx[3], delta = tx[4];
        B2_REQUIRE(!gamma.is_zero() && !delta.is_zero(), B2G16_E_INVALID_ARG, "gamma and delta must be non-zero");
        cudaStream_t st = c->stream;
        {
            const uint64_t m = pk->m, nv = pk->n_vars;
            // Lagrange basis at tau
            Fr w = host_root_of_unity((int)pk->log_m);
            Fr tm = tau; for (uint32_t i = 0; i < pk->log_m; ++i) tm = tm.sqr();
            const Fr zt = tm - Fr::one();
            uint32_t mc[8] = {(uint32_t)m, (uint32_t)(m >> 32), 0, 0, 0, 0, 0, 0};
            const Fr pref = zt * Fr::from_canonical(mc).inv();
            DevBuf<Fr> lag(m, st), u(nv, st), v(nv, st), ww(nv, st);
            lagrange_kernel<<<grid_for(m, 128), 128, 0, st>>>(lag.get(), tau, pref, w, m); c->launches++;
            // u, v, w = A^T lag, B^T lag, C^T lag via a column grouping (counting sort by column)
            const DevCsr* mats[3] = {&pk->A, &pk->B, &pk->C};
            Fr* outs[3] = {u.get(), v.get(), ww.get()};
            const int n_tiles = (int)((nv + SCAN_TILE - 1) / SCAN_TILE);
            for (int t = 0; t < 3; ++t) {
                const DevCsr& M = *mats[t];
                DevBuf<uint32_t> hist(nv + 1, st), off(nv + 1, st), cursor(nv + 1, st), tile_sums((size_t)n_tiles + 1, st), ent_row(M.nnz + 1, st), ent_k(M.nnz + 1, st);
                B2_CUDA(cudaMemsetAsync(hist.get(), 0, (nv + 1) * sizeof(uint32_t), st));
                if (M.nnz) { csc_hist_kernel<<<grid_for(M.nnz, 256), 256, 0, st>>>(M.col, M.nnz, hist.get()); c->launches++; }
                scan_u32(c, st, hist.get(), off.get(), nv, 0, 1, tile_sums.get());
                B2_CUDA(cudaMemcpyAsync(cursor.get(), off.get(), (nv + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
                if (M.nnz) { csc_scatter_kernel<<<grid_for(M.n_rows, 256), 256, 0, st>>>(M.row_ptr, M.col, M.n_rows, cursor.get(), ent_row.get(), ent_k.get()); c->launches++; }
                qap_column_kernel<<<grid_for(nv, 128), 128, 0, st>>>(off.get(), ent_row.get(), ent_k.get(), M.val, lag.get(), nv, pk->n_cons, pk->n_in, t == 0, outs[t]); c->launches++;
                B2_CUDA(cudaGetLastError());
            }
            // scalars of every query (canonical), then fixed-base multiplications
            DevBuf<uint32_t> su(nv * 8, st), sv(nv * 8, st), sk(nv * 8, st), sh((m ? m : 1) * 8, st);
            fr_to_bytes_dev(c, u.get(), (uint8_t*)su.get(), nv);
            fr_to_bytes_dev(c, v.get(), (uint8_t*)sv.get(), nv);
            const Fr ginv = gamma.inv(), dinv = delta.inv();
            setup_k_kernel<<<grid_for(nv, 128), 128, 0, st>>>(u.get(), v.get(), ww.get(), alpha, beta, ginv, dinv, nv, pk->n_in, sk.get()); c->launches++;
            if (m > 1) { setup_h_kernel<<<grid_for(m - 1, 128), 128, 0, st>>>(tau, zt * dinv, m - 1, sh.get()); c->launches++; }
            B2_CUDA(cudaGetLastError());
            B2_CUDA(cudaMalloc((void**)&pk->a, (nv + 1) * sizeof(G1Affine)));
            B2_CUDA(cudaMalloc((void**)&pk->b_g1, (nv + 1) * sizeof(G1Affine)));
            B2_CUDA(cudaMalloc((void**)&pk->b_g2, (nv + 1) * sizeof(G2Affine)));
            B2_CUDA(cudaMalloc((void**)&pk->l, (pk->n_aux + 1) * sizeof(G1Affine)));
            B2_CUDA(cudaMalloc((void**)&pk->h, m * sizeof(G1Affine)));
            gen_mul_g1_dev(c, su.get(), nv, pk->a);
            gen_mul_g1_dev(c, sv.get(), nv, pk->b_g1);
            gen_mul_g2_dev(c, sv.get(), nv, pk->b_g2);
            gen_mul_g1_dev(c, sk.get() + 8 * (1 + pk->n_in), pk->n_aux, pk->l);
            gen_mul_g1_dev(c, sh.get(), m - 1, pk->h);
            // single elements and the verifying key: alpha, beta, delta in G1; beta, gamma, delta in G2; IC
            const size_t n_ic = pk->n_in + 1;
            DevBuf<uint32_t> s1(8 * 8, st);
            uint32_t hs[6][8];
            alpha.to_canonical(hs[0]); beta.to_canonical(hs[1]); delta.to_canonical(hs[2]); gamma.to_canonical(hs[3]);
            B2_CUDA(cudaMemcpyAsync(s1.get(), hs, sizeof(uint32_t) * 32, cudaMemcpyHostToDevice, st));
            DevBuf<G1Affine> e1(3, st), ic(n_ic, st);
            DevBuf<G2Affine> e2(3, st);
            gen_mul_g1_dev(c, s1.get(), 3, e1.get());                 // alpha, beta, delta
            gen_mul_g2_dev(c, s1.get() + 8, 3, e2.get());             // beta, delta, gamma
            gen_mul_g1_dev(c, sk.get(), n_ic, ic.get());
            G1Affine h1[3]; G2Affine h2[3];
            B2_CUDA(cudaMemcpyAsync(h1, e1.get(), sizeof(h1), cudaMemcpyDeviceToHost, st));
            B2_CUDA(cudaMemcpyAsync(h2, e2.get(), sizeof(h2), cudaMemcpyDeviceToHost, st));
            std::vector<uint8_t> icb(96 * n_ic);
            if (vk_out) {
                DevBuf<uint8_t> raw(96 * n_ic, st);
                g1_to_bytes_dev(c, ic.get(), raw.get(), n_ic, st);
                B2_CUDA(cudaMemcpyAsync(icb.data(), raw.get(), 96 * n_ic, cudaMemcpyDeviceToHost, st));
                B2_CUDA(cudaStreamSynchronize(st));
            }
            bad.check("constraint system");
            pk->alpha_g1 = h1[0]; pk->beta_g1 = h1[1]; pk->delta_g1 = h1[2];
            pk->beta_g2 = h2[0]; pk->delta_g2 = h2[1];
            if (vk_out) {
                host_g1_to_bytes(h1[0], vk_out); host_g2_to_bytes(h2[0], vk_out + 96);
                host_g2_to_bytes(h2[2], vk_out + 288); host_g2_to_bytes(h2[1], vk_out + 480);
                memcpy(vk_out + 672, icb.data(), 96 * n_ic);
            }
        }
}
}  // namespace b2

using namespace b2;

extern "C" {

int b2g16_pk_load(b2g16_ctx* hctx, const b2g16_pk_view* v, b2g16_pk** out) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !v || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        B2_REQUIRE(v->alpha_g1 && v->beta_g1 && v->delta_g1 && v->beta_g2 && v->delta_g2 && v->a && v->b_g1 && v->b_g2 && v->h && v->cs,
                   B2G16_E_INVALID_ARG, "proving key view has null members");
        DevFlag bad(c);
        b2g16_pk* pk = new_pk_with_cs(c, v->cs, bad.ptr());
        try {
            B2_REQUIRE(pk->n_aux == 0 || v->l, B2G16_E_INVALID_ARG, "l query missing");
            pk->a = upload_points<G1Affine>(c, v->a, pk->n_vars, bad.ptr());
            pk->b_g1 = upload_points<G1Affine>(c, v->b_g1, pk->n_vars, bad.ptr());
            pk->b_g2 = upload_points<G2Affine>(c, v->b_g2, pk->n_vars, bad.ptr());
            pk->l = upload_points<G1Affine>(c, v->l, pk->n_aux, bad.ptr());
            pk->h = upload_points<G1Affine>(c, v->h, pk->m - 1, bad.ptr());
            pk_set_single_points(c, pk, v->alpha_g1, v->beta_g1, v->delta_g1, v->beta_g2, v->delta_g2, bad.ptr());
            bad.check("proving key");
        } catch (...) { cudaStreamSynchronize(c->stream); free_pk(pk); throw; }
        *out = pk;
    });
}

int b2g16_pk_free(b2g16_ctx* hctx, b2g16_pk* pk) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { B2_CUDA(cudaStreamSynchronize(c->stream)); free_pk(pk); });
}

int b2g16_pk_export(b2g16_ctx* hctx, const b2g16_pk* pk, int which, uint8_t* out, size_t out_bytes) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !out || which < 0 || which > 9) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        if (which >= 5) {
            const size_t need = which >= 8 ? 192 : 96;
            B2_REQUIRE(out_bytes >= need, B2G16_E_INVALID_ARG, "output buffer too small");
            switch (which) {
                case 5: host_g1_to_bytes(pk->alpha_g1, out); break;
                case 6: host_g1_to_bytes(pk->beta_g1, out); break;
                case 7: host_g1_to_bytes(pk->delta_g1, out); break;
                case 8: host_g2_to_bytes(pk->beta_g2, out); break;
                default: host_g2_to_bytes(pk->delta_g2, out);
            }
            return;
        }
        const size_t counts[5] = {pk->n_vars, pk->n_vars, pk->n_vars, pk->n_aux, pk->m - 1};
        const size_t enc = which == 2 ? 192 : 96, n = counts[which];
        B2_REQUIRE(out_bytes >= n * enc, B2G16_E_INVALID_ARG, "output buffer too small");
        if (!n) return;
        DevBuf<uint8_t> raw(n * enc, c->stream);
        switch (which) {
            case 0: g1_to_bytes_dev(c, pk->a, raw.get(), n, c->stream); break;
            case 1: g1_to_bytes_dev(c, pk->b_g1, raw.get(), n, c->stream); break;
            case 2: g2_to_bytes_dev(c, pk->b_g2, raw.get(), n, c->stream); break;
            case 3: g1_to_bytes_dev(c, pk->l, raw.get(), n, c->stream); break;
            default: g1_to_bytes_dev(c, pk->h, raw.get(), n, c->stream);
        }
        B2_CUDA(cudaMemcpyAsync(out, raw.get(), n * enc, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
}

int b2g16_prove_device(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* d_assignment32, const uint8_t r32[32], const uint8_t s32[32],
                       int priority, uint8_t proof_out[192], b2g16_trace* trace) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !d_assignment32 || !r32 || !s32 || !proof_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { prove_one(c, pk, d_assignment32, r32, s32, proof_out, trace); }, priority != 0);
}

int b2g16_prove(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* assignment32, const uint8_t r32[32], const uint8_t s32[32],
                int priority, uint8_t proof_out[192], b2g16_trace* trace) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !assignment32 || !r32 || !s32 || !proof_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevBuf<uint8_t> dz(pk->n_vars * 32, c->stream);
        B2_CUDA(cudaMemcpyAsync(dz.get(), assignment32, pk->n_vars * 32, cudaMemcpyHostToDevice, c->stream));
        prove_one(c, pk, dz.get(), r32, s32, proof_out, trace);
    }, priority != 0);
}

int b2g16_prove_abc(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* a32, const uint8_t* b32, const uint8_t* c32, const uint8_t* assignment32,
                    const uint8_t r32[32], const uint8_t s32[32], int priority, uint8_t proof_out[192], b2g16_trace* trace) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !a32 || !b32 || !c32 || !assignment32 || !r32 || !s32 || !proof_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevBuf<uint8_t> dz(pk->n_vars * 32, c->stream);
        B2_CUDA(cudaMemcpyAsync(dz.get(), assignment32, pk->n_vars * 32, cudaMemcpyHostToDevice, c->stream));
        const uint8_t* abc[3] = {a32, b32, c32};
        prove_one(c, pk, dz.get(), r32, s32, proof_out, trace, abc);
    }, priority != 0);
}

// One proof over `parts` GPUs (latency mode): each rank, holding the whole key and witness,
// runs the witness map and reduces its slice of the five queries; the host gathers parts x 576 bytes
// (msm_a 96, msm_b1 96, msm_l 96, msm_h 96, msm_b2 192) and any rank assembles.
int b2g16_prove_partial(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* d_assignment32, unsigned part, unsigned parts, uint8_t partials_out[576]) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !d_assignment32 || !partials_out || parts == 0 || part >= parts) return B2G16_E_INVALID_ARG;
    static_assert(sizeof(MsmOut) == 576, "partial layout");
    return guard(c, [&] {
        MsmOut out;
        prove_one(c, pk, d_assignment32, nullptr, nullptr, nullptr, nullptr, nullptr, part, parts, &out);
        memcpy(partials_out, &out, sizeof out);
    });
}
// Pieces of a proof whose witness map is spread over the ranks as well (b2g16.dist.prove_sharded_full):
//   abc_evals_strided: this rank's rows row0 + stride k of a, b, c as canonical scalars
//   h_pointwise:       (a b - c) / Z on the coset, elementwise on canonical scalars
//   prove_partial_h:   b2g16_prove_partial with the H coefficients [h_first, h_first + h_count) supplied by the caller
int b2g16_abc_evals_strided_device(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* d_assignment32, uint64_t row0, uint64_t stride, uint64_t count,
                                   uint8_t* d_a32, uint8_t* d_b32, uint8_t* d_c32) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !d_assignment32 || !d_a32 || !d_b32 || !d_c32 || stride == 0) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        B2_REQUIRE(pk->A.row_ptr && pk->B.row_ptr && pk->C.row_ptr, B2G16_E_INVALID_ARG, "abc_evals: key was loaded without constraint matrices");
        B2_REQUIRE(count == 0 || row0 + stride * (count - 1) < pk->m, B2G16_E_INVALID_ARG, "abc_evals: rows outside the domain");
        DevFlag bad(c);
        DevBuf<Fr> z(pk->n_vars, c->stream);
        fr_from_bytes_dev(c, d_assignment32, z.get(), pk->n_vars, bad.ptr());
        abc_spmv_strided_kernel<<<grid_for(count, 256), 256, 0, c->stream>>>(pk->A, pk->B, pk->C, z.get(), pk->n_cons, pk->n_in, row0, stride, count,
                                                                              reinterpret_cast<Fr*>(d_a32), reinterpret_cast<Fr*>(d_b32), reinterpret_cast<Fr*>(d_c32));
        c->launches++;
        B2_CUDA(cudaGetLastError());
        bad.check("assignment");
    });
}
int b2g16_h_pointwise_device(b2g16_ctx* hctx, const uint8_t* d_a32, const uint8_t* d_b32, const uint8_t* d_c32, uint64_t count, unsigned log_m, uint8_t* d_out32) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !d_a32 || !d_b32 || !d_c32 || !d_out32 || log_m > 30) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        Fr g = Fr::one(), seven = Fr::zero();
        for (int i = 0; i < 7; ++i) seven = seven + Fr::one();
        g = seven;
        for (unsigned i = 0; i < log_m; ++i) g = g.sqr();                                   // g^(2^log_m)
        const Fr zinv = (g - Fr::one()).inv();
        h_pointwise_kernel<<<grid_for(count, 256), 256, 0, c->stream>>>(reinterpret_cast<const Fr*>(d_a32), reinterpret_cast<const Fr*>(d_b32),
                                                                         reinterpret_cast<const Fr*>(d_c32), count, zinv, reinterpret_cast<Fr*>(d_out32));
        c->launches++;
        B2_CUDA(cudaGetLastError());
    });
}
int b2g16_prove_partial_h(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* d_assignment32, unsigned part, unsigned parts, const uint8_t* d_h_slice32,
                          uint64_t h_first, uint64_t h_count, uint8_t partials_out[576]) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !d_assignment32 || !partials_out || !d_h_slice32 || parts == 0 || part >= parts) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        B2_REQUIRE(h_first + h_count <= pk->m, B2G16_E_INVALID_ARG, "prove_partial_h: H range outside the domain");
        DevFlag bad(c);
        scalars_check_dev(c, d_h_slice32, h_count, bad.ptr());
        bad.check("H coefficients");
        MsmOut out;
        prove_one(c, pk, d_assignment32, nullptr, nullptr, nullptr, nullptr, nullptr, part, parts, &out, d_h_slice32, h_first, h_count);
        memcpy(partials_out, &out, sizeof out);
    });
}
int b2g16_prove_assemble(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* partials, unsigned parts, const uint8_t r32[32], const uint8_t s32[32],
                         uint8_t proof_out[192]) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !partials || parts == 0 || !r32 || !s32 || !proof_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        G1Xyzz g1[4] = {G1Xyzz::identity(), G1Xyzz::identity(), G1Xyzz::identity(), G1Xyzz::identity()};
        G2Xyzz g2 = G2Xyzz::identity();
        for (unsigned k = 0; k < parts; ++k) {
            const uint8_t* p = partials + (size_t)k * sizeof(MsmOut);
            for (int q = 0; q < 4; ++q) g1[q].add_mixed(host_g1_from_bytes(p + 96 * q));
            g2.add_mixed(host_g2_from_bytes(p + 384));
        }
        MsmOut sum;
        host_g1_to_bytes(g1[0].to_affine(), sum.a); host_g1_to_bytes(g1[1].to_affine(), sum.b1);
        host_g1_to_bytes(g1[2].to_affine(), sum.l); host_g1_to_bytes(g1[3].to_affine(), sum.h);
        host_g2_to_bytes(g2.to_affine(), sum.b2);
        assemble_proof(pk, sum, r32, s32, proof_out, nullptr);
    });
}

int b2g16_prove_batch(b2g16_ctx* hctx, const b2g16_pk* pk, const uint8_t* assignments32, size_t n_partitions,
                      const uint8_t* r32s, const uint8_t* s32s, int priority, uint8_t* proofs_out) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || (n_partitions && (!assignments32 || !r32s || !s32s || !proofs_out))) return B2G16_E_INVALID_ARG;
    if (n_partitions == 0) return B2G16_OK;
    const size_t zb = pk->n_vars * 32;
    // Two witness buffers and a copy stream of their own: partition k + 1 is uploaded by a helper thread while partition k
    // is proved (a caller's witnesses are ordinary pageable memory, whose upload is a host-driven staged copy: 10-15 ms of
    // a 150 ms proof at 2^22 when it sits in front of every proof).  The buffers are plain cudaMalloc memory, so the copy
    // stream and the proving stream need no allocator hand-over; ordering is by the host (the upload is synchronised
    // before the proof that reads it is launched, a proof is complete before its buffer is written again).
    struct Bufs {
        int device; uint8_t* dz[2] = {nullptr, nullptr}; cudaStream_t copy = nullptr;
        ~Bufs() { cudaSetDevice(device); for (auto p : dz) if (p) cudaFree(p); if (copy) cudaStreamDestroy(copy); }
    } bufs{c->device};
    int rc = guard(c, [&] {
        for (auto& p : bufs.dz) B2_CUDA(cudaMalloc((void**)&p, zb));
        B2_CUDA(cudaStreamCreateWithFlags(&bufs.copy, cudaStreamNonBlocking));
    }, priority != 0);
    if (rc != B2G16_OK) return rc;
    const int device = c->device;
    auto upload = [&bufs, assignments32, zb, device](size_t k) -> cudaError_t {
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess) e = cudaMemcpyAsync(bufs.dz[k & 1], assignments32 + k * zb, zb, cudaMemcpyHostToDevice, bufs.copy);
        if (e == cudaSuccess) e = cudaStreamSynchronize(bufs.copy);
        return e;
    };
    cudaError_t up = upload(0);
    // one admission per partition: a priority proof submitted from another thread (a PoSt proof while a sector's
    // partitions are being proved, compound_proof.hpp:134) gets the device between two partitions of the batch
    for (size_t k = 0; k < n_partitions; ++k) {
        if (up != cudaSuccess) {
            cudaGetLastError();
            guard(c, [&] { throw CudaError{up, "witness upload (b2g16_prove_batch)", __FILE__, __LINE__}; });
            return B2G16_E_CUDA;
        }
        std::future<cudaError_t> next;
        if (k + 1 < n_partitions) next = std::async(std::launch::async, upload, k + 1);
        rc = guard(c, [&] { prove_one(c, pk, bufs.dz[k & 1], r32s + 32 * k, s32s + 32 * k, proofs_out + 192 * k, nullptr); }, priority != 0);
        if (next.valid()) up = next.get();                 // always joined, also when the proof failed
        if (rc != B2G16_OK) return rc;
    }
    return B2G16_OK;
}

int b2g16_synth_witness(b2g16_ctx* hctx, const b2g16_pk* pk, uint8_t* d_assignment32, uint64_t n_free) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !pk || !d_assignment32) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        const uint64_t base = 1 + pk->n_in + n_free;
        B2_REQUIRE(base + pk->n_cons == pk->n_vars, B2G16_E_INVALID_ARG, "synthetic layout: n_aux must equal n_free + n_cons");
        DevFlag bad(c);
        DevBuf<Fr> z(pk->n_vars, c->stream);
        B2_CUDA(cudaMemsetAsync(z.get(), 0, pk->n_vars * sizeof(Fr), c->stream));
        fr_from_bytes_dev(c, d_assignment32, z.get(), base, bad.ptr());
        uint64_t row = 0;
        while (row < pk->n_cons) {                       // same layering as python/b2g16/synth.py layers()
            uint64_t visible = base + row;
            uint64_t size = visible / 2 > 1 ? visible / 2 : 1;
            if (size > pk->n_cons - row) size = pk->n_cons - row;
            synth_layer_kernel<<<grid_for(size, 256), 256, 0, c->stream>>>(pk->A, pk->B, z.get(), row, row + size, base);
            c->launches++;
            row += size;
        }
        B2_CUDA(cudaGetLastError());
        fr_to_bytes_dev(c, z.get(), d_assignment32, pk->n_vars);
        bad.check("assignment head");
    });
}

int b2g16_pk_generate(b2g16_ctx* hctx, const b2g16_r1cs_view* cs, const uint8_t toxic[160], b2g16_pk** out, uint8_t* vk_out) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !cs || !toxic || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        DevFlag bad(c);
        b2g16_pk* pk = new_pk_with_cs(c, cs, bad.ptr());
        try { setup_on_device(c, pk, toxic, vk_out, bad); }
        catch (...) { cudaStreamSynchronize(c->stream); free_pk(pk); throw; }
        *out = pk;
    });
}

int b2g16_pk_generate_synth(b2g16_ctx* hctx, uint32_t log_m, uint64_t n_in, uint64_t n_free, uint64_t seed, const uint8_t toxic[160],
                            b2g16_pk** out, uint8_t* vk_out) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !toxic || !out || log_m > 28) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        const uint64_t m = 1ull << log_m;
        B2_REQUIRE(n_in + 2 <= m, B2G16_E_INVALID_ARG, "domain too small for the public inputs");
        b2g16_pk* pk = new b2g16_pk();
        try {
            pk->log_m = log_m; pk->m = m; pk->n_in = n_in; pk->n_cons = m - n_in - 1;      // domain exactly 2^log_m
            pk->n_aux = n_free + pk->n_cons; pk->n_vars = 1 + n_in + pk->n_aux;
            B2_REQUIRE(3 * pk->n_cons < (1ull << 32), B2G16_E_INVALID_ARG, "non-zero count must fit 32 bits");
            SynthLayers L; L.n = 0;
            const uint64_t base = 1 + n_in + n_free;
            for (uint64_t row = 0; row < pk->n_cons;) {                                    // same layering as synth.layers()
                B2_REQUIRE(L.n < SYNTH_MAX_LAYERS, B2G16_E_INTERNAL, "too many layers");
                L.first[L.n++] = row;
                uint64_t visible = base + row, size = visible / 2 > 1 ? visible / 2 : 1;
                row += size < pk->n_cons - row ? size : pk->n_cons - row;
            }
            L.first[L.n] = pk->n_cons;
            DevCsr* M[3] = {&pk->A, &pk->B, &pk->C};
            for (int t = 0; t < 3; ++t) {
                M[t]->n_rows = pk->n_cons; M[t]->nnz = (t < 2 ? 3 : 1) * pk->n_cons;
                B2_CUDA(cudaMalloc((void**)&M[t]->row_ptr, (pk->n_cons + 1) * sizeof(uint32_t)));
                B2_CUDA(cudaMalloc((void**)&M[t]->col, (M[t]->nnz + 1) * sizeof(uint32_t)));
                B2_CUDA(cudaMalloc((void**)&M[t]->val, (M[t]->nnz + 1) * sizeof(Fr)));
            }
            synth_rows_kernel<<<grid_for(pk->n_cons + 1, 256), 256, 0, c->stream>>>(seed, pk->n_cons, base, L, pk->A.row_ptr, pk->A.col, pk->A.val,
                                                                                   pk->B.row_ptr, pk->B.col, pk->B.val, pk->C.row_ptr, pk->C.col, pk->C.val);
            c->launches++;
            B2_CUDA(cudaGetLastError());
            DevFlag bad(c);
            setup_on_device(c, pk, toxic, vk_out, bad);
        } catch (...) { cudaStreamSynchronize(c->stream); free_pk(pk); throw; }
        *out = pk;
    });
}

int b2g16_pk_shape(const b2g16_pk* pk, uint64_t out[5]) {
    if (!pk || !out) return B2G16_E_INVALID_ARG;
    out[0] = pk->log_m; out[1] = pk->n_cons; out[2] = pk->n_in; out[3] = pk->n_aux; out[4] = pk->n_vars;
    return B2G16_OK;
}

}  // extern "C"
