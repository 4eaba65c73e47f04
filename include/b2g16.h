/* libb2g16 -- B200-native Groth16 (r1cs_gg_ppzksnark, BLS12-381) proving backend.
 *
 * C ABI that stands where crypto3-fil-proofs calls into the (header-only, un-vendored)
 * crypto3 suite for its SNARK hot path.  The reference has no FFI layer: its callers
 * name C++ templates directly, so each entry point below cites the reference call site
 * or type it replaces (paths relative to the reference root;
 *   S/ = libs/storage/include/nil/filecoin/storage/proofs/
 *   F/ = libs/filecoin/include/nil/filecoin/proofs/).
 * INTEGRATION.md shows the C++ shim (`nil::crypto3::...` names) a maintainer binds.
 *
 * Data formats (Montgomery form never crosses the ABI):
 *   Fr  : 32 bytes little-endian, canonical (< r)            S/core/fr32.hpp:36-52
 *   G1  : affine x||y, 48-byte LE canonical each (96 B); identity = 96 zero bytes
 *   G2  : affine x.c0||x.c1||y.c0||y.c1 (192 B); identity = 192 zero bytes
 *   proof : 192 bytes = compressed A(48)||B(96)||C(48), big-endian with flag bits
 *           (F/constants.hpp:93 SINGLE_PARTITION_PROOF_LEN; F/api/seal.hpp:306-308)
 * Every function returns 0 on success or a negative B2G16_E_* code; the message is
 * available from b2g16_last_error().  No exceptions cross the ABI.  There is no CPU
 * fallback: without a CUDA device b2g16_init fails with B2G16_E_CUDA.
 * Host pointers unless a parameter name starts with d_ (device pointer).
 */
#ifndef B2G16_H
#define B2G16_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2G16_OK 0
#define B2G16_E_INVALID_ARG (-1)
#define B2G16_E_NOT_CANONICAL (-2)  /* Fr >= r or coordinate >= p (fr32 contract)            */
#define B2G16_E_NOT_ON_CURVE (-3)
#define B2G16_E_CUDA (-4)
#define B2G16_E_NCCL (-5)
#define B2G16_E_OOM (-6)
#define B2G16_E_INTERNAL (-7)

typedef struct b2g16_ctx b2g16_ctx;
typedef struct b2g16_pk b2g16_pk;

/* One context per process/GPU (one process per GPU; ranks are independent).
 * `stream` is a cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream) or NULL for a
 * library-owned stream.  All work is ordered on that stream. */
int b2g16_init(int device, void* stream, b2g16_ctx** out);
int b2g16_shutdown(b2g16_ctx* ctx);
const char* b2g16_last_error(b2g16_ctx* ctx);
int b2g16_synchronize(b2g16_ctx* ctx);
/* number of kernels this library has launched on ctx so far */
uint64_t b2g16_launch_count(b2g16_ctx* ctx);

/* ---- multiexp: replaces algebra::multiexp / multiexp_with_mixed_addition,
 *      used for the a / b_g1 / b_g2 / l / h queries of S/core/crypto/scheme_params.hpp:46-66.
 *      out = sum_i scalars[i] * bases[i] in affine ABI encoding. */
int b2g16_msm_g1(b2g16_ctx*, const uint8_t* bases96, const uint8_t* scalars32, size_t n, uint8_t out96[96]);
int b2g16_msm_g2(b2g16_ctx*, const uint8_t* bases192, const uint8_t* scalars32, size_t n, uint8_t out192[192]);

/* Device-resident bases (the proving-key residency of F/caches.hpp:48-67
 * GROTH_PARAM_MEMORY_CACHE): upload once, multiply many times. */
typedef struct b2g16_bases b2g16_bases;
int b2g16_bases_upload_g1(b2g16_ctx*, const uint8_t* bases96, size_t n, b2g16_bases** out);
int b2g16_bases_upload_g2(b2g16_ctx*, const uint8_t* bases192, size_t n, b2g16_bases** out);
/* bases[i] = k_i * G for the canonical scalars d_k (synthetic inputs for benchmarks) */
int b2g16_bases_generate(b2g16_ctx*, int group /*1|2*/, const uint8_t* d_scalars32, size_t n, b2g16_bases** out);
int b2g16_bases_download(b2g16_ctx*, const b2g16_bases*, size_t first, size_t count, uint8_t* out);
size_t b2g16_bases_count(const b2g16_bases*);
int b2g16_bases_free(b2g16_ctx*, b2g16_bases*);
/* scalars already in device memory (32-byte canonical each); result to host.
 * Uses points [first, first+n) of the resident set (MSM range-sharding across ranks). */
int b2g16_msm_resident(b2g16_ctx*, const b2g16_bases*, size_t first, const uint8_t* d_scalars32, size_t n, uint8_t* out_affine);
/* same, asynchronous: writes the affine result (96/192 B) to device memory d_out.  PRECONDITION: the scalars are
 * canonical (< r).  Nothing here can wait for a device-side check, so unlike b2g16_msm_resident this entry point
 * does not validate them: the signed-digit recoding is exact only below 2^255, and a non-canonical scalar yields a
 * wrong point, not an error.  Validate once where the scalars are produced, or use the synchronous form. */
int b2g16_msm_resident_async(b2g16_ctx*, const b2g16_bases*, size_t first, const uint8_t* d_scalars32, size_t n, uint8_t* d_out_affine);
/* out = sum of n affine points (combining per-GPU partial MSM results after all-gather) */
int b2g16_g1_sum(b2g16_ctx*, const uint8_t* pts96, size_t n, uint8_t out96[96]);
int b2g16_g2_sum(b2g16_ctx*, const uint8_t* pts192, size_t n, uint8_t out192[192]);

/* ---- evaluation_domain: replaces math::evaluation_domain<Fr>::fft / inverse_fft and the
 *      coset variants.  Natural order in and out, size 2^log_n, in place.
 *      coset uses the multiplicative generator g = 7: forward evaluates on g*D,
 *      inverse interpolates from g*D. */
int b2g16_ntt(b2g16_ctx*, uint8_t* fr32_inout, unsigned log_n, int inverse, int coset);
/* device-resident variant: d_data holds 2^log_n canonical 32-byte scalars */
int b2g16_ntt_device(b2g16_ctx*, uint8_t* d_fr32_inout, unsigned log_n, int inverse, int coset);
/* One transform of size 2^log_n sharded over G = 2^log_g GPUs (rank r holds elements
 * [r M, (r+1) M), M = 2^log_n / G).  The exchanges are all-to-alls done by the host over NCCL
 * (python/b2g16/dist.py::ntt_sharded shows the sequence); these are the device steps in between:
 *   cross:     per column of a G x (M/G) row-major matrix, a G-point transform and the twiddle
 *              w_N^(j k1), j = chunk * M/G + column  (forward: transform then twiddle; inverse: twiddle,
 *              inverse transform, 1/G).  With `coset` the inverse leaves the factor to the g^-e table.
 *   coset_scale: x[i] *= g^(offset+i)  (inverse: n^-1 g^-(offset+i)) for the rank's block of the domain.
 *   transpose: rows x cols -> cols x rows of 32-byte elements (cyclic <-> block order of a rank's spectrum). */
int b2g16_ntt_cross_device(b2g16_ctx*, const uint8_t* d_in, uint8_t* d_out, unsigned log_n, unsigned log_g, unsigned chunk, int inverse, int coset);
/* The cross step fused with both exchanges around it: d_in_peers[r] / d_out_peers[k] are the bases of rank
 * r's input block and rank k's output buffer (M scalars each) as mapped into this process; the kernel loads
 * its column chunk from every peer and stores row k straight into peer k over NVLink.  The caller orders it
 * with a barrier before (inputs written, outputs free) and after (stores landed).  Peer buffers:
 * b2g16_peer_alloc (cudaMalloc + IPC handle, 64 bytes, to be exchanged by the host), b2g16_peer_open /
 * _close on the other ranks, b2g16_peer_free by the owner. */
int b2g16_peer_alloc(b2g16_ctx*, size_t bytes, uint8_t** d_ptr, uint8_t handle_out[64]);
int b2g16_peer_open(b2g16_ctx*, const uint8_t handle[64], uint8_t** d_peer_ptr);
int b2g16_peer_close(b2g16_ctx*, uint8_t* d_peer_ptr);
int b2g16_peer_free(b2g16_ctx*, uint8_t* d_ptr);
int b2g16_ntt_cross_p2p_device(b2g16_ctx*, const uint8_t* const* d_in_peers, uint8_t* const* d_out_peers, unsigned log_n, unsigned log_g,
                               unsigned rank, int inverse, int coset);
int b2g16_fr_coset_scale_device(b2g16_ctx*, uint8_t* d_fr32_inout, size_t count, unsigned log_n, uint64_t offset, int inverse);
int b2g16_fr_transpose_device(b2g16_ctx*, const uint8_t* d_in, uint8_t* d_out, size_t rows, size_t cols);

/* ---- r1cs_to_qap witness map tail: a,b,c evaluations on the domain ->
 *      coefficients of H = (A*B - C)/Z  (m values, h[m-1] = 0). */
int b2g16_h_from_abc(b2g16_ctx*, const uint8_t* a32, const uint8_t* b32, const uint8_t* c32, unsigned log_m, uint8_t* h32_out);

/* ---- constraint system + proving key -------------------------------------------------- */
typedef struct {
    uint64_t n_rows, nnz;
    const uint32_t* row_ptr; /* n_rows + 1 */
    const uint32_t* col;     /* nnz  */
    const uint8_t* val;      /* nnz * 32, canonical Fr */
} b2g16_csr;

typedef struct {
    uint32_t log_m;          /* domain size 2^log_m >= n_cons + n_in + 1 (radix-2 only) */
    uint64_t n_cons, n_in, n_aux; /* variables z = (1, x_1..x_nin, w_1..w_naux) */
    b2g16_csr A, B, C;
} b2g16_r1cs_view;

/* Mirrors scheme_params<r1cs_gg_ppzksnark<bls12<381>>> (S/core/crypto/scheme_params.hpp:38-67)
 * in the dense crypto3 layout: one base per variable, identity where the polynomial is zero. */
typedef struct {
    const uint8_t *alpha_g1, *beta_g1, *delta_g1; /* 96 B   */
    const uint8_t *beta_g2, *delta_g2;            /* 192 B  */
    const uint8_t* a;    /* n_vars * 96  */
    const uint8_t* b_g1; /* n_vars * 96  */
    const uint8_t* b_g2; /* n_vars * 192 */
    const uint8_t* l;    /* n_aux  * 96  */
    const uint8_t* h;    /* (m-1)  * 96  */
    const b2g16_r1cs_view* cs;
} b2g16_pk_view;

/* Copies key + constraint system to the device (caller keeps its buffers). */
int b2g16_pk_load(b2g16_ctx*, const b2g16_pk_view*, b2g16_pk** out);
int b2g16_pk_free(b2g16_ctx*, b2g16_pk*);

/* Groth16 setup from explicit toxic waste (tau, alpha, beta, gamma, delta; 5*32 B), computed
 * on the device (fixed-base multiplications).  Benchmark/test support for synthetic keys
 * (S/core/parameter_cache.hpp:185-200 generates throw-away params the same way).
 * vk_out (optional): alpha_g1(96) beta_g2(192) gamma_g2(192) delta_g2(192) then (n_in+1) IC (96 each). */
int b2g16_pk_generate(b2g16_ctx*, const b2g16_r1cs_view* cs, const uint8_t toxic[160], b2g16_pk** out, uint8_t* vk_out);
/* Same, for the hashed synthetic constraint system generated on the device (python/b2g16/synth.py
 * HashedR1cs regenerates it bit-identically on the host at small sizes): n_cons = 2^log_m - n_in - 1,
 * n_aux = n_free + n_cons.  Lets 2^27-constraint instances (BASELINE configs 4 and 5) exist without
 * 33 GB of host arrays.  Bench/test support. */
int b2g16_pk_generate_synth(b2g16_ctx*, uint32_t log_m, uint64_t n_in, uint64_t n_free, uint64_t seed, const uint8_t toxic[160],
                            b2g16_pk** out, uint8_t* vk_out);
/* out = { log_m, n_cons, n_in, n_aux, n_vars } */
int b2g16_pk_shape(const b2g16_pk*, uint64_t out[5]);
/* read back one query for parity tests: which = 0 a, 1 b_g1, 2 b_g2, 3 l, 4 h,
 * 5 alpha_g1, 6 beta_g1, 7 delta_g1, 8 beta_g2, 9 delta_g2 */
int b2g16_pk_export(b2g16_ctx*, const b2g16_pk*, int which, uint8_t* out, size_t out_bytes);

/* Optional trace of every intermediate (any pointer may be NULL). Host buffers. */
typedef struct {
    uint8_t *a, *b, *c;                      /* m*32 evaluations             */
    uint8_t* h;                              /* m*32 coefficients            */
    uint8_t *msm_a, *msm_b1, *msm_l, *msm_h; /* 96  */
    uint8_t* msm_b2;                         /* 192 */
    uint8_t *A, *C;                          /* 96 affine proof points */
    uint8_t* B;                              /* 192 */
    float stage_ms[8]; /* out: ingest, abc(spmv), ntt/h, msm_a, msm_b1, msm_b2, msm_l, msm_h+final */
} b2g16_trace;

/* The prover: replaces r1cs_gg_ppzksnark prove(pk, primary, aux) as called from
 * compound_proof::circuit_proofs (S/core/proof/compound_proof.hpp:127-137) on behalf of
 * seal_commit_phase2 (F/api/seal.hpp:298-301) and generate_{window,winning}_post
 * (F/api/post.hpp:225,345).  assignment32 = (1, inputs, aux) = 1+n_in+n_aux scalars.
 * r, s are explicit (bit-exact parity).  priority mirrors compound_proof.hpp:134: a context runs one call at a
 * time, and a waiting call with priority != 0 is admitted before every waiting call without it (b2g16_prove_batch asks
 * for admission once per partition, so a priority proof from another thread runs between two partitions of a batch).
 * b2g16_prove_batch uploads the witness of partition k + 1 on its own stream while partition k is proved. */
int b2g16_prove(b2g16_ctx*, const b2g16_pk*, const uint8_t* assignment32, const uint8_t r32[32],
                const uint8_t s32[32], int priority, uint8_t proof_out[192], b2g16_trace* trace);
/* assignment already on the device; proof written to host */
int b2g16_prove_device(b2g16_ctx*, const b2g16_pk*, const uint8_t* d_assignment32, const uint8_t r32[32],
                       const uint8_t s32[32], int priority, uint8_t proof_out[192], b2g16_trace* trace);
/* One proof spread over `parts` GPUs (latency mode; what C4's single WindowPoSt partition needs
 * to use more than one GPU): every rank holds the key and the witness, runs the witness map and reduces only the
 * contiguous slice [part n/parts, (part+1) n/parts) of each of the five queries.  partials_out: msm_a(96) msm_b1(96)
 * msm_l(96) msm_h(96) msm_b2(192).  The host gathers parts x 576 bytes (an all-gather) and b2g16_prove_assemble
 * sums them and builds the same 192 bytes b2g16_prove would. */
int b2g16_prove_partial(b2g16_ctx*, const b2g16_pk*, const uint8_t* d_assignment32, unsigned part, unsigned parts, uint8_t partials_out[576]);
/* The same with the witness map spread over the ranks too (b2g16.dist.prove_sharded_full shows the sequence): each
 * rank evaluates its rows row0 + stride k of a, b, c (canonical scalars, cyclic distribution), the seven transforms
 * run sharded (b2g16_ntt_cross_p2p_device + local transforms), (a b - c) / Z is pointwise, and the rank's block of H
 * coefficients feeds its share of the H query directly. */
int b2g16_abc_evals_strided_device(b2g16_ctx*, const b2g16_pk*, const uint8_t* d_assignment32, uint64_t row0, uint64_t stride, uint64_t count,
                                   uint8_t* d_a32, uint8_t* d_b32, uint8_t* d_c32);
int b2g16_h_pointwise_device(b2g16_ctx*, const uint8_t* d_a32, const uint8_t* d_b32, const uint8_t* d_c32, uint64_t count, unsigned log_m, uint8_t* d_out32);
int b2g16_prove_partial_h(b2g16_ctx*, const b2g16_pk*, const uint8_t* d_assignment32, unsigned part, unsigned parts, const uint8_t* d_h_slice32,
                          uint64_t h_first, uint64_t h_count, uint8_t partials_out[576]);
int b2g16_prove_assemble(b2g16_ctx*, const b2g16_pk*, const uint8_t* partials, unsigned parts, const uint8_t r32[32],
                         const uint8_t s32[32], uint8_t proof_out[192]);
/* n_partitions independent witnesses for one key -> 192*n bytes (multi_proof, F/api/seal.hpp:304-308) */
int b2g16_prove_batch(b2g16_ctx*, const b2g16_pk*, const uint8_t* assignments32, size_t n_partitions,
                      const uint8_t* r32s, const uint8_t* s32s, int priority, uint8_t* proofs_out);

/* ---- several GPUs of one box from one process.  The reference's caller is a single C++ process that hands every
 *      partition of a proof to one prover call (compound_proof::circuit_proofs, S/core/proof/compound_proof.hpp:89-95,
 *      127-137; F/api/seal.hpp:298-308, F/api/post.hpp:305-348): these entry points spread that call over the devices
 *      given to b2g16_init_multi, one context and one host thread per device, no torch.distributed and no NCCL (the only
 *      exchange is 576 bytes of partial sums per device and proof, which come back through the host).
 *   b2g16_prove_batch_multi: with w = n_partitions / n full waves, device i proves partitions [i w, (i + 1) w) one after
 *      the other (b2g16_prove_batch: the next witness is uploaded while the current proof runs); the last, partial wave splits the devices
 *      into one group per remaining partition and every group proves its partition in latency mode.  Proofs are written
 *      in partition order; each equals what b2g16_prove returns for the same inputs.
 *   b2g16_prove_multi: one proof over all n devices in latency mode (b2g16_prove_partial on every device + assemble).
 * The same device may be listed more than once (two contexts share it): that is how a single-GPU box tests the schedule.
 * Keys are replicated: b2g16_mpk_load / _generate_synth build one resident key per context. */
typedef struct b2g16_mctx b2g16_mctx;
typedef struct b2g16_mpk b2g16_mpk;
int b2g16_init_multi(const int* devices, int n, b2g16_mctx** out);
int b2g16_shutdown_multi(b2g16_mctx*);
int b2g16_multi_count(const b2g16_mctx*);
b2g16_ctx* b2g16_multi_ctx(b2g16_mctx*, int slot);           /* the per-device context, e.g. for b2g16_vk_prepare */
const char* b2g16_multi_last_error(b2g16_mctx*);              /* first failure of the last multi call, with its device */
int b2g16_mpk_load(b2g16_mctx*, const b2g16_pk_view*, b2g16_mpk** out);
int b2g16_mpk_generate_synth(b2g16_mctx*, uint32_t log_m, uint64_t n_in, uint64_t n_free, uint64_t seed, const uint8_t toxic[160],
                             b2g16_mpk** out, uint8_t* vk_out);
b2g16_pk* b2g16_mpk_device_key(const b2g16_mpk*, int slot);
int b2g16_mpk_free(b2g16_mctx*, b2g16_mpk*);
int b2g16_prove_multi(b2g16_mctx*, const b2g16_mpk*, const uint8_t* assignment32, const uint8_t r32[32], const uint8_t s32[32], int priority,
                      uint8_t proof_out[192]);
/* the plan of b2g16_prove_batch_multi as data: plan[3 p ..] = { wave, first device slot, slots } of partition p (host-only) */
int b2g16_multi_schedule(int n_devices, size_t n_partitions, uint32_t* plan);
int b2g16_prove_batch_multi(b2g16_mctx*, const b2g16_mpk*, const uint8_t* assignments32, size_t n_partitions, const uint8_t* r32s,
                            const uint8_t* s32s, int priority, uint8_t* proofs_out);

/* Same prover when the caller's circuit synthesis already produced the constraint evaluations
 * (the bellperson-style flow of compound_proof::circuit_proofs, S/core/proof/compound_proof.hpp:127-137:
 * synthesize -> a, b, c on the domain -> prove): a32 / b32 / c32 are 2^log_m canonical scalars each
 * (rows >= n_cons must already hold the input-consistency values of SURVEY App. B step 2).
 * The key may have been loaded with empty matrices (cs->A.row_ptr == NULL). */
int b2g16_prove_abc(b2g16_ctx*, const b2g16_pk*, const uint8_t* a32, const uint8_t* b32, const uint8_t* c32,
                    const uint8_t* assignment32, const uint8_t r32[32], const uint8_t s32[32], int priority,
                    uint8_t proof_out[192], b2g16_trace* trace);

/* ---- verifier: replaces verify<r1cs_gg_ppzksnark>(pvk, primary_input, proof) as reached from
 *      verify_seal / verify_batch_seal (F/api/seal.hpp:339-395, 409-485) through compound_proof::verify
 *      (S/core/proof/compound_proof.hpp:97-102).  vk bytes: alpha_g1(96) beta_g2(192) gamma_g2(192)
 *      delta_g2(192) then n_ic = n_in + 1 IC points (96 each) -- the vk_out layout of b2g16_pk_generate.
 *      The prepared key caches the alpha/beta Miller value and keeps the IC bases on the device
 *      (the `pvk` of S/core/crypto/mapped_scheme_params.hpp:61). */
typedef struct b2g16_pvk b2g16_pvk;
int b2g16_vk_prepare(b2g16_ctx*, const uint8_t* vk, size_t n_ic, b2g16_pvk** out);
int b2g16_pvk_free(b2g16_ctx*, b2g16_pvk*);
/* inputs32: n_in canonical scalars (without the leading ONE).  *ok = 1 accept, 0 reject (a malformed
 * proof is a reject, not an error). */
int b2g16_verify(b2g16_ctx*, const b2g16_pvk*, const uint8_t* inputs32, size_t n_in, const uint8_t proof[192], int* ok);
/* n_proofs proofs (192 B each) with n_in inputs each, one random linear combination, one final
 * exponentiation.  rand32: n_proofs canonical scalars, or NULL to draw them inside. */
int b2g16_verify_batch(b2g16_ctx*, const b2g16_pvk*, const uint8_t* inputs32, size_t n_in, const uint8_t* proofs,
                       size_t n_proofs, const uint8_t* rand32, int* ok);
/* 192-byte proof (F/constants.hpp:93: compressed A(48) B(96) C(48)) <-> ABI affine points; decode
 * validates the encoding, the curve equation and prime-order subgroup membership. */
int b2g16_proof_decode(const uint8_t proof[192], uint8_t a96[96], uint8_t b192[192], uint8_t c96[96]);
int b2g16_proof_encode(const uint8_t a96[96], const uint8_t b192[192], const uint8_t c96[96], uint8_t proof[192]);

/* ---- parameter files: the `.params` / `.vk` cache entries of S/core/parameter_cache.hpp:50-94,125-153
 *      and the arrays S/core/crypto/mapped_scheme_params.hpp:65-81 indexes, in the bellperson
 *      Parameters::write layout.  Host-only byte shuffles; points arrive in the ABI
 *      encoding and are validated when handed to b2g16_pk_load / b2g16_vk_prepare. */
typedef struct {
    uint8_t alpha_g1[96], beta_g1[96], beta_g2[192], gamma_g2[192], delta_g1[96], delta_g2[192];
    uint64_t n_ic, n_h, n_l, n_a, n_b_g1, n_b_g2;
    uint8_t *ic, *h, *l, *a, *b_g1; /* G1, 96 B each  */
    uint8_t* b_g2;                  /* G2, 192 B each */
} b2g16_params;
int b2g16_params_read(const char* path, b2g16_params** out); /* buffers owned by the library */
int b2g16_params_free(b2g16_params*);
int b2g16_params_write(const char* path, const b2g16_params*);
/* vk block in the b2g16_vk_prepare layout; vk_bytes must be 672 + 96 * n_ic */
int b2g16_params_vk_bytes(const b2g16_params*, uint8_t* vk_out, size_t vk_bytes);
/* bellman / bellperson parameter files hold the a, b_g1 and b_g2 queries with the points at infinity
 * filtered out (a variable that appears in no row of A, resp. B, has a zero polynomial), and their
 * prover walks them with the density bitmaps it records during synthesis.  b2g16_pk_view wants the
 * dense crypto3 layout, so a caller holding such a file derives the bitmap from the constraint
 * system it loads anyway and expands:
 *   b2g16_query_density: density[j] = 1 iff j < n_always or column j (variable j of z) occurs in the matrix;
 *                       *count_out = popcount.  The a query carries the input-consistency rows, so ONE and every
 *                       public input are always present there: pass n_always = n_in + 1 with A, 0 with B.
 *   b2g16_csr_density : the n_always = 0 form (plain column occupancy; right for b_g1 / b_g2 only)
 *   b2g16_query_expand: dense[j] = next filtered point if density[j] else identity (zero bytes);
 *                       B2G16_E_INVALID_ARG unless n_filtered equals the number of set entries
 *   b2g16_query_filter: the inverse (drops all-zero points), for writing bellman-style files;
 *                       filtered_out may be NULL to count only
 * point_bytes = 96 (G1) or 192 (G2).  Host-only. */
int b2g16_query_density(const b2g16_csr* m, uint64_t n_vars, uint64_t n_always, uint8_t* density_out, uint64_t* count_out);
int b2g16_csr_density(const b2g16_csr* m, uint64_t n_vars, uint8_t* density_out, uint64_t* count_out);
int b2g16_query_expand(const uint8_t* filtered, uint64_t n_filtered, const uint8_t* density, uint64_t n_vars, size_t point_bytes,
                       uint8_t* dense_out);
int b2g16_query_filter(const uint8_t* dense, uint64_t n_vars, size_t point_bytes, uint8_t* filtered_out, uint8_t* density_out,
                       uint64_t* count_out);

/* ---- mapped / streamed parameter load (S/core/crypto/mapped_scheme_params.hpp:43-84, S/core/parameter_cache.hpp:125-129).
 *   b2g16_params_map   : counts and byte offsets of the six arrays of a `.params` file in file order
 *                        { ic, h, l, a, b_g1, b_g2 } -- the offset tables build_mapped_parameters keeps; reads the
 *                        length prefixes only.
 *   b2g16_pk_load_file : file -> resident key.  The queries stream through a ring of pinned host buffers
 *                        (tuning key "key_chunk_mib", default 32) and are converted on the device (big-endian
 *                        coordinates -> Montgomery limbs, range / curve checks); density-filtered a / b_g1 / b_g2
 *                        queries are expanded on the device with the column occupancy of cs->A / cs->B (plus ONE and
 *                        the public inputs for a), so the key never exists in host memory.  vk_out (optional,
 *                        672 + 96 (n_in + 1) bytes): the vk block in the b2g16_vk_prepare layout. */
int b2g16_params_map(const char* path, uint64_t counts[6], uint64_t offsets[6]);
int b2g16_pk_load_file(b2g16_ctx*, const char* path, const b2g16_r1cs_view* cs, b2g16_pk** out, uint8_t* vk_out, size_t vk_bytes);

/* ---- parameter-cache entries: the naming and the two small files of S/core/parameter_cache.hpp.
 *   b2g16_cache_dir        : $FIL_PROOFS_PARAMETER_CACHE, else /var/tmp/filecoin-proof-parameters        (:52-53,65-71)
 *   b2g16_cache_identifier : "<cache_prefix>-<hex sha256(parameter_set_identifier)>"                     (:166-173)
 *   b2g16_cache_path       : "<dir>/v28-<identifier>.<ext>", kind 0 = params, 1 = vk, 2 = meta            (:50,54-56,73-94)
 *   b2g16_meta_write/_read : cache_entry_metadata { sector_size } as {"sector_size":N}                   (:108-123)
 *   b2g16_vk_file_write/_read : the verifying key alone, = the vk block that opens a .params file        (:131-145);
 *                            read returns a b2g16_params with only the header and ic filled (free with b2g16_params_free)
 * out / cap: caller's buffer for the NUL-terminated string; B2G16_E_INVALID_ARG if it does not fit.  Host-only. */
#define B2G16_PARAMETER_CACHE_VERSION 28
int b2g16_sha256(const uint8_t* data, size_t len, uint8_t out32[32]);
int b2g16_cache_dir(char* out, size_t cap);
int b2g16_cache_identifier(const char* cache_prefix, const char* parameter_set_identifier, char* out, size_t cap);
int b2g16_cache_path(const char* identifier, int kind, char* out, size_t cap);
int b2g16_meta_write(const char* path, uint64_t sector_size);
int b2g16_meta_read(const char* path, uint64_t* sector_size);
int b2g16_vk_file_write(const char* path, const b2g16_params*);
int b2g16_vk_file_read(const char* path, b2g16_params** out);

/* ---- synthetic-instance support (bench/test): evaluate one layer of a forward-built R1CS on
 *      the device: for rows [row0,row1): z[out_base + i] = (A_i . z) * (B_i . z).  d_z canonical. */
int b2g16_synth_witness(b2g16_ctx*, const b2g16_pk*, uint8_t* d_assignment32, uint64_t n_free);

/* ---- measurement: sustained 32x32+64 multiply-add rate of this GPU in MAD32/s
 *      (the MSM roofline denominator) */
int b2g16_measure_mad32_peak(b2g16_ctx*, double* mad32_per_s, double* elapsed_ms);
/* last MSM's per-stage device times in ms: digits+histogram, scan, scatter, accumulate, reduce */
int b2g16_last_msm_timing(b2g16_ctx*, float out_ms[5]);
int b2g16_set_msm_timing(b2g16_ctx*, int enabled);
/* force the Pippenger window width (8..23 bits; 0 = choose by size) */
int b2g16_set_msm_window(b2g16_ctx*, int window_bits);
/* knobs by name: "serial_above_log" (proofs with a larger domain run their multiexps one after the other),
 * "red_log", "acc_variant", "kernel_log" (1 = record CUDA events around every bucket-accumulation launch) */
int b2g16_set_tuning(b2g16_ctx*, const char* key, int value);
/* drains the launch log: out[5 i ..] = { group 1|2, window bits, points, kernel ms, (point, window) entries added };
 * returns the entry count */
int b2g16_kernel_log(b2g16_ctx*, double* out, int max_entries);

/* ---- elementwise field kernels exposed for parity tests (device does the arithmetic).
 *      field: 0 = Fr (32 B), 1 = Fq (48 B); op: 0 add 1 sub 2 mul 3 inv 4 sqr 5 neg */
int b2g16_field_op(b2g16_ctx*, int field, int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out);
/* out = sum_i a_i b_i mod r over device-resident canonical scalars: the scalar side of the multiexp linearity
 * identity sum_i s_i (k_i G) = (sum_i s_i k_i) G, a size-independent parity property checked at full size */
int b2g16_fr_dot_device(b2g16_ctx*, const uint8_t* d_a32, const uint8_t* d_b32, size_t n, uint8_t out32[32]);

/* ---- an entry point that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added
 *      for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 *      forbid allocation).  The no-allocation variant of b2g16_fr_dot_device: d_out32 (DEVICE memory, 32 bytes) =
 *      sum_i a_i b_i mod r over n canonical scalars; d_status (device int, may be NULL) = 0, or B2G16_E_NOT_CANONICAL
 *      when an operand was >= r.  Its scratch -- one 32-byte partial sum per 4096 elements and the bad-input flag --
 *      is a file-scope device array of 256 + 32 * B2G16_NOALLOC_MAX_BLOCKS = 131,328 bytes that every GPU loading the
 *      library holds until exit; n > B2G16_NOALLOC_MAX_N (2^24) is refused with B2G16_E_INVALID_ARG.  The call enqueues
 *      a memset and two kernels on the context's stream and returns without synchronising; it calls no cudaMalloc /
 *      cudaFree / cudaMallocAsync.  All calls on one device share the array (it keeps the last call's contents; what
 *      must start at zero is cleared by each call): calls that overlap on the GPU race -- serialise them. */
#define B2G16_NOALLOC_MAX_BLOCKS 4096
#define B2G16_NOALLOC_MAX_N (4096ull * B2G16_NOALLOC_MAX_BLOCKS)
int b2g16_fr_dot_device_noalloc(b2g16_ctx*, const uint8_t* d_a32, const uint8_t* d_b32, size_t n, uint8_t* d_out32, int* d_status);

#ifdef __cplusplus
}
#endif
#endif
