// C ABI of libb2g16 (include/b2g16.h): context, multiexp, evaluation-domain transforms,
// witness-map tail, resident bases.  The prover entry points live in prove.cu.
// No CPU fallback anywhere: without a CUDA device b2g16_init fails.
#include <new>
#include <cstdlib>
#include <cstring>
#include <string>
#include "common.cuh"
#include "api_util.cuh"

using namespace b2;

extern "C" {

int b2g16_init(int device, void* stream, b2g16_ctx** out) {
    if (!out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count) return B2G16_E_CUDA;
    Ctx* c = new (std::nothrow) Ctx();
    if (!c) return B2G16_E_OOM;
    int rc = guard(c, [&] {
        c->device = device;
        B2_CUDA(cudaSetDevice(device));
        if (stream) c->stream = (cudaStream_t)stream;
        else { B2_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
        for (int i = 0; i < 4; ++i) {
            B2_CUDA(cudaStreamCreateWithFlags(&c->aux[i], cudaStreamNonBlocking));
            B2_CUDA(cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming));
        }
        B2_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
        for (auto& e : c->ev_stage) B2_CUDA(cudaEventCreate(&e));
        cudaMemPool_t pool;
        B2_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        uint64_t keep = ~0ull;                                      // keep freed workspace in the pool
        B2_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        if (const char* e = getenv("B2G16_MSM_C")) c->msm_window_override = atoi(e);
        if (const char* e = getenv("B2G16_ACC_VARIANT")) c->acc_variant = atoi(e);
        if (const char* e = getenv("B2G16_SERIAL_LOG")) c->serial_above_log = atoi(e);
        if (const char* e = getenv("B2G16_RED_BITS")) c->red_bits = atoi(e) != 0;
        if (const char* e = getenv("B2G16_TREE_LEVELS")) c->tree_levels = atoi(e);
        if (const char* e = getenv("B2G16_TREE_K")) { int v = atoi(e); if (v >= 2 && v <= 64) c->tree_k = v; }
        if (const char* e = getenv("B2G16_TREE_K2")) { int v = atoi(e); if (v >= 2 && v <= 64) c->tree_k2 = v; }
        if (const char* e = getenv("B2G16_TREE_WGROUP")) c->tree_wgroup = atoi(e);
        if (const char* e = getenv("B2G16_RED_LOG")) { int v = atoi(e); if (v >= 1 && v <= 6) c->red_log = v; }
        ntt_init_roots(c);
    });
    if (rc != B2G16_OK) { delete c; return rc; }
    *out = reinterpret_cast<b2g16_ctx*>(c);
    return B2G16_OK;
}

int b2g16_shutdown(b2g16_ctx* h) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c) return B2G16_E_INVALID_ARG;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    ntt_free_domains(c);
    for (int i = 0; i < 2; ++i) cudaFree(c->fixed_base[i]);
    for (int i = 0; i < 4; ++i) { cudaStreamDestroy(c->aux[i]); cudaEventDestroy(c->ev_join[i]); }
    cudaEventDestroy(c->ev_fork);
    for (auto& e : c->ev_stage) cudaEventDestroy(e);
    for (auto& e : c->klog) { cudaEventDestroy(e.e0); cudaEventDestroy(e.e1); }
    cudaFree(c->d_klog_entries);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
    return B2G16_OK;
}

const char* b2g16_last_error(b2g16_ctx* h) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    return c ? c->last_error.c_str() : "null context";
}
int b2g16_synchronize(b2g16_ctx* h) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { B2_CUDA(cudaStreamSynchronize(c->stream)); });
}
uint64_t b2g16_launch_count(b2g16_ctx* h) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    return c ? c->launches : 0;
}

// ------------------------------------------------------------------------ resident bases
struct b2g16_bases { int group; size_t n; void* d_pts; };

static int bases_upload(b2g16_ctx* h, int group, const uint8_t* bytes, size_t n, b2g16_bases** out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out || (n && !bytes)) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        const size_t enc = group == 1 ? 96 : 192;
        void* d_pts = nullptr;
        if (n) B2_CUDA(cudaMalloc(&d_pts, n * enc));
        try {
            DevFlag bad(c);
            if (n) {
                // ABI bytes and Montgomery limbs have the same size: upload into place, convert in place (one thread
                // reads its whole point before it writes it) -- no second n * enc staging buffer for 72 GiB-class keys
                B2_CUDA(cudaMemcpyAsync(d_pts, bytes, n * enc, cudaMemcpyHostToDevice, c->stream));
                if (group == 1) g1_from_bytes_dev(c, (const uint8_t*)d_pts, (G1Affine*)d_pts, n, bad.ptr(), true);
                else g2_from_bytes_dev(c, (const uint8_t*)d_pts, (G2Affine*)d_pts, n, bad.ptr(), true);
            }
            bad.check("bases");
        } catch (...) { cudaFree(d_pts); throw; }
        *out = new b2g16_bases{group, n, d_pts};
    });
}
int b2g16_bases_upload_g1(b2g16_ctx* h, const uint8_t* b, size_t n, b2g16_bases** out) { return bases_upload(h, 1, b, n, out); }
int b2g16_bases_upload_g2(b2g16_ctx* h, const uint8_t* b, size_t n, b2g16_bases** out) { return bases_upload(h, 2, b, n, out); }

int b2g16_bases_generate(b2g16_ctx* h, int group, const uint8_t* d_scalars32, size_t n, b2g16_bases** out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out || (group != 1 && group != 2) || (n && !d_scalars32)) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        void* d_pts = nullptr;
        if (n) B2_CUDA(cudaMalloc(&d_pts, n * (group == 1 ? sizeof(G1Affine) : sizeof(G2Affine))));
        try {
            DevFlag bad(c);
            scalars_check_dev(c, d_scalars32, n, bad.ptr());
            if (group == 1) gen_mul_g1_dev(c, (const uint32_t*)d_scalars32, n, (G1Affine*)d_pts);
            else gen_mul_g2_dev(c, (const uint32_t*)d_scalars32, n, (G2Affine*)d_pts);
            bad.check("scalars");
        } catch (...) { cudaFree(d_pts); throw; }
        *out = new b2g16_bases{group, n, d_pts};
    });
}

int b2g16_bases_download(b2g16_ctx* h, const b2g16_bases* b, size_t first, size_t count, uint8_t* out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !b || (count && !out) || first > b->n || count > b->n - first) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        if (!count) return;
        const size_t enc = b->group == 1 ? 96 : 192;
        DevBuf<uint8_t> raw(count * enc, c->stream);
        if (b->group == 1) g1_to_bytes_dev(c, (const G1Affine*)b->d_pts + first, raw.get(), count, c->stream);
        else g2_to_bytes_dev(c, (const G2Affine*)b->d_pts + first, raw.get(), count, c->stream);
        B2_CUDA(cudaMemcpyAsync(out, raw.get(), count * enc, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
}
size_t b2g16_bases_count(const b2g16_bases* b) { return b ? b->n : 0; }
int b2g16_bases_free(b2g16_ctx* h, b2g16_bases* b) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !b) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        B2_CUDA(cudaStreamSynchronize(c->stream));
        B2_CUDA(cudaFree(b->d_pts));
        delete b;
    });
}

// ------------------------------------------------------------------------------ multiexp
static void msm_any(Ctx* c, const b2g16_bases* b, size_t first, const uint8_t* d_scalars, size_t n, uint8_t* d_out, cudaStream_t st) {
    if (b->group == 1) msm_g1_dev(c, (const G1Affine*)b->d_pts + first, (const uint32_t*)d_scalars, n, d_out, st, 0);
    else msm_g2_dev(c, (const G2Affine*)b->d_pts + first, (const uint32_t*)d_scalars, n, d_out, st, 0);
}

int b2g16_msm_resident_async(b2g16_ctx* h, const b2g16_bases* b, size_t first, const uint8_t* d_scalars32, size_t n, uint8_t* d_out_affine) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !b || !d_out_affine || (n && !d_scalars32) || first > b->n || n > b->n - first) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { msm_any(c, b, first, d_scalars32, n, d_out_affine, c->stream); });
}

int b2g16_msm_resident(b2g16_ctx* h, const b2g16_bases* b, size_t first, const uint8_t* d_scalars32, size_t n, uint8_t* out_affine) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !b || !out_affine || (n && !d_scalars32) || first > b->n || n > b->n - first) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        const size_t enc = b->group == 1 ? 96 : 192;
        DevBuf<uint8_t> d_out(enc, c->stream);
        DevFlag bad(c);
        scalars_check_dev(c, d_scalars32, n, bad.ptr());
        msm_any(c, b, first, d_scalars32, n, d_out.get(), c->stream);
        B2_CUDA(cudaMemcpyAsync(out_affine, d_out.get(), enc, cudaMemcpyDeviceToHost, c->stream));
        bad.check("scalars");
    });
}

static int msm_host(b2g16_ctx* h, int group, const uint8_t* bases, const uint8_t* scalars32, size_t n, uint8_t* out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out || (n && (!bases || !scalars32))) return B2G16_E_INVALID_ARG;
    b2g16_bases* b = nullptr;
    int rc = bases_upload(h, group, bases, n, &b);
    if (rc != B2G16_OK) return rc;
    rc = guard(c, [&] {
        const size_t enc = group == 1 ? 96 : 192;
        DevBuf<uint8_t> d_s(n * 32, c->stream), d_out(enc, c->stream);
        DevFlag bad(c);
        if (n) B2_CUDA(cudaMemcpyAsync(d_s.get(), scalars32, n * 32, cudaMemcpyHostToDevice, c->stream));
        scalars_check_dev(c, d_s.get(), n, bad.ptr());
        msm_any(c, b, 0, d_s.get(), n, d_out.get(), c->stream);
        B2_CUDA(cudaMemcpyAsync(out, d_out.get(), enc, cudaMemcpyDeviceToHost, c->stream));
        bad.check("scalars");
    });
    b2g16_bases_free(h, b);
    return rc;
}
int b2g16_msm_g1(b2g16_ctx* h, const uint8_t* bases96, const uint8_t* scalars32, size_t n, uint8_t out96[96]) { return msm_host(h, 1, bases96, scalars32, n, out96); }
int b2g16_msm_g2(b2g16_ctx* h, const uint8_t* bases192, const uint8_t* scalars32, size_t n, uint8_t out192[192]) { return msm_host(h, 2, bases192, scalars32, n, out192); }

static int sum_host(b2g16_ctx* h, int group, const uint8_t* pts, size_t n, uint8_t* out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out || (n && !pts)) return B2G16_E_INVALID_ARG;
    b2g16_bases* b = nullptr;
    int rc = bases_upload(h, group, pts, n, &b);
    if (rc != B2G16_OK) return rc;
    rc = guard(c, [&] {
        const size_t enc = group == 1 ? 96 : 192;
        DevBuf<uint8_t> d_out(enc, c->stream);
        if (group == 1) sum_g1_dev(c, (const G1Affine*)b->d_pts, n, d_out.get());
        else sum_g2_dev(c, (const G2Affine*)b->d_pts, n, d_out.get());
        B2_CUDA(cudaMemcpyAsync(out, d_out.get(), enc, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
    b2g16_bases_free(h, b);
    return rc;
}
int b2g16_g1_sum(b2g16_ctx* h, const uint8_t* pts96, size_t n, uint8_t out96[96]) { return sum_host(h, 1, pts96, n, out96); }
int b2g16_g2_sum(b2g16_ctx* h, const uint8_t* pts192, size_t n, uint8_t out192[192]) { return sum_host(h, 2, pts192, n, out192); }

// --------------------------------------------------------------------- evaluation domain
int b2g16_ntt_device(b2g16_ctx* h, uint8_t* d_fr32_inout, unsigned log_n, int inverse, int coset) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_fr32_inout || log_n > 30) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevBuf<Fr> tmp;
        if (ntt_needs_tmp((int)log_n)) tmp.alloc((size_t)1 << log_n, c->stream);
        DevFlag bad(c);
        Fr* d = reinterpret_cast<Fr*>(d_fr32_inout);
        ntt_natural_ex(c, d, d, tmp.get(), (int)log_n, inverse != 0, coset != 0, true, true, bad.ptr());
        bad.check("scalars");
    });
}
// ---- one transform sharded over 2^log_g ranks: the per-rank device steps between the all-to-all
//      exchanges that python/b2g16/dist.py (or any NCCL host) performs.  All buffers hold canonical scalars.
int b2g16_ntt_cross_device(b2g16_ctx* h, const uint8_t* d_in, uint8_t* d_out, unsigned log_n, unsigned log_g, unsigned chunk, int inverse, int coset) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_in || !d_out || d_in == d_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevFlag bad(c);
        ntt_cross_dev(c, reinterpret_cast<const Fr*>(d_in), reinterpret_cast<Fr*>(d_out), (int)log_n, (int)log_g, chunk, inverse != 0, coset != 0, bad.ptr());
        bad.check("scalars");
    });
}
// ---- peer memory: buffers other ranks' kernels may load from / store to over NVLink (CUDA IPC)
int b2g16_peer_alloc(b2g16_ctx* h, size_t bytes, uint8_t** d_ptr, uint8_t handle_out[64]) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_ptr || !handle_out || bytes == 0) return B2G16_E_INVALID_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    return guard(c, [&] {
        void* p = nullptr;
        B2_CUDA(cudaMalloc(&p, bytes));                      // a whole allocation: IPC handles name allocations, not sub-ranges
        cudaIpcMemHandle_t hd;
        cudaError_t e = cudaIpcGetMemHandle(&hd, p);
        if (e != cudaSuccess) { cudaFree(p); B2_CUDA(e); }
        memcpy(handle_out, &hd, 64);
        *d_ptr = (uint8_t*)p;
    });
}
int b2g16_peer_open(b2g16_ctx* h, const uint8_t handle[64], uint8_t** d_peer_ptr) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !handle || !d_peer_ptr) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        cudaIpcMemHandle_t hd; memcpy(&hd, handle, 64);
        void* p = nullptr;
        B2_CUDA(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
        *d_peer_ptr = (uint8_t*)p;
    });
}
int b2g16_peer_close(b2g16_ctx* h, uint8_t* d_peer_ptr) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_peer_ptr) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { B2_CUDA(cudaStreamSynchronize(c->stream)); B2_CUDA(cudaIpcCloseMemHandle(d_peer_ptr)); });
}
int b2g16_peer_free(b2g16_ctx* h, uint8_t* d_ptr) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_ptr) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { B2_CUDA(cudaStreamSynchronize(c->stream)); B2_CUDA(cudaFree(d_ptr)); });
}
// cross step fused with both of its exchanges: loads straight from every rank's block, stores straight
// into every rank's output buffer.  The caller brackets it with barriers (inputs ready / outputs landed).
int b2g16_ntt_cross_p2p_device(b2g16_ctx* h, const uint8_t* const* d_in_peers, uint8_t* const* d_out_peers, unsigned log_n, unsigned log_g,
                               unsigned rank, int inverse, int coset) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_in_peers || !d_out_peers) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevFlag bad(c);
        ntt_cross_p2p_dev(c, reinterpret_cast<const Fr* const*>(d_in_peers), reinterpret_cast<Fr* const*>(d_out_peers), (int)log_n, (int)log_g, rank,
                          inverse != 0, coset != 0, bad.ptr());
        bad.check("scalars");
    });
}
int b2g16_fr_coset_scale_device(b2g16_ctx* h, uint8_t* d_fr32_inout, size_t count, unsigned log_n, uint64_t offset, int inverse) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || (count && !d_fr32_inout) || log_n > 30) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevFlag bad(c);
        fr_coset_scale_dev(c, reinterpret_cast<Fr*>(d_fr32_inout), count, (int)log_n, offset, inverse != 0, bad.ptr());
        bad.check("scalars");
    });
}
int b2g16_fr_transpose_device(b2g16_ctx* h, const uint8_t* d_in, uint8_t* d_out, size_t rows, size_t cols) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !d_in || !d_out || d_in == d_out) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { fr_transpose_dev(c, reinterpret_cast<const Fr*>(d_in), reinterpret_cast<Fr*>(d_out), rows, cols); });
}

int b2g16_ntt(b2g16_ctx* h, uint8_t* fr32_inout, unsigned log_n, int inverse, int coset) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !fr32_inout || log_n > 30) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        const size_t n = (size_t)1 << log_n;
        DevBuf<Fr> d(n, c->stream), tmp;
        if (ntt_needs_tmp((int)log_n)) tmp.alloc(n, c->stream);
        DevFlag bad(c);
        B2_CUDA(cudaMemcpyAsync(d.get(), fr32_inout, n * 32, cudaMemcpyHostToDevice, c->stream));
        ntt_natural_ex(c, d.get(), d.get(), tmp.get(), (int)log_n, inverse != 0, coset != 0, true, true, bad.ptr());
        bad.check("scalars");                          // synchronises; nothing is written back on error
        B2_CUDA(cudaMemcpyAsync(fr32_inout, d.get(), n * 32, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
}

int b2g16_h_from_abc(b2g16_ctx* h, const uint8_t* a32, const uint8_t* b32, const uint8_t* c32, unsigned log_m, uint8_t* h32_out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !a32 || !b32 || !c32 || !h32_out || log_m > 30) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        const size_t m = (size_t)1 << log_m;
        DevBuf<Fr> da(m, c->stream), db(m, c->stream), dc(m, c->stream), dh(m, c->stream);
        DevFlag bad(c);
        B2_CUDA(cudaMemcpyAsync(da.get(), a32, m * 32, cudaMemcpyHostToDevice, c->stream));
        B2_CUDA(cudaMemcpyAsync(db.get(), b32, m * 32, cudaMemcpyHostToDevice, c->stream));
        B2_CUDA(cudaMemcpyAsync(dc.get(), c32, m * 32, cudaMemcpyHostToDevice, c->stream));
        h_from_abc_dev(c, da.get(), db.get(), dc.get(), (int)log_m, reinterpret_cast<uint32_t*>(dh.get()), true, bad.ptr());
        bad.check("scalars");
        B2_CUDA(cudaMemcpyAsync(h32_out, dh.get(), m * 32, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
}

// --------------------------------------------------------------- parity / measurement aids
int b2g16_field_op(b2g16_ctx* h, int field, int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || (field != 0 && field != 1) || op < 0 || op > 5 || (n && (!a || !b || !out))) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        const size_t sz = field == 0 ? 32 : 48;
        DevBuf<uint8_t> da(n * sz, c->stream), db(n * sz, c->stream), dout(n * sz, c->stream);
        DevFlag bad(c);
        if (n) {
            B2_CUDA(cudaMemcpyAsync(da.get(), a, n * sz, cudaMemcpyHostToDevice, c->stream));
            B2_CUDA(cudaMemcpyAsync(db.get(), b, n * sz, cudaMemcpyHostToDevice, c->stream));
        }
        field_op_dev(c, field, op, da.get(), db.get(), n, dout.get(), bad.ptr());
        bad.check("field elements");
        if (n) B2_CUDA(cudaMemcpyAsync(out, dout.get(), n * sz, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA(cudaStreamSynchronize(c->stream));
    });
}

int b2g16_fr_dot_device(b2g16_ctx* h, const uint8_t* d_a32, const uint8_t* d_b32, size_t n, uint8_t out32[32]) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out32 || (n && (!d_a32 || !d_b32))) return B2G16_E_INVALID_ARG;
    return guard(c, [&] {
        DevBuf<uint8_t> d_out(32, c->stream);
        DevFlag bad(c);
        fr_dot_dev(c, d_a32, d_b32, n, d_out.get(), bad.ptr());
        B2_CUDA(cudaMemcpyAsync(out32, d_out.get(), 32, cudaMemcpyDeviceToHost, c->stream));
        bad.check("scalars");
    });
}

int b2g16_measure_mad32_peak(b2g16_ctx* h, double* mad32_per_s, double* elapsed_ms) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !mad32_per_s) return B2G16_E_INVALID_ARG;
    return guard(c, [&] { *mad32_per_s = mad32_peak_dev(c, elapsed_ms); });
}
int b2g16_last_msm_timing(b2g16_ctx* h, float out_ms[5]) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !out_ms) return B2G16_E_INVALID_ARG;
    memcpy(out_ms, c->msm_ms, sizeof(c->msm_ms));
    return B2G16_OK;
}
int b2g16_set_msm_timing(b2g16_ctx* h, int enabled) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c) return B2G16_E_INVALID_ARG;
    c->time_msm = enabled != 0;
    return B2G16_OK;
}
// Tuning / measurement knobs by name (the same ones b2g16_init reads from B2G16_* environment variables).
int b2g16_set_tuning(b2g16_ctx* h, const char* key, int value) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || !key) return B2G16_E_INVALID_ARG;
    std::lock_guard<std::mutex> lock(c->mu);
    const std::string k(key);
    if (k == "serial_above_log") c->serial_above_log = value;
    else if (k == "red_log" && value >= 1 && value <= 6) c->red_log = value;
    else if (k == "acc_variant") c->acc_variant = value;
    else if (k == "red_bits") c->red_bits = value != 0;
    else if (k == "tree_levels") c->tree_levels = value;
    else if (k == "tree_k" && value >= 2 && value <= 64) c->tree_k = value;
    else if (k == "tree_k2" && value >= 2 && value <= 64) c->tree_k2 = value;
    else if (k == "tree_wgroup") c->tree_wgroup = value;
    else if (k == "key_chunk_mib" && value >= 1 && value <= 1024) c->key_chunk_mib = value;
    else if (k == "kernel_log") {
        c->klog_on = value != 0;
        cudaSetDevice(c->device);
        if (c->klog_on && !c->d_klog_entries && cudaMalloc((void**)&c->d_klog_entries, 64 * sizeof(uint32_t)) != cudaSuccess) return B2G16_E_OOM;
        for (auto& e : c->klog) { cudaEventDestroy(e.e0); cudaEventDestroy(e.e1); }
        c->klog.clear();
    } else return B2G16_E_INVALID_ARG;
    return B2G16_OK;
}
// Drains the accumulate-kernel launch log: out[5 * i ..] = { group (1|2), window bits, points, kernel milliseconds,
// sorted (point, window) entries = additions the launch performed };
// returns the number of entries written (at most max_entries) or a negative error.  Synchronises the device.
int b2g16_kernel_log(b2g16_ctx* h, double* out, int max_entries) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || (max_entries > 0 && !out)) return B2G16_E_INVALID_ARG;
    int n = 0;
    int rc = guard(c, [&] {
        B2_CUDA(cudaDeviceSynchronize());
        uint32_t entries[64] = {};
        if (c->d_klog_entries) B2_CUDA(cudaMemcpy(entries, c->d_klog_entries, sizeof entries, cudaMemcpyDeviceToHost));
        for (auto& e : c->klog) {
            float ms = 0.f;
            if (n < max_entries) {
                B2_CUDA(cudaEventElapsedTime(&ms, e.e0, e.e1));
                out[5 * n] = e.group; out[5 * n + 1] = e.c; out[5 * n + 2] = (double)e.n; out[5 * n + 3] = ms; out[5 * n + 4] = n < 64 ? entries[n] : 0;
                ++n;
            }
            cudaEventDestroy(e.e0); cudaEventDestroy(e.e1);
        }
        c->klog.clear();
    });
    return rc == B2G16_OK ? n : rc;
}
int b2g16_set_msm_window(b2g16_ctx* h, int window_bits) {
    Ctx* c = reinterpret_cast<Ctx*>(h);
    if (!c || (window_bits != 0 && (window_bits < 8 || window_bits > 23))) return B2G16_E_INVALID_ARG;
    c->msm_window_override = window_bits;
    return B2G16_OK;
}

}  // extern "C"
