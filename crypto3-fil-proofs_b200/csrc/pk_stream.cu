// Parameter file -> resident proving key without a host copy of the key.
//
// The reference maps its `.params` cache entry and keeps per-query offset tables so that the prover pulls points
// from the file on demand (libs/storage/include/nil/filecoin/storage/proofs/core/crypto/mapped_scheme_params.hpp:43-84,
// read path core/parameter_cache.hpp:125-129).  On a GPU the key must be resident, so the equivalent is:
//   b2g16_params_map   : the offset / count table of the five queries and the vk block (what build_mapped_parameters
//                        computes), from the length prefixes alone -- no point is read;
//   b2g16_pk_load_file : streams each query through a ring of pinned host buffers (the read of chunk k+1 -- positioned
//                        reads split over four threads -- overlaps the cudaMemcpyAsync + conversion kernel of chunk k)
//                        and converts ON THE DEVICE: big-endian file
//                        coordinates -> little-endian limbs -> Montgomery form, range + curve checks, G2 coefficient
//                        order, infinity flag.  Density-filtered a / b_g1 / b_g2 queries (bellman's layout) are expanded
//                        on the device into the dense layout through a position table, so neither the filtered nor the
//                        dense query ever exists in host memory.
// File layout: params_io.cu (bellperson Parameters::write; by recall).
#include <stdio.h>
#include <string.h>

#include <unistd.h>

#include <thread>
#include <vector>

#include "api_util.cuh"
#include "pk.cuh"

namespace b2 {

// 48 big-endian bytes (already in 12 words as loaded) -> 12 little-endian limbs
B2_D Fq fq_from_be_words(const uint32_t* w) {
    Fq c;
#pragma unroll
    for (int i = 0; i < 12; ++i) c.v[i] = __byte_perm(w[11 - i], 0, 0x0123);
    return c;
}
B2_D Fq fq_file_to_mont(const uint32_t* w, int& bad) {
    Fq c = fq_from_be_words(w);
    if (Fq::geq_p(c.v)) bad |= 1;
    return c * Fq::r2();
}
B2_D Fq fq_small_k(uint32_t k) { Fq r = Fq::zero(); const Fq one = Fq::one(); for (uint32_t i = 0; i < k; ++i) r = r + one; return r; }

// One file record per thread: G1 = x y (96 B), G2 = x.c1 x.c0 y.c1 y.c0 (192 B); bit 7 of byte 0 = compressed (not
// allowed here), bit 6 = infinity.  pos (optional): dense index of filtered entry f0 + i.
template <bool G2> __global__ void __launch_bounds__(128)
points_from_file_kernel(const uint8_t* __restrict__ in, size_t n, const uint32_t* __restrict__ pos, size_t f0, uint8_t* __restrict__ out_base, int* d_bad) {
    constexpr int SZ = G2 ? 192 : 96, WORDS = SZ / 4;
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t w[WORDS];
        const uint4* q = reinterpret_cast<const uint4*>(in + i * SZ);
#pragma unroll
        for (int k = 0; k < WORDS / 4; ++k) { const uint4 t = q[k]; w[4 * k] = t.x; w[4 * k + 1] = t.y; w[4 * k + 2] = t.z; w[4 * k + 3] = t.w; }
        const uint32_t flags = w[0] & 0xffu;                // byte 0 of the record (words are little-endian loads)
        const size_t dst = pos ? (size_t)pos[f0 + i] : f0 + i;
        if (flags & 0x80u) bad |= 1;
        if (G2) {
            G2Affine p;
            if (flags & 0x40u) p = G2Affine::identity();
            else {
                p.x.c1 = fq_file_to_mont(w, bad); p.x.c0 = fq_file_to_mont(w + 12, bad);
                p.y.c1 = fq_file_to_mont(w + 24, bad); p.y.c0 = fq_file_to_mont(w + 36, bad);
                const Fq four = fq_small_k(4);
                const Fq2 b{four, four};
                if (p.y.sqr() != p.x.sqr() * p.x + b) bad |= 2;
            }
            reinterpret_cast<G2Affine*>(out_base)[dst] = p;
        } else {
            G1Affine p;
            if (flags & 0x40u) p = G1Affine::identity();
            else {
                p.x = fq_file_to_mont(w, bad); p.y = fq_file_to_mont(w + 12, bad);
                if (p.y.sqr() != p.x.sqr() * p.x + fq_small_k(4)) bad |= 2;
            }
            reinterpret_cast<G1Affine*>(out_base)[dst] = p;
        }
    }
    if (bad) atomicOr(d_bad, bad);
}

namespace {

struct FileIndex { uint64_t count[6]; uint64_t offset[6]; uint64_t file_size; };   // ic, h, l, a, b_g1, b_g2

bool rd_u32_be(FILE* f, uint64_t* v) { uint8_t b[4]; if (fread(b, 1, 4, f) != 4) return false; *v = ((uint32_t)b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3]; return true; }

// walks the length prefixes; false when the file is truncated or a prefix points past its end
bool index_file(FILE* f, FileIndex* ix) {
    if (fseeko(f, 0, SEEK_END) != 0) return false;
    const off_t end = ftello(f);
    if (end < 864) return false;
    ix->file_size = (uint64_t)end;
    uint64_t at = 864;
    for (int q = 0; q < 6; ++q) {
        if (fseeko(f, (off_t)at, SEEK_SET) != 0 || !rd_u32_be(f, &ix->count[q])) return false;
        at += 4;
        ix->offset[q] = at;
        const uint64_t bytes = ix->count[q] * (q == 5 ? 192ull : 96ull);
        if (bytes > ix->file_size - at) return false;
        at += bytes;
    }
    return true;
}

// ring of pinned host buffers with a device staging twin each
struct Ring {
    static constexpr int SLOTS = 3;
    size_t bytes = 0;
    uint8_t* host[SLOTS] = {}; uint8_t* dev[SLOTS] = {}; cudaEvent_t done[SLOTS] = {};
    explicit Ring(size_t b) : bytes(b) {
        for (int i = 0; i < SLOTS; ++i) {
            B2_CUDA(cudaMallocHost((void**)&host[i], bytes));
            B2_CUDA(cudaMalloc((void**)&dev[i], bytes));
            B2_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
        }
    }
    ~Ring() { for (int i = 0; i < SLOTS; ++i) { if (host[i]) cudaFreeHost(host[i]); if (dev[i]) cudaFree(dev[i]); if (done[i]) cudaEventDestroy(done[i]); } }
};

// bytes [at, at + len) of the file into dst, split over a few threads (one thread's copy out of the page cache tops out
// near 6 GB/s; the PCIe link behind the pinned buffer takes several times that)
bool read_parallel(FILE* f, uint64_t at, uint8_t* dst, size_t len) {
    const int fd = fileno(f);
    const int T = len >= (8u << 20) ? 4 : 1;
    bool ok[4] = {true, true, true, true};
    auto part = [&](int t) {
        size_t lo = len * t / T, hi = len * (t + 1) / T;
        while (lo < hi) {
            const ssize_t got = pread(fd, dst + lo, hi - lo, (off_t)(at + lo));
            if (got <= 0) { ok[t] = false; return; }
            lo += (size_t)got;
        }
    };
    std::thread th[3];
    for (int t = 1; t < T; ++t) th[t - 1] = std::thread(part, t);
    part(0);
    for (int t = 1; t < T; ++t) th[t - 1].join();
    return ok[0] && ok[1] && ok[2] && ok[3];
}

// file records [0, n) at `offset` -> out_base (dense positions through pos, or identity mapping)
template <bool G2> void stream_query(Ctx* c, FILE* f, uint64_t offset, uint64_t n, const uint32_t* d_pos, uint8_t* out_base, Ring& ring, int* d_bad, int& slot) {
    constexpr size_t SZ = G2 ? 192 : 96;
    const size_t per = ring.bytes / SZ;
    for (uint64_t f0 = 0; f0 < n; f0 += per) {                             // positioned reads (pread): the FILE's own offset is not used
        const size_t k = (size_t)(n - f0 < per ? n - f0 : per);
        B2_CUDA(cudaEventSynchronize(ring.done[slot]));                 // the slot's previous chunk has been consumed
        B2_REQUIRE(read_parallel(f, offset + f0 * SZ, ring.host[slot], k * SZ), B2G16_E_INVALID_ARG, "parameter file: short read");
        B2_CUDA(cudaMemcpyAsync(ring.dev[slot], ring.host[slot], k * SZ, cudaMemcpyHostToDevice, c->stream));
        points_from_file_kernel<G2><<<grid_for(k, 128), 128, 0, c->stream>>>(ring.dev[slot], k, d_pos, (size_t)f0, out_base, d_bad);
        c->launches++;
        B2_CUDA(cudaGetLastError());
        B2_CUDA(cudaEventRecord(ring.done[slot], c->stream));
        slot = (slot + 1) % Ring::SLOTS;
    }
}

// positions of the set entries of a density bitmap built from the CSR columns of `m` (+ the first n_always variables)
std::vector<uint32_t> dense_positions(const b2g16_csr& m, uint64_t n_vars, uint64_t n_always, uint64_t expect, const char* what) {
    B2_REQUIRE(m.row_ptr && (m.nnz == 0 || m.col), B2G16_E_INVALID_ARG, std::string(what) + ": a density-filtered query needs the constraint matrices");
    std::vector<uint8_t> dens(n_vars, 0);
    for (uint64_t j = 0; j < n_always; ++j) dens[j] = 1;
    for (uint64_t k = 0; k < m.nnz; ++k) { B2_REQUIRE(m.col[k] < n_vars, B2G16_E_INVALID_ARG, "constraint matrix: column index out of range"); dens[m.col[k]] = 1; }
    std::vector<uint32_t> pos;
    pos.reserve(expect);
    for (uint64_t j = 0; j < n_vars; ++j) if (dens[j]) pos.push_back((uint32_t)j);
    B2_REQUIRE(pos.size() == expect, B2G16_E_INVALID_ARG, std::string(what) + ": filtered query does not match the constraint system's density");
    return pos;
}

void file_g1_to_abi(const uint8_t* f, uint8_t* a) {
    if (f[0] & 0x40) { memset(a, 0, 96); return; }
    for (int k = 0; k < 2; ++k) for (int i = 0; i < 48; ++i) a[48 * k + i] = f[48 * k + 47 - i];
}
void file_g2_to_abi(const uint8_t* f, uint8_t* a) {
    if (f[0] & 0x40) { memset(a, 0, 192); return; }
    const int src[4] = {1, 0, 3, 2};                                    // file order c1 c0 -> ABI c0 c1
    for (int k = 0; k < 4; ++k) for (int i = 0; i < 48; ++i) a[48 * k + i] = f[48 * src[k] + 47 - i];
}

}  // namespace
}  // namespace b2

using namespace b2;

extern "C" {

int b2g16_params_map(const char* path, uint64_t counts[6], uint64_t offsets[6]) {
    if (!path || !counts || !offsets) return B2G16_E_INVALID_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return B2G16_E_INVALID_ARG;
    FileIndex ix;
    const bool ok = index_file(f, &ix);
    fclose(f);
    if (!ok) return B2G16_E_INVALID_ARG;
    for (int q = 0; q < 6; ++q) { counts[q] = ix.count[q]; offsets[q] = ix.offset[q]; }
    return B2G16_OK;
}

int b2g16_pk_load_file(b2g16_ctx* hctx, const char* path, const b2g16_r1cs_view* cs, b2g16_pk** out, uint8_t* vk_out, size_t vk_bytes) {
    Ctx* c = reinterpret_cast<Ctx*>(hctx);
    if (!c || !path || !cs || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    return guard(c, [&] {
        FILE* f = fopen(path, "rb");
        B2_REQUIRE(f, B2G16_E_INVALID_ARG, std::string("cannot open ") + path);
        struct Closer { FILE* f; ~Closer() { fclose(f); } } closer{f};
        FileIndex ix;
        B2_REQUIRE(index_file(f, &ix), B2G16_E_INVALID_ARG, "parameter file: truncated or inconsistent length prefixes");
        const uint64_t nv = 1 + cs->n_in + cs->n_aux, m = 1ull << cs->log_m;
        B2_REQUIRE(ix.count[0] == cs->n_in + 1 && ix.count[1] + 1 == m && ix.count[2] == cs->n_aux, B2G16_E_INVALID_ARG,
                   "parameter file: parameters of another circuit (ic / h / l lengths)");
        B2_REQUIRE(ix.count[3] <= nv && ix.count[4] <= nv && ix.count[5] <= nv && ix.count[4] == ix.count[5], B2G16_E_INVALID_ARG,
                   "parameter file: a / b query lengths do not fit the constraint system");
        B2_REQUIRE(!vk_out || vk_bytes == 672 + 96 * ix.count[0], B2G16_E_INVALID_ARG, "vk buffer must be 672 + 96 n_ic bytes");
        DevFlag bad(c);
        b2g16_pk* pk = new_pk_with_cs(c, cs, bad.ptr());
        try {
            // header: six single points (864 B) and the IC points, converted on the host (a few KB)
            uint8_t hdr[864], abi[864];
            B2_REQUIRE(fseeko(f, 0, SEEK_SET) == 0 && fread(hdr, 1, sizeof hdr, f) == sizeof hdr, B2G16_E_INVALID_ARG, "parameter file: header");
            for (int o : {0, 96, 192, 384, 576, 672}) B2_REQUIRE(!(hdr[o] & 0x80), B2G16_E_NOT_CANONICAL, "parameter file: compressed point in the header");
            file_g1_to_abi(hdr, abi); file_g1_to_abi(hdr + 96, abi + 96); file_g2_to_abi(hdr + 192, abi + 192);           // alpha_g1 beta_g1 beta_g2
            file_g2_to_abi(hdr + 384, abi + 384); file_g1_to_abi(hdr + 576, abi + 576); file_g2_to_abi(hdr + 672, abi + 672);  // gamma_g2 delta_g1 delta_g2
            pk_set_single_points(c, pk, abi, abi + 96, abi + 576, abi + 192, abi + 672, bad.ptr());
            if (vk_out) {
                memcpy(vk_out, abi, 96); memcpy(vk_out + 96, abi + 192, 192); memcpy(vk_out + 288, abi + 384, 192); memcpy(vk_out + 480, abi + 672, 192);
                std::vector<uint8_t> ic(96 * (size_t)ix.count[0]);
                B2_REQUIRE(fseeko(f, (off_t)ix.offset[0], SEEK_SET) == 0 && fread(ic.data(), 96, ix.count[0], f) == ix.count[0], B2G16_E_INVALID_ARG, "parameter file: ic");
                for (uint64_t i = 0; i < ix.count[0]; ++i) file_g1_to_abi(ic.data() + 96 * i, vk_out + 672 + 96 * i);
            }
            // the five queries, streamed
            const size_t chunk = (size_t)(c->key_chunk_mib > 0 ? c->key_chunk_mib : 32) << 20;
            Ring ring(chunk - chunk % 192);
            int slot = 0;
            auto dense = [&](int q, const b2g16_csr& mat, uint64_t n_always, const char* what, DevBuf<uint32_t>& d_pos) -> const uint32_t* {
                if (ix.count[q] == nv) return nullptr;
                std::vector<uint32_t> pos = dense_positions(mat, nv, n_always, ix.count[q], what);
                d_pos.alloc(pos.size() + 1, c->stream);
                B2_CUDA(cudaMemcpyAsync(d_pos.get(), pos.data(), pos.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
                B2_CUDA(cudaStreamSynchronize(c->stream));                // pos is a local
                return d_pos.get();
            };
            B2_CUDA(cudaMalloc((void**)&pk->h, m * sizeof(G1Affine)));
            stream_query<false>(c, f, ix.offset[1], ix.count[1], nullptr, (uint8_t*)pk->h, ring, bad.ptr(), slot);
            B2_CUDA(cudaMalloc((void**)&pk->l, (cs->n_aux + 1) * sizeof(G1Affine)));
            stream_query<false>(c, f, ix.offset[2], ix.count[2], nullptr, (uint8_t*)pk->l, ring, bad.ptr(), slot);
            DevBuf<uint32_t> pos_a, pos_b;
            B2_CUDA(cudaMalloc((void**)&pk->a, (nv + 1) * sizeof(G1Affine)));
            B2_CUDA(cudaMemsetAsync(pk->a, 0, nv * sizeof(G1Affine), c->stream));        // absent entries = identity
            const uint32_t* pa = dense(3, cs->A, cs->n_in + 1, "a query", pos_a);
            stream_query<false>(c, f, ix.offset[3], ix.count[3], pa, (uint8_t*)pk->a, ring, bad.ptr(), slot);
            B2_CUDA(cudaMalloc((void**)&pk->b_g1, (nv + 1) * sizeof(G1Affine)));
            B2_CUDA(cudaMemsetAsync(pk->b_g1, 0, nv * sizeof(G1Affine), c->stream));
            const uint32_t* pb = dense(4, cs->B, 0, "b_g1 query", pos_b);
            stream_query<false>(c, f, ix.offset[4], ix.count[4], pb, (uint8_t*)pk->b_g1, ring, bad.ptr(), slot);
            B2_CUDA(cudaMalloc((void**)&pk->b_g2, (nv + 1) * sizeof(G2Affine)));
            B2_CUDA(cudaMemsetAsync(pk->b_g2, 0, nv * sizeof(G2Affine), c->stream));
            stream_query<true>(c, f, ix.offset[5], ix.count[5], pb, (uint8_t*)pk->b_g2, ring, bad.ptr(), slot);
            bad.check("parameter file");
        } catch (...) { cudaStreamSynchronize(c->stream); free_pk(pk); throw; }
        *out = pk;
    });
}

}  // extern "C"
