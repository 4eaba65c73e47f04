// Parameter-file and multi-proof byte formats.
//
// The reference caches Groth16 parameters on disk as  <cache>/v28-<prefix>-<sha256(id)>.params
// and the verification key as .vk (libs/storage/.../core/parameter_cache.hpp:50-94,125-153) and
// memory-maps them through offset tables for h / l / a / b_g1 / b_g2
// (.../core/crypto/mapped_scheme_params.hpp:52-83).  The serialisers themselves live in the
// un-vendored crypto3-zk-marshalling module; the layout restated here is the published
// bellperson `Parameters::write` one that Filecoin's parameter files use:
//   vk:  alpha_g1(96) beta_g1(96) beta_g2(192) gamma_g2(192) delta_g1(96) delta_g2(192)
//        u32_be n_ic, n_ic * G1(96)
//   then five u32_be-length-prefixed arrays: h (G1), l (G1), a (G1), b_g1 (G1), b_g2 (G2)
// with points uncompressed big-endian (G2 coordinates c1 then c0; bit 6 of byte 0 = infinity).
// Both this and the C ABI's point format are canonical (non-Montgomery) integers, so conversion
// is a byte shuffle; range / curve validation happens on the device when the key is loaded.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <new>
#include <thread>
#include <vector>

#include "../../include/b2g16.h"

namespace {

// reverse 48 bytes (BE <-> LE field element) a word at a time
inline void rev48(const uint8_t* in, uint8_t* out) {
    uint64_t w[6];
    memcpy(w, in, 48);
    uint64_t r[6];
    for (int i = 0; i < 6; ++i) r[i] = __builtin_bswap64(w[5 - i]);
    memcpy(out, r, 48);
}
inline bool all_zero(const uint8_t* a, size_t n) {
    uint64_t acc = 0, w;
    for (size_t i = 0; i < n; i += 8) { memcpy(&w, a + i, 8); acc |= w; }
    return acc == 0;
}

// file G1 (x y, BE) -> ABI (x y, LE); infinity -> all zero.  false on a malformed flag byte.  In place is fine (in == out).
bool g1_file_to_abi(const uint8_t* f, uint8_t* a) {
    if (f[0] & 0x80) return false;
    if (f[0] & 0x40) { memset(a, 0, 96); return true; }
    rev48(f, a); rev48(f + 48, a + 48); return true;
}
void g1_abi_to_file(const uint8_t* a, uint8_t* f) {
    if (all_zero(a, 96)) { memset(f, 0, 96); f[0] = 0x40; return; }
    rev48(a, f); rev48(a + 48, f + 48);
}
// file G2 (x.c1 x.c0 y.c1 y.c0, BE) -> ABI (x.c0 x.c1 y.c0 y.c1, LE)
bool g2_file_to_abi(const uint8_t* f, uint8_t* a) {
    if (f[0] & 0x80) return false;
    if (f[0] & 0x40) { memset(a, 0, 192); return true; }
    uint8_t t[192];
    rev48(f, t + 48); rev48(f + 48, t); rev48(f + 96, t + 144); rev48(f + 144, t + 96);
    memcpy(a, t, 192); return true;
}
void g2_abi_to_file(const uint8_t* a, uint8_t* f) {
    if (all_zero(a, 192)) { memset(f, 0, 192); f[0] = 0x40; return; }
    uint8_t t[192];
    rev48(a + 48, t); rev48(a, t + 48); rev48(a + 144, t + 96); rev48(a + 96, t + 144);
    memcpy(f, t, 192);
}

struct File {
    FILE* f;
    explicit File(const char* path, const char* mode) : f(fopen(path, mode)) {}
    ~File() { if (f) fclose(f); }
};

bool read_u32_be(FILE* f, uint64_t* v) { uint8_t b[4]; if (fread(b, 1, 4, f) != 4) return false; *v = ((uint32_t)b[0] << 24) | (b[1] << 16) | (b[2] << 8) | b[3]; return true; }
bool write_u32_be(FILE* f, uint64_t v) { uint8_t b[4] = {(uint8_t)(v >> 24), (uint8_t)(v >> 16), (uint8_t)(v >> 8), (uint8_t)v}; return fwrite(b, 1, 4, f) == 4; }

// A 32 GiB-sector key is ~80 GB of points: split the byte shuffles over the host cores.
template <class Fn> void parallel_ranges(uint64_t n, Fn fn) {
    unsigned t = std::thread::hardware_concurrency();
    t = t ? (t > 32 ? 32 : t) : 1;
    if (n < (1u << 16) || t == 1) { fn(0, n); return; }
    std::vector<std::thread> th;
    const uint64_t per = (n + t - 1) / t;
    for (unsigned k = 0; k < t; ++k) {
        const uint64_t lo = k * per, hi = lo + per < n ? lo + per : n;
        if (lo >= hi) break;
        th.emplace_back([=] { fn(lo, hi); });
    }
    for (auto& x : th) x.join();
}

// one array: n points -> freshly malloc'ed ABI buffer (read whole, converted in place)
int read_points(FILE* f, uint64_t n, int g2, uint8_t** out) {
    const size_t sz = g2 ? 192 : 96;
    // the length prefix comes from the file: refuse one the file cannot hold before allocating for it
    const long at = ftell(f);
    if (at < 0 || fseek(f, 0, SEEK_END) != 0) return B2G16_E_INVALID_ARG;
    const long end = ftell(f);
    if (end < 0 || fseek(f, at, SEEK_SET) != 0 || (uint64_t)(end - at) < n * sz) return B2G16_E_INVALID_ARG;
    uint8_t* buf = (uint8_t*)malloc(n ? n * sz : 1);
    if (!buf) return B2G16_E_OOM;
    if (n && fread(buf, sz, n, f) != n) { free(buf); return B2G16_E_INVALID_ARG; }
    std::atomic<int> bad(0);
    parallel_ranges(n, [&](uint64_t lo, uint64_t hi) {
        for (uint64_t i = lo; i < hi; ++i) {
            uint8_t* p = buf + i * sz;
            if (!(g2 ? g2_file_to_abi(p, p) : g1_file_to_abi(p, p))) { bad.store(1); return; }
        }
    });
    if (bad.load()) { free(buf); return B2G16_E_NOT_CANONICAL; }
    *out = buf;
    return B2G16_OK;
}
bool write_points(FILE* f, const uint8_t* abi, uint64_t n, int g2) {
    const size_t sz = g2 ? 192 : 96;
    const uint64_t CH = 1u << 20;                                  // points per write
    std::vector<uint8_t> rec((size_t)((n < CH ? n : CH) * sz));
    for (uint64_t i = 0; i < n; i += CH) {
        const uint64_t k = n - i < CH ? n - i : CH;
        parallel_ranges(k, [&](uint64_t lo, uint64_t hi) {
            for (uint64_t j = lo; j < hi; ++j) {
                if (g2) g2_abi_to_file(abi + (i + j) * sz, rec.data() + j * sz); else g1_abi_to_file(abi + (i + j) * sz, rec.data() + j * sz);
            }
        });
        if (fwrite(rec.data(), sz, k, f) != k) return false;
    }
    return true;
}

}  // namespace

// The vk block that opens a .params file and is the whole of a .vk file (param_cache.cu):
// alpha_g1 beta_g1 beta_g2 gamma_g2 delta_g1 delta_g2 u32_be n_ic ic[]
namespace b2io {
bool read_vk_block(FILE* f, b2g16_params* p, int* rc) {
    uint8_t hdr[864];
    if (fread(hdr, 1, sizeof hdr, f) != sizeof hdr) { *rc = B2G16_E_INVALID_ARG; return false; }
    if (!(g1_file_to_abi(hdr, p->alpha_g1) && g1_file_to_abi(hdr + 96, p->beta_g1) && g2_file_to_abi(hdr + 192, p->beta_g2) &&
          g2_file_to_abi(hdr + 384, p->gamma_g2) && g1_file_to_abi(hdr + 576, p->delta_g1) && g2_file_to_abi(hdr + 672, p->delta_g2))) {
        *rc = B2G16_E_NOT_CANONICAL; return false;
    }
    if (!read_u32_be(f, &p->n_ic)) { *rc = B2G16_E_INVALID_ARG; return false; }
    *rc = read_points(f, p->n_ic, 0, &p->ic);
    return *rc == B2G16_OK;
}
bool write_vk_block(FILE* f, const b2g16_params* p) {
    uint8_t hdr[864];
    g1_abi_to_file(p->alpha_g1, hdr); g1_abi_to_file(p->beta_g1, hdr + 96); g2_abi_to_file(p->beta_g2, hdr + 192);
    g2_abi_to_file(p->gamma_g2, hdr + 384); g1_abi_to_file(p->delta_g1, hdr + 576); g2_abi_to_file(p->delta_g2, hdr + 672);
    return fwrite(hdr, 1, sizeof hdr, f) == sizeof hdr && write_u32_be(f, p->n_ic) && write_points(f, p->ic, p->n_ic, 0);
}
}  // namespace b2io

extern "C" {

int b2g16_params_free(b2g16_params* p) {
    if (!p) return B2G16_OK;
    free(p->ic); free(p->h); free(p->l); free(p->a); free(p->b_g1); free(p->b_g2);
    delete p;
    return B2G16_OK;
}

int b2g16_params_read(const char* path, b2g16_params** out) {
    if (!path || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    File fh(path, "rb");
    if (!fh.f) return B2G16_E_INVALID_ARG;
    b2g16_params* p = new (std::nothrow) b2g16_params();
    if (!p) return B2G16_E_OOM;
    memset(p, 0, sizeof *p);
    int rc = B2G16_OK;
    b2io::read_vk_block(fh.f, p, &rc);                   // header + IC points
    struct { uint64_t* n; uint8_t** buf; int g2; } arr[5] = {{&p->n_h, &p->h, 0}, {&p->n_l, &p->l, 0},
                                                            {&p->n_a, &p->a, 0}, {&p->n_b_g1, &p->b_g1, 0}, {&p->n_b_g2, &p->b_g2, 1}};
    for (int i = 0; i < 5 && rc == B2G16_OK; ++i) {
        if (!read_u32_be(fh.f, arr[i].n)) { rc = B2G16_E_INVALID_ARG; break; }
        rc = read_points(fh.f, *arr[i].n, arr[i].g2, arr[i].buf);
    }
    if (rc != B2G16_OK) { b2g16_params_free(p); return rc; }
    *out = p;
    return B2G16_OK;
}

int b2g16_params_write(const char* path, const b2g16_params* p) {
    if (!path || !p) return B2G16_E_INVALID_ARG;
    const uint64_t counts[6] = {p->n_ic, p->n_h, p->n_l, p->n_a, p->n_b_g1, p->n_b_g2};
    for (uint64_t n : counts) if (n >> 32) return B2G16_E_INVALID_ARG;   // u32 length prefixes
    File fh(path, "wb");
    if (!fh.f) return B2G16_E_INVALID_ARG;
    bool ok = b2io::write_vk_block(fh.f, p);
    const uint8_t* bufs[5] = {p->h, p->l, p->a, p->b_g1, p->b_g2};
    for (int i = 0; i < 5 && ok; ++i) ok = write_u32_be(fh.f, counts[i + 1]) && write_points(fh.f, bufs[i], counts[i + 1], i == 4);
    return ok ? B2G16_OK : B2G16_E_INTERNAL;
}

// .vk file = the vk block of a .params file alone (parameter_cache.hpp:56,131-145) -> the layout
// b2g16_vk_prepare takes: alpha_g1 beta_g2 gamma_g2 delta_g2 ic[]   (672 + 96 n_ic bytes)
int b2g16_params_vk_bytes(const b2g16_params* p, uint8_t* vk_out, size_t vk_bytes) {
    if (!p || !vk_out || vk_bytes != 672 + 96 * (size_t)p->n_ic) return B2G16_E_INVALID_ARG;
    memcpy(vk_out, p->alpha_g1, 96); memcpy(vk_out + 96, p->beta_g2, 192); memcpy(vk_out + 288, p->gamma_g2, 192);
    memcpy(vk_out + 480, p->delta_g2, 192); memcpy(vk_out + 672, p->ic, 96 * (size_t)p->n_ic);
    return B2G16_OK;
}

// Density-filtered queries (bellman generator: points at infinity are dropped from a / b_g1 / b_g2;
// its prover skips them through the a_aux / b_input / b_aux density bitmaps) <-> the dense layout of
// scheme_params.hpp:46-66 that b2g16_pk_view takes.  A column's polynomial is zero exactly when the
// variable occurs in no row, so the bitmap is the column occupancy of the matrix.
int b2g16_query_density(const b2g16_csr* m, uint64_t n_vars, uint64_t n_always, uint8_t* density_out, uint64_t* count_out) {
    if (!m || !density_out || (m->nnz && !m->col) || n_always > n_vars) return B2G16_E_INVALID_ARG;
    memset(density_out, 0, n_vars);
    // The a query carries the input-consistency rows (u_j += L[n_cons + j] for j <= n_in), so the
    // polynomial of ONE and of every public input is never zero there, whether or not the variable occurs in a row
    // of A -- bellman keeps all n_in + 1 of them (its a_inputs density is full).  n_always = n_in + 1 for a, 0 for b.
    memset(density_out, 1, n_always);
    for (uint64_t k = 0; k < m->nnz; ++k) {
        if (m->col[k] >= n_vars) return B2G16_E_INVALID_ARG;
        density_out[m->col[k]] = 1;
    }
    if (count_out) {
        uint64_t c = 0;
        for (uint64_t j = 0; j < n_vars; ++j) c += density_out[j];
        *count_out = c;
    }
    return B2G16_OK;
}
int b2g16_csr_density(const b2g16_csr* m, uint64_t n_vars, uint8_t* density_out, uint64_t* count_out) {
    return b2g16_query_density(m, n_vars, 0, density_out, count_out);
}

int b2g16_query_expand(const uint8_t* filtered, uint64_t n_filtered, const uint8_t* density, uint64_t n_vars, size_t point_bytes,
                       uint8_t* dense_out) {
    if ((!filtered && n_filtered) || !density || !dense_out || (point_bytes != 96 && point_bytes != 192)) return B2G16_E_INVALID_ARG;
    uint64_t set = 0;
    for (uint64_t j = 0; j < n_vars; ++j) set += density[j] != 0;
    if (set != n_filtered) return B2G16_E_INVALID_ARG;          // file and constraint system disagree: refuse before writing
    uint64_t k = 0;
    for (uint64_t j = 0; j < n_vars; ++j) {
        if (density[j]) memcpy(dense_out + j * point_bytes, filtered + (k++) * point_bytes, point_bytes);
        else memset(dense_out + j * point_bytes, 0, point_bytes);
    }
    return B2G16_OK;
}

int b2g16_query_filter(const uint8_t* dense, uint64_t n_vars, size_t point_bytes, uint8_t* filtered_out, uint8_t* density_out,
                       uint64_t* count_out) {
    if ((!dense && n_vars) || (point_bytes != 96 && point_bytes != 192)) return B2G16_E_INVALID_ARG;
    uint64_t k = 0;
    for (uint64_t j = 0; j < n_vars; ++j) {
        const uint8_t* p = dense + j * point_bytes;
        uint8_t any = 0;
        for (size_t b = 0; b < point_bytes; ++b) any |= p[b];
        if (density_out) density_out[j] = any != 0;
        if (any) {
            if (filtered_out) memcpy(filtered_out + k * point_bytes, p, point_bytes);
            ++k;
        }
    }
    if (count_out) *count_out = k;
    return B2G16_OK;
}

}  // extern "C"
