// Parameter-cache entry names and the small cache files.
//
// Restates the path logic of libs/storage/include/nil/filecoin/storage/proofs/core/parameter_cache.hpp:
//   :50-56   VERSION = 28, FIL_PROOFS_PARAMETER_CACHE, /var/tmp/filecoin-proof-parameters/, extensions params / meta / vk
//   :69-94   <dir>/v28-<identifier>.<ext>
//   :108-123 cache_entry_metadata { sector_size } as a one-field JSON object (serde_json in the original)
//   :131-145 the verifying key cached alone in a .vk file
//   :166-173 identifier = <cache_prefix>-<hex sha256(parameter-set identifier string)>
// Host-only; nothing here touches the device.  The .vk layout is the vk block that opens a .params file
// (params_io.cu): alpha_g1 beta_g1 beta_g2 gamma_g2 delta_g1 delta_g2 u32_be n_ic ic[].
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <string>

#include "../../include/b2g16.h"

namespace b2io {
// shared with params_io.cu
bool read_vk_block(FILE* f, b2g16_params* p, int* rc);
bool write_vk_block(FILE* f, const b2g16_params* p);
}  // namespace b2io

namespace {

// ---- SHA-256 (FIPS 180-4), one-shot ----------------------------------------------------------
struct Sha256 {
    uint32_t h[8];
    static uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
    static void block(uint32_t* h, const uint8_t* p) {
        static const uint32_t K[64] = {
            0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, 0xd807aa98u, 0x12835b01u, 0x243185beu,
            0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, 0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau,
            0x5cb0a9dcu, 0x76f988dau, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, 0x27b70a85u,
            0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, 0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u,
            0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, 0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu,
            0x682e6ff3u, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};
        uint32_t w[64];
        for (int i = 0; i < 16; ++i) w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
        for (int i = 16; i < 64; ++i) {
            const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3), s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16] + s0 + w[i - 7] + s1;
        }
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        for (int i = 0; i < 64; ++i) {
            const uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
            const uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
            hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    }
    static void digest(const uint8_t* data, size_t len, uint8_t out[32]) {
        uint32_t h[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au, 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
        size_t i = 0;
        for (; i + 64 <= len; i += 64) block(h, data + i);
        uint8_t tail[128] = {0};
        const size_t rem = len - i;
        if (rem) memcpy(tail, data + i, rem);
        tail[rem] = 0x80;
        const size_t padded = rem + 9 <= 64 ? 64 : 128;
        const uint64_t bits = (uint64_t)len * 8;
        for (int k = 0; k < 8; ++k) tail[padded - 1 - k] = (uint8_t)(bits >> (8 * k));
        block(h, tail);
        if (padded == 128) block(h, tail + 64);
        for (int k = 0; k < 8; ++k) { out[4 * k] = (uint8_t)(h[k] >> 24); out[4 * k + 1] = (uint8_t)(h[k] >> 16); out[4 * k + 2] = (uint8_t)(h[k] >> 8); out[4 * k + 3] = (uint8_t)h[k]; }
    }
};

int put(const std::string& s, char* out, size_t cap) {
    if (!out || cap < s.size() + 1) return B2G16_E_INVALID_ARG;
    memcpy(out, s.c_str(), s.size() + 1);
    return B2G16_OK;
}
std::string cache_dir() {
    const char* e = getenv("FIL_PROOFS_PARAMETER_CACHE");                // parameter_cache.hpp:52,65-67
    std::string d = (e && *e) ? e : "/var/tmp/filecoin-proof-parameters/";  // :53 (the settings default of the original)
    while (d.size() > 1 && d.back() == '/') d.pop_back();
    return d;
}

}  // namespace

extern "C" {

int b2g16_sha256(const uint8_t* data, size_t len, uint8_t out32[32]) {
    if ((!data && len) || !out32) return B2G16_E_INVALID_ARG;
    Sha256::digest(data, len, out32);
    return B2G16_OK;
}

int b2g16_cache_dir(char* out, size_t cap) { return put(cache_dir(), out, cap); }

int b2g16_cache_identifier(const char* cache_prefix, const char* parameter_set_identifier, char* out, size_t cap) {
    if (!cache_prefix || !parameter_set_identifier) return B2G16_E_INVALID_ARG;
    uint8_t d[32];
    Sha256::digest((const uint8_t*)parameter_set_identifier, strlen(parameter_set_identifier), d);
    static const char* hex = "0123456789abcdef";
    std::string s = std::string(cache_prefix) + "-";
    for (int i = 0; i < 32; ++i) { s.push_back(hex[d[i] >> 4]); s.push_back(hex[d[i] & 15]); }
    return put(s, out, cap);
}

int b2g16_cache_path(const char* identifier, int kind, char* out, size_t cap) {
    static const char* ext[3] = {"params", "vk", "meta"};                // :54-56
    if (!identifier || kind < 0 || kind > 2) return B2G16_E_INVALID_ARG;
    return put(cache_dir() + "/v" + std::to_string(B2G16_PARAMETER_CACHE_VERSION) + "-" + identifier + "." + ext[kind], out, cap);
}

int b2g16_meta_write(const char* path, uint64_t sector_size) {
    if (!path) return B2G16_E_INVALID_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return B2G16_E_INVALID_ARG;
    const bool ok = fprintf(f, "{\"sector_size\":%llu}", (unsigned long long)sector_size) > 0;
    return (fclose(f) == 0 && ok) ? B2G16_OK : B2G16_E_INTERNAL;
}

int b2g16_meta_read(const char* path, uint64_t* sector_size) {
    if (!path || !sector_size) return B2G16_E_INVALID_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return B2G16_E_INVALID_ARG;
    char buf[256];
    const size_t n = fread(buf, 1, sizeof buf - 1, f);
    fclose(f);
    buf[n] = 0;
    const char* k = strstr(buf, "\"sector_size\"");
    if (!k) return B2G16_E_INVALID_ARG;
    k = strchr(k, ':');
    if (!k) return B2G16_E_INVALID_ARG;
    ++k;
    while (*k == ' ' || *k == '\t' || *k == '\n' || *k == '\r') ++k;
    if (*k < '0' || *k > '9') return B2G16_E_INVALID_ARG;
    *sector_size = strtoull(k, nullptr, 10);
    return B2G16_OK;
}

int b2g16_vk_file_write(const char* path, const b2g16_params* p) {
    if (!path || !p || (p->n_ic >> 32)) return B2G16_E_INVALID_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return B2G16_E_INVALID_ARG;
    const bool ok = b2io::write_vk_block(f, p);
    return (fclose(f) == 0 && ok) ? B2G16_OK : B2G16_E_INTERNAL;
}

int b2g16_vk_file_read(const char* path, b2g16_params** out) {
    if (!path || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return B2G16_E_INVALID_ARG;
    b2g16_params* p = new (std::nothrow) b2g16_params();
    if (!p) { fclose(f); return B2G16_E_OOM; }
    memset(p, 0, sizeof *p);
    int rc = B2G16_OK;
    b2io::read_vk_block(f, p, &rc);
    fclose(f);
    if (rc != B2G16_OK) { b2g16_params_free(p); return rc; }
    *out = p;
    return B2G16_OK;
}

}  // extern "C"
