// Several GPUs of one box behind the C ABI: one process, one context per device, host threads for the fan-out.
//
// The reference's caller is a single C++ process: compound_proof::circuit_proofs
// (libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137) hands ALL partition
// circuits of a proof to one prover call, and seal_commit_phase2 / generate_window_post
// (libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:298-308, api/post.hpp:305-348) write the proofs back in
// partition order.  These entry points give that caller every GPU of the box without torch.distributed:
//   * b2g16_prove_batch_multi: partitions -> devices.  Full waves run one partition per device at a time (independent
//     units, the key replicated, no data-path exchange; device i takes a contiguous run of partitions through
//     b2g16_prove_batch, which overlaps the next witness upload with the current proof); the last partial wave splits
//     the devices into groups and each group proves ONE partition in latency mode (every member reduces its slice of the five queries,
//     b2g16_prove_partial; 576 bytes per member come back to the host and are assembled), so no device idles.
//   * b2g16_prove_multi: one proof over all devices in latency mode (a WindowPoSt / WinningPoSt proof is one partition).
// Built on the single-device entry points of include/b2g16.h only; python/b2g16/dist.py is the same schedule for the
// one-process-per-GPU (torchrun) deployment.
#include <cuda_runtime.h>
#include <string.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b2g16.h"

struct b2g16_mctx {
    std::vector<int> devs;
    std::vector<b2g16_ctx*> ctx;
    std::mutex mu;
    std::string last_error;
};
struct b2g16_mpk {
    std::vector<b2g16_pk*> pk;       // one replica per context, same order
    uint64_t shape[5];               // log_m, n_cons, n_in, n_aux, n_vars
};

namespace {

void note_error(b2g16_mctx* m, int slot, int rc, const char* what) {
    std::lock_guard<std::mutex> lock(m->mu);
    if (!m->last_error.empty()) return;                      // keep the first
    const char* inner = (slot >= 0 && slot < (int)m->ctx.size() && m->ctx[slot]) ? b2g16_last_error(m->ctx[slot]) : "";
    m->last_error = std::string(what) + " on device " + (slot >= 0 ? std::to_string(m->devs[slot]) : std::string("?")) + " (slot " + std::to_string(slot) +
                    "): code " + std::to_string(rc) + (inner && *inner ? std::string(": ") + inner : std::string());
}

// fn(slot) on one host thread per context; returns the first non-zero code
template <class Fn> int fan_out(b2g16_mctx* m, Fn fn) {
    const int n = (int)m->ctx.size();
    std::vector<int> rc(n, B2G16_OK);
    std::vector<std::thread> th;
    for (int i = 1; i < n; ++i) th.emplace_back([&, i] { rc[i] = fn(i); });
    rc[0] = fn(0);
    for (auto& t : th) t.join();
    for (int i = 0; i < n; ++i) if (rc[i] != B2G16_OK) return rc[i];
    return B2G16_OK;
}

// this member's share of one proof in latency mode: witness to the device, partial sums back
int partial_on(b2g16_mctx* m, const b2g16_mpk* k, int slot, const uint8_t* assignment32, unsigned part, unsigned parts, uint8_t* partial576) {
    const size_t zb = (size_t)k->shape[4] * 32;
    if (cudaSetDevice(m->devs[slot]) != cudaSuccess) { cudaGetLastError(); note_error(m, slot, B2G16_E_CUDA, "cudaSetDevice"); return B2G16_E_CUDA; }
    uint8_t* dz = nullptr;
    if (cudaMalloc((void**)&dz, zb) != cudaSuccess) { cudaGetLastError(); note_error(m, slot, B2G16_E_OOM, "witness allocation"); return B2G16_E_OOM; }
    int rc = B2G16_OK;
    if (cudaMemcpy(dz, assignment32, zb, cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); rc = B2G16_E_CUDA; note_error(m, slot, rc, "witness upload"); }
    if (rc == B2G16_OK) {
        rc = b2g16_prove_partial(m->ctx[slot], k->pk[slot], dz, part, parts, partial576);
        if (rc != B2G16_OK) note_error(m, slot, rc, "b2g16_prove_partial");
    }
    cudaFree(dz);
    return rc;
}

}  // namespace

extern "C" {

int b2g16_init_multi(const int* devices, int n, b2g16_mctx** out) {
    if (!devices || n <= 0 || n > 64 || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr;
    b2g16_mctx* m = new (std::nothrow) b2g16_mctx();
    if (!m) return B2G16_E_OOM;
    for (int i = 0; i < n; ++i) {
        b2g16_ctx* c = nullptr;
        const int rc = b2g16_init(devices[i], nullptr, &c);          // library-owned stream per context
        if (rc != B2G16_OK) {
            for (b2g16_ctx* x : m->ctx) b2g16_shutdown(x);
            delete m;
            return rc;                                                // B2G16_E_CUDA without a device: no CPU fallback
        }
        m->devs.push_back(devices[i]);
        m->ctx.push_back(c);
    }
    *out = m;
    return B2G16_OK;
}

int b2g16_shutdown_multi(b2g16_mctx* m) {
    if (!m) return B2G16_OK;
    for (b2g16_ctx* c : m->ctx) b2g16_shutdown(c);
    delete m;
    return B2G16_OK;
}

int b2g16_multi_count(const b2g16_mctx* m) { return m ? (int)m->ctx.size() : 0; }
b2g16_ctx* b2g16_multi_ctx(b2g16_mctx* m, int slot) { return (m && slot >= 0 && slot < (int)m->ctx.size()) ? m->ctx[slot] : nullptr; }
const char* b2g16_multi_last_error(b2g16_mctx* m) { return m ? m->last_error.c_str() : "null context"; }

int b2g16_mpk_free(b2g16_mctx* m, b2g16_mpk* k) {
    if (!m || !k) return B2G16_OK;
    for (size_t i = 0; i < k->pk.size(); ++i) if (k->pk[i]) b2g16_pk_free(m->ctx[i], k->pk[i]);
    delete k;
    return B2G16_OK;
}

static int finish_mpk(b2g16_mctx* m, b2g16_mpk* k, int rc, b2g16_mpk** out) {
    if (rc == B2G16_OK) rc = b2g16_pk_shape(k->pk[0], k->shape);
    if (rc != B2G16_OK) { b2g16_mpk_free(m, k); return rc; }
    *out = k;
    return B2G16_OK;
}

int b2g16_mpk_load(b2g16_mctx* m, const b2g16_pk_view* v, b2g16_mpk** out) {
    if (!m || !v || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr; m->last_error.clear();
    b2g16_mpk* k = new (std::nothrow) b2g16_mpk();
    if (!k) return B2G16_E_OOM;
    k->pk.assign(m->ctx.size(), nullptr);
    const int rc = fan_out(m, [&](int i) {
        const int r = b2g16_pk_load(m->ctx[i], v, &k->pk[i]);
        if (r != B2G16_OK) note_error(m, i, r, "b2g16_pk_load");
        return r;
    });
    return finish_mpk(m, k, rc, out);
}

int b2g16_mpk_generate_synth(b2g16_mctx* m, uint32_t log_m, uint64_t n_in, uint64_t n_free, uint64_t seed, const uint8_t toxic[160], b2g16_mpk** out,
                             uint8_t* vk_out) {
    if (!m || !toxic || !out) return B2G16_E_INVALID_ARG;
    *out = nullptr; m->last_error.clear();
    b2g16_mpk* k = new (std::nothrow) b2g16_mpk();
    if (!k) return B2G16_E_OOM;
    k->pk.assign(m->ctx.size(), nullptr);
    const int rc = fan_out(m, [&](int i) {
        const int r = b2g16_pk_generate_synth(m->ctx[i], log_m, n_in, n_free, seed, toxic, &k->pk[i], i == 0 ? vk_out : nullptr);
        if (r != B2G16_OK) note_error(m, i, r, "b2g16_pk_generate_synth");
        return r;
    });
    return finish_mpk(m, k, rc, out);
}

b2g16_pk* b2g16_mpk_device_key(const b2g16_mpk* k, int slot) { return (k && slot >= 0 && slot < (int)k->pk.size()) ? k->pk[slot] : nullptr; }

int b2g16_prove_multi(b2g16_mctx* m, const b2g16_mpk* k, const uint8_t* assignment32, const uint8_t r32[32], const uint8_t s32[32], int priority,
                      uint8_t proof_out[192]) {
    (void)priority;
    if (!m || !k || !assignment32 || !r32 || !s32 || !proof_out || k->pk.size() != m->ctx.size()) return B2G16_E_INVALID_ARG;
    m->last_error.clear();
    const unsigned parts = (unsigned)m->ctx.size();
    if (parts == 1) {
        const int rc = b2g16_prove(m->ctx[0], k->pk[0], assignment32, r32, s32, priority, proof_out, nullptr);
        if (rc != B2G16_OK) note_error(m, 0, rc, "b2g16_prove");
        return rc;
    }
    std::vector<uint8_t> partials((size_t)parts * 576);
    int rc = fan_out(m, [&](int i) { return partial_on(m, k, i, assignment32, (unsigned)i, parts, partials.data() + (size_t)i * 576); });
    if (rc != B2G16_OK) return rc;
    rc = b2g16_prove_assemble(m->ctx[0], k->pk[0], partials.data(), parts, r32, s32, proof_out);
    if (rc != B2G16_OK) note_error(m, 0, rc, "b2g16_prove_assemble");
    return rc;
}

// which contexts prove partition `p` of the last, partial wave: `rem` partitions over `g` contexts, group sizes differ by at most one
static void remainder_group(int g, int rem, int p, int* first, int* size) {
    const int base = g / rem, extra = g % rem;
    *size = base + (p < extra ? 1 : 0);
    *first = p * base + (p < extra ? p : extra);
}

// The schedule b2g16_prove_batch_multi follows, as data (host-only; lets the plan be checked without a device):
// plan[3 p .. 3 p + 2] = { wave, first device slot, number of slots } of partition p.
int b2g16_multi_schedule(int n_devices, size_t n_partitions, uint32_t* plan) {
    if (n_devices <= 0 || (n_partitions && !plan)) return B2G16_E_INVALID_ARG;
    const size_t waves = n_partitions / (size_t)n_devices;
    const int rem = (int)(n_partitions % (size_t)n_devices);
    for (size_t p = 0; p < n_partitions; ++p) {
        // full waves: device i proves the contiguous run [i waves, (i + 1) waves) -- one b2g16_prove_batch per device, which
        // uploads the next witness while the current partition is proved
        if (p < waves * (size_t)n_devices) { plan[3 * p] = (uint32_t)(p % waves); plan[3 * p + 1] = (uint32_t)(p / waves); plan[3 * p + 2] = 1; }
        else {
            int first, size;
            remainder_group(n_devices, rem, (int)(p - waves * (size_t)n_devices), &first, &size);
            plan[3 * p] = (uint32_t)waves; plan[3 * p + 1] = (uint32_t)first; plan[3 * p + 2] = (uint32_t)size;
        }
    }
    return B2G16_OK;
}

int b2g16_prove_batch_multi(b2g16_mctx* m, const b2g16_mpk* k, const uint8_t* assignments32, size_t n_partitions, const uint8_t* r32s,
                            const uint8_t* s32s, int priority, uint8_t* proofs_out) {
    if (!m || !k || k->pk.size() != m->ctx.size() || (n_partitions && (!assignments32 || !r32s || !s32s || !proofs_out))) return B2G16_E_INVALID_ARG;
    m->last_error.clear();
    const int g = (int)m->ctx.size();
    const size_t zb = (size_t)k->shape[4] * 32;
    const size_t waves = n_partitions / (size_t)g;
    const int rem = (int)(n_partitions % (size_t)g);
    std::vector<uint8_t> partials((size_t)g * 576);              // remainder wave: one slot per context
    int rc = fan_out(m, [&](int i) {
        if (waves) {                                                // full waves: partitions [i waves, (i + 1) waves) on context i
            const size_t p = (size_t)i * waves;
            const int r = b2g16_prove_batch(m->ctx[i], k->pk[i], assignments32 + p * zb, waves, r32s + 32 * p, s32s + 32 * p, priority, proofs_out + 192 * p);
            if (r != B2G16_OK) { note_error(m, i, r, "b2g16_prove_batch"); return r; }
        }
        for (int q = 0; q < rem; ++q) {                             // remainder wave: context i is member (i - first) of one group
            int first, size;
            remainder_group(g, rem, q, &first, &size);
            if (i < first || i >= first + size) continue;
            const size_t p = waves * (size_t)g + (size_t)q;
            if (size == 1) {
                const int r = b2g16_prove(m->ctx[i], k->pk[i], assignments32 + p * zb, r32s + 32 * p, s32s + 32 * p, priority, proofs_out + 192 * p, nullptr);
                if (r != B2G16_OK) { note_error(m, i, r, "b2g16_prove"); return r; }
            } else {
                const int r = partial_on(m, k, i, assignments32 + p * zb, (unsigned)(i - first), (unsigned)size, partials.data() + (size_t)i * 576);
                if (r != B2G16_OK) return r;
            }
        }
        return (int)B2G16_OK;
    });
    if (rc != B2G16_OK) return rc;
    for (int q = 0; q < rem; ++q) {
        int first, size;
        remainder_group(g, rem, q, &first, &size);
        if (size == 1) continue;
        const size_t p = waves * (size_t)g + (size_t)q;
        rc = b2g16_prove_assemble(m->ctx[first], k->pk[first], partials.data() + (size_t)first * 576, (unsigned)size, r32s + 32 * p, s32s + 32 * p, proofs_out + 192 * p);
        if (rc != B2G16_OK) { note_error(m, first, rc, "b2g16_prove_assemble"); return rc; }
    }
    return B2G16_OK;
}

}  // extern "C"
