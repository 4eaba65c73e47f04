// Host-side planning of the multi-pass Fr NTT (see ntt_kernels.cuh for the pass math).
// Pure C++ so that tests/host_emu can run the same plans through the same phase
// functions on the CPU.  Replaces the sizing logic of crypto3-math
// `make_evaluation_domain` / `basic_radix2_domain` (un-vendored): the
// domain is always radix-2 of size exactly 2^log_n (SURVEY §8c "domain trap").
#pragma once
#include <vector>
#include "ntt_kernels.cuh"

namespace b2 {

struct NttPlan {
    NttDigits digits;                 // k per pass, top to bottom
    int L[NTT_MAX_PASSES];            // block log-size seen by pass p
    int logc[NTT_MAX_PASSES];         // tile columns (log)
    int contiguous[NTT_MAX_PASSES];
    int lo_bits;                      // split of the two-level power tables
};

inline NttPlan ntt_make_plan(int log_n) {
    NttPlan pl{};
    pl.digits.log_n = log_n;
    const int K_STRIDED = 9;          // strided passes keep >= 4 adjacent columns (128 B segments)
    int P = 1;
    while (K_STRIDED * (P - 1) + NTT_TILE_LOG < log_n) ++P;
    pl.digits.n_pass = P;
    int rest = log_n;
    for (int p = 0; p < P; ++p) {
        int k = (rest + (P - p) - 1) / (P - p);          // even split of what is left
        if (p < P - 1 && k > K_STRIDED) k = K_STRIDED;
        if (p == P - 1) k = rest;
        pl.digits.k[p] = k;
        rest -= k;
    }
    int L = log_n;
    for (int p = 0; p < P; ++p) {
        int k = pl.digits.k[p];
        pl.L[p] = L;
        pl.contiguous[p] = (L == k);
        int avail = pl.contiguous[p] ? (log_n - k) : (L - k);   // blocks resp. columns there are
        int want = log_n - 8 - k;                               // aim at >= 256 tiles
        int hi = NTT_TILE_LOG - k;
        int lc = want < 2 ? 2 : want;
        if (lc > hi) lc = hi;
        if (lc > avail) lc = avail;
        if (lc < 0) lc = 0;
        pl.logc[p] = lc;
        L -= k;
    }
    pl.lo_bits = (log_n + 1) / 2;
    return pl;
}

struct NttTables {                    // device (or host-emulation) pointers, all Montgomery
    const Fr* roots;                  // w_{2^11}^j, j < 1024   (forward or inverse set)
    const Fr* tlo; const Fr* thi;     // w_N^e two-level        (forward or inverse set)
    const Fr* glo; const Fr* ghi;     // coset powers two-level (g^e, or n^-1 * g^-e)
    const Fr* glo_raw;                // glo as plain (non-Montgomery) residues: a Montgomery product with it also leaves Montgomery form
    // per strided pass p: w_{2^L_p}^x, x < 2^(L_p - 1) (always the FORWARD roots; an inverse transform reads them negated
    // and mirrored), or null -> that pass rebuilds its inter-pass twiddles from tlo / thi (fallback when the table does not fit)
    const Fr* tw[NTT_MAX_PASSES];
    int inverse;                      // this set serves an inverse transform
};

// Passes of one transform in execution order.  rn = false: natural -> digit-reversed
// (passes 0..P-1, twiddle after); rn = true: digit-reversed -> natural (P-1..0, before).
inline std::vector<NttPass> ntt_build_passes(const NttPlan& pl, bool rn, const NttTables& tb) {
    std::vector<NttPass> out;
    const int P = pl.digits.n_pass;
    for (int i = 0; i < P; ++i) {
        int p = rn ? P - 1 - i : i;
        NttPass q{};
        q.log_n = pl.digits.log_n; q.L = pl.L[p]; q.k = pl.digits.k[p]; q.logc = pl.logc[p];
        q.contiguous = pl.contiguous[p];
        q.tw_mode = q.contiguous ? 0 : (rn ? 2 : 1);
        q.dit = 0; q.tw = nullptr; q.tw_half = 0; q.tw_inv = 0;
        if (!q.contiguous && tb.tw[p]) {                 // combined butterfly twiddles: no inter-pass twiddle product at all
            q.tw_mode = 3; q.tw = tb.tw[p]; q.tw_half = 1u << (q.L - 1); q.tw_inv = tb.inverse; q.dit = rn ? 1 : 0;
        }
        q.scale_mode = 0; q.scale_kind = 0;
        q.roots = tb.roots; q.tlo = tb.tlo; q.thi = tb.thi; q.lo_bits = pl.lo_bits;
        q.glo = tb.glo; q.ghi = tb.ghi; q.g_lo_bits = pl.lo_bits;
        q.cscale = Fr::one();
        q.digits = pl.digits;
        q.in_canon = q.check_canon = q.out_canon = q.perm_store = q.fuse_abc = 0;
        q.src_b = q.src_c = nullptr; q.zinv = Fr::one(); q.bad = nullptr; q.no_swz = 0;
        out.push_back(q);
    }
    return out;
}

// ---- transform chains as lists of (src, dst, pass); executed by ntt.cu on the device and by
// tests/host_emu on the CPU.
struct NttStep { const Fr* src; Fr* dst; NttPass pass; };
struct NttConsts { Fr n_inv; Fr zinv; };         // n^-1 ; (g^n - 1)^-1

// Natural order in and out (evaluation_domain fft / inverse_fft / coset variants).  src may
// equal dst; tmp (n elements) is needed when the plan has more than one pass.
inline std::vector<NttStep> ntt_steps_natural(const NttPlan& pl, const NttTables& tb, const NttConsts& k, const Fr* src, Fr* dst, Fr* tmp,
                                              bool inverse, bool coset, bool in_canon, bool out_canon, int* bad) {
    std::vector<NttPass> ps = ntt_build_passes(pl, false, tb);
    NttPass& first = ps.front();
    NttPass& last = ps.back();
    // The transform is linear and every constant it multiplies by is in Montgomery form, so canonical
    // input taken as it is (x = Montgomery form of x/R) comes out as canonical output: when both ends
    // are canonical neither conversion product is needed, only the range check of the input.
    const bool passthrough = in_canon && out_canon;
    first.in_canon = in_canon && !passthrough; first.check_canon = passthrough; first.bad = bad;
    if (coset && !inverse) { first.scale_mode = 1; first.scale_kind = 1; }       // x_i *= g^i
    if (inverse) {
        last.scale_mode = 2;
        if (coset) last.scale_kind = 2;                                          // n^-1 g^-idx
        else { last.scale_kind = 0; last.cscale = k.n_inv; }
    }
    last.out_canon = out_canon && !passthrough;
    last.perm_store = 1;
    std::vector<NttStep> out;
    const size_t P = ps.size();
    for (size_t i = 0; i < P; ++i) out.push_back({i == 0 ? src : tmp, i == P - 1 ? dst : tmp, ps[i]});
    return out;
}

// Witness-map tail: a, b, c evaluations on the domain (natural order;
// destroyed) -> H coefficients, canonical, natural order, in h.  Per vector: iNTT (natural ->
// digit-reversed coefficients), coset NTT back (digit-reversed -> natural, scaled by g^idx on
// load); then one iNTT whose first pass computes (a*b - c)/Z on load and whose last pass scales
// by n^-1 g^-idx, converts to canonical and stores in natural order.  7 transforms, no
// reorder passes.
inline std::vector<NttStep> ntt_steps_h(const NttPlan& pl, const NttTables& fwd_tb, const NttTables& inv_tb, const NttConsts& k,
                                        Fr* a, Fr* b, Fr* c, Fr* h, bool in_canon, int* bad) {
    std::vector<NttStep> out;
    Fr* v[3] = {a, b, c};
    for (int t = 0; t < 3; ++t) {
        std::vector<NttPass> inv = ntt_build_passes(pl, false, inv_tb);
        inv.front().in_canon = in_canon; inv.front().bad = bad;
        inv.back().scale_mode = 2; inv.back().scale_kind = 0; inv.back().cscale = k.n_inv;
        for (auto& p : inv) out.push_back({v[t], v[t], p});
        std::vector<NttPass> fwd = ntt_build_passes(pl, true, fwd_tb);
        fwd.front().scale_mode = 1; fwd.front().scale_kind = 2;                  // coefficient index at this position
        for (auto& p : fwd) out.push_back({v[t], v[t], p});
    }
    std::vector<NttPass> fin = ntt_build_passes(pl, false, inv_tb);
    fin.front().fuse_abc = 1; fin.front().src_b = b; fin.front().src_c = c; fin.front().zinv = k.zinv;
    fin.back().scale_mode = 2; fin.back().scale_kind = 2;
    fin.back().glo = inv_tb.glo_raw;                        // scaling by plain residues converts to canonical for free
    fin.back().out_canon = 0; fin.back().perm_store = 1;
    const size_t P = fin.size();
    for (size_t i = 0; i < P; ++i) out.push_back({a, i == P - 1 ? h : a, fin[i]});
    return out;
}

inline uint64_t ntt_pass_tiles(const NttPass& p) { return (1ull << p.log_n) >> (p.k + p.logc); }

}  // namespace b2
