// Host build of the *device* field/curve headers (fp.cuh, ec.cuh) with the PTX carry
// primitives emulated, so the exact limb algorithms are checked on the CPU against the
// oracle before running on a B200.  Test-only; built by tests/test_host_emulation.py.
#include <cstring>
#include <cstddef>
#include "fp.cuh"
#include "ec.cuh"
using namespace b2;

template <class F> static int field_op(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
    const int B = 4 * F::N;
    for (size_t i = 0; i < n; ++i) {
        uint32_t la[F::N], lb[F::N], lo[F::N];
        memcpy(la, a + B * i, B); memcpy(lb, b + B * i, B);
        if (F::geq_p(la) || F::geq_p(lb)) return -2;
        F x = F::from_canonical(la), y = F::from_canonical(lb), r;
        switch (op) { case 0: r = x + y; break; case 1: r = x - y; break; case 2: r = x * y; break;
            case 3: r = x.inv(); break; case 4: r = x.sqr(); break; case 6: r = x.inv_gcd(); break; default: r = x.neg(); }
        r.to_canonical(lo); memcpy(out + B * i, lo, B);
    }
    return 0;
}
static bool load_fq(Fq& f, const uint8_t* b) { uint32_t l[12]; memcpy(l, b, 48); if (Fq::geq_p(l)) return false; f = Fq::from_canonical(l); return true; }
static void store_fq(const Fq& f, uint8_t* b) { uint32_t l[12]; f.to_canonical(l); memcpy(b, l, 48); }
static bool load(G1Affine& p, const uint8_t* b) { return load_fq(p.x, b) && load_fq(p.y, b + 48); }
static bool load(G2Affine& p, const uint8_t* b) { return load_fq(p.x.c0, b) && load_fq(p.x.c1, b + 48) && load_fq(p.y.c0, b + 96) && load_fq(p.y.c1, b + 144); }
static void store(const G1Affine& p, uint8_t* b) { store_fq(p.x, b); store_fq(p.y, b + 48); }
static void store(const G2Affine& p, uint8_t* b) { store_fq(p.x.c0, b); store_fq(p.x.c1, b + 48); store_fq(p.y.c0, b + 96); store_fq(p.y.c1, b + 144); }

// sum of n affine points: mode 0 = chain of add_mixed; 1 = pairwise tree of full adds;
// 2 = add_mixed chain then 3 doublings (x8)
template <class F, int SZ> static int sum_points(int mode, const uint8_t* pts, size_t n, uint8_t* out) {
    Xyzz<F> acc = Xyzz<F>::identity();
    if (mode == 1) {
        Xyzz<F> a = Xyzz<F>::identity(), b = Xyzz<F>::identity();
        for (size_t i = 0; i < n; ++i) { Affine<F> p; if (!load(p, pts + SZ * i)) return -2; ((i & 1) ? b : a).add_mixed(p); }
        a.add(b); acc = a;
    } else {
        for (size_t i = 0; i < n; ++i) { Affine<F> p; if (!load(p, pts + SZ * i)) return -2; acc.add_mixed(p); }
        if (mode == 2) for (int k = 0; k < 3; ++k) acc = acc.dbl();
    }
    store(acc.to_affine(), out);
    return 0;
}
extern "C" {
int emu_fr_op(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) { return field_op<Fr>(op, a, b, n, out); }
int emu_fq_op(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) { return field_op<Fq>(op, a, b, n, out); }
int emu_g1_sum(int mode, const uint8_t* p, size_t n, uint8_t* out) { return sum_points<Fq, 96>(mode, p, n, out); }
int emu_g2_sum(int mode, const uint8_t* p, size_t n, uint8_t* out) { return sum_points<Fq2, 192>(mode, p, n, out); }
// fp_wide.cuh: which = 0 split product + reduction, 1 Karatsuba product + reduction (Fq, 48-byte canonical operands);
// which = 2 / 3: Fq2 product with lazy reduction, schoolbook / Karatsuba (96-byte operands c0 || c1).  -1 on a non-canonical operand.
int emu_wide_mul(int which, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
    const int comps = (which == 2 || which == 3) ? 2 : 1;
    for (size_t i = 0; i < n; ++i) {
        Fq x[2], y[2];
        for (int k = 0; k < comps; ++k) {
            uint32_t ca[12], cb[12];
            memcpy(ca, a + (i * comps + k) * 48, 48); memcpy(cb, b + (i * comps + k) * 48, 48);
            if (Fq::geq_p(ca) || Fq::geq_p(cb)) return -1;
            x[k] = Fq::from_canonical(ca); y[k] = Fq::from_canonical(cb);
        }
        Fq r[2];
        if (which == 4) r[0] = wide::fq_sqr_halves(x[0]);                     // b ignored
        else if (which == 0) r[0] = wide::fq_mul<false>(x[0], y[0]);
        else if (which == 1) r[0] = wide::fq_mul<true>(x[0], y[0]);
        else {
            const Fq2 z = which == 2 ? wide::fq2_mul_lazy<false>(Fq2{x[0], x[1]}, Fq2{y[0], y[1]}) : wide::fq2_mul_lazy<true>(Fq2{x[0], x[1]}, Fq2{y[0], y[1]});
            r[0] = z.c0; r[1] = z.c1;
        }
        for (int k = 0; k < comps; ++k) { uint32_t c[12]; r[k].to_canonical(c); memcpy(out + (i * comps + k) * 48, c, 48); }
    }
    return 0;
}
}

// ---------------------------------------------------------------------------------------
// NTT: the device pass kernel body (ntt_kernels.cuh phases) and the chain builders
// (ntt_plan.cuh) executed tile by tile on the CPU, tables built with the same formulas as
// ntt.cu.  Checks plan / tile / twiddle index arithmetic against the oracle without a GPU.
#include <vector>
#include "ntt_plan.cuh"
#include "msm_digits.cuh"

static Fr fr_u64(uint64_t x) { uint32_t c[8] = {(uint32_t)x, (uint32_t)(x >> 32), 0, 0, 0, 0, 0, 0}; return Fr::from_canonical(c); }
static Fr fr_pow(Fr b, uint64_t e) { Fr r = Fr::one(); while (e) { if (e & 1) r = r * b; b = b.sqr(); e >>= 1; } return r; }
static Fr root_of_unity(int log_n) {
    uint32_t e[8];
    for (int i = 0; i < 8; ++i) e[i] = FrParams::P(i);
    e[0] -= 1;
    for (int s = 0; s < log_n; ++s) for (int i = 0; i < 8; ++i) e[i] = (e[i] >> 1) | (i < 7 ? e[i + 1] << 31 : 0);
    return fr_u64(7).pow_limbs(e, 8);
}
static std::vector<Fr> pow_table(Fr base, Fr pre, int shift, uint32_t count) {
    std::vector<Fr> t(count);
    for (uint32_t e = 0; e < count; ++e) t[e] = pre * fr_pow(base, (uint64_t)e << shift);
    return t;
}
struct EmuDomain {
    NttPlan plan; NttConsts k;
    std::vector<Fr> roots_f, roots_i, tlo_f, thi_f, tlo_i, thi_i, glo_f, ghi_f, glo_i, ghi_i, glo_f_raw, glo_i_raw;
    std::vector<Fr> tw[NTT_MAX_PASSES];
    explicit EmuDomain(int log_n, bool tw_tables = true) {
        plan = ntt_make_plan(log_n);
        int lo = plan.lo_bits, hi = log_n - lo;
        uint32_t n_lo = 1u << lo, n_hi = 1u << (hi > 0 ? hi : 0);
        Fr w = root_of_unity(log_n), wi = w.inv(), g = fr_u64(7), gi = g.inv();
        k.n_inv = fr_u64(1ull << log_n).inv();
        k.zinv = (fr_pow(g, 1ull << log_n) - Fr::one()).inv();
        Fr w11 = root_of_unity(NTT_ROOT_LOG);
        roots_f = pow_table(w11, Fr::one(), 0, 1u << (NTT_ROOT_LOG - 1));
        roots_i = pow_table(w11.inv(), Fr::one(), 0, 1u << (NTT_ROOT_LOG - 1));
        tlo_f = pow_table(w, Fr::one(), 0, n_lo); thi_f = pow_table(w, Fr::one(), lo, n_hi);
        tlo_i = pow_table(wi, Fr::one(), 0, n_lo); thi_i = pow_table(wi, Fr::one(), lo, n_hi);
        glo_f = pow_table(g, Fr::one(), 0, n_lo); ghi_f = pow_table(g, Fr::one(), lo, n_hi);
        glo_i = pow_table(gi, k.n_inv, 0, n_lo); ghi_i = pow_table(gi, Fr::one(), lo, n_hi);
        Fr raw_one = Fr::zero(); raw_one.v[0] = 1;
        glo_f_raw = pow_table(g, raw_one, 0, n_lo); glo_i_raw = pow_table(gi, k.n_inv * raw_one, 0, n_lo);
        if (tw_tables) for (int p = 0; p < plan.digits.n_pass; ++p) if (!plan.contiguous[p]) tw[p] = pow_table(root_of_unity(plan.L[p]), Fr::one(), 0, 1u << (plan.L[p] - 1));
    }
    NttTables tables(bool inverse) const {
        NttTables t;
        t.roots = (inverse ? roots_i : roots_f).data();
        t.tlo = (inverse ? tlo_i : tlo_f).data(); t.thi = (inverse ? thi_i : thi_f).data();
        t.glo = (inverse ? glo_i : glo_f).data(); t.ghi = (inverse ? ghi_i : ghi_f).data();
        t.glo_raw = (inverse ? glo_i_raw : glo_f_raw).data();
        for (int p = 0; p < NTT_MAX_PASSES; ++p) t.tw[p] = tw[p].empty() ? nullptr : tw[p].data();
        t.inverse = inverse ? 1 : 0;
        return t;
    }
};
static bool g_emu_ntt_rounds = true;    // register rounds (what the kernel runs) or one stage at a time (the reference form)
static void run_step(const NttStep& s) {
    const NttPass& p = s.pass;
    const int T = 1 << (p.k + p.logc);
    std::vector<uint4> lo(T), hi(T);
    for (uint64_t tile = 0; tile < ntt_pass_tiles(p); ++tile) {
        NttTile t = ntt_make_tile(p, tile);
        for (int e = 0; e < T; ++e) ntt_phase_load(p, t, s.src, lo.data(), hi.data(), e);
        if (g_emu_ntt_rounds) {
            for (int r = 0; r < ntt_round_count(p); ++r) for (int sl = 0; sl < (T >> ntt_round_size(p, r)); ++sl) ntt_phase_round_any(p, t, lo.data(), hi.data(), r, sl);
        } else {
            for (int i = 0; i < p.k; ++i) for (int b = 0; b < T / 2; ++b) ntt_phase_stage(p, t, lo.data(), hi.data(), ntt_stage_hs(p, i), b);
        }
        for (int e = 0; e < T; ++e) ntt_phase_store(p, t, s.dst, lo.data(), hi.data(), e);
    }
}
extern "C" {
static bool g_emu_tw_tables = true;
void emu_set_tw_tables(int on) { g_emu_tw_tables = on != 0; }
void emu_set_ntt_rounds(int on) { g_emu_ntt_rounds = on != 0; }
int emu_ntt(uint8_t* data, unsigned log_n, int inverse, int coset) {
    EmuDomain d((int)log_n, g_emu_tw_tables);
    size_t n = (size_t)1 << log_n;
    std::vector<Fr> buf(n), tmp(n);
    memcpy(buf.data(), data, 32 * n);
    for (auto& s : ntt_steps_natural(d.plan, d.tables(inverse != 0), d.k, buf.data(), buf.data(), tmp.data(), inverse != 0, coset != 0, true, true, nullptr)) run_step(s);
    memcpy(data, buf.data(), 32 * n);
    return d.plan.digits.n_pass;
}
int emu_h_from_abc(const uint8_t* a, const uint8_t* b, const uint8_t* c, unsigned log_m, uint8_t* h) {
    EmuDomain d((int)log_m, g_emu_tw_tables);
    size_t m = (size_t)1 << log_m;
    std::vector<Fr> va(m), vb(m), vc(m), vh(m);
    memcpy(va.data(), a, 32 * m); memcpy(vb.data(), b, 32 * m); memcpy(vc.data(), c, 32 * m);
    for (auto& s : ntt_steps_h(d.plan, d.tables(false), d.tables(true), d.k, va.data(), vb.data(), vc.data(), vh.data(), true, nullptr)) run_step(s);
    memcpy(h, vh.data(), 32 * m);
    return 0;
}
// signed digits of one canonical scalar: out[w], w < msm_windows(c)
int emu_msm_digits(const uint8_t* scalar, int c, int* out) {
    uint32_t s[8]; memcpy(s, scalar, 32);
    int W = msm_windows(c);
    for (int w = 0; w < W; ++w) out[w] = msm_signed_digit(s, w, c);
    return W;
}
// Tail of the bucket reduction by bit decomposition (msm_reduce_bits_kernel + msm_reduce_bits_final_kernel), sequentially:
// V = sum_t a_t + 2^log_s sum_t t r_t = sum_t a_t + 2^log_s sum_b 2^b S_b,  S_b = sum of the r_t whose index has bit b set,
// enumerated with the kernel's own msm_bits_index.  a, r: count affine G1 points each (ABI bytes); out: affine.
int emu_reduce_bits_g1(const uint8_t* a, const uint8_t* r, uint32_t count, int log_s, uint8_t* out) {
    int NB = 0;
    while ((1u << NB) < count) ++NB;
    Xyzz<Fq> total = Xyzz<Fq>::identity();
    for (int b = NB - 1; b >= 0; --b) {                       // Horner over the bits: total = 2 total + S_b
        total = total.dbl();
        Xyzz<Fq> S = Xyzz<Fq>::identity();
        for (uint32_t j = 0;; ++j) {
            const uint32_t t = msm_bits_index(j, b);
            if (t >= count) break;
            Affine<Fq> p; if (!load(p, r + 96 * (size_t)t)) return -2;
            S.add_mixed(p);
        }
        total.add(S);
    }
    for (int k = 0; k < log_s; ++k) total = total.dbl();
    for (uint32_t t = 0; t < count; ++t) { Affine<Fq> p; if (!load(p, a + 96 * (size_t)t)) return -2; total.add_mixed(p); }
    store(total.to_affine(), out);
    return NB;
}
// The MSM pipeline of msm_impl.cuh restated sequentially with the same bucket layout,
// segment split, chunked running sum (+ chunk-offset lift) and Horner tail.
int emu_msm_g1(const uint8_t* bases, const uint8_t* scalars, size_t n, int c, uint32_t S, uint8_t* out) {
    const int W = msm_windows(c);
    const uint32_t B = 1u << (c - 1), T = msm_reduce_chunks(c), per = B / T;
    std::vector<G1Affine> pts(n);
    for (size_t i = 0; i < n; ++i) if (!load(pts[i], bases + 96 * i)) return -2;
    std::vector<std::vector<uint32_t>> bucket((size_t)W * B);
    for (size_t i = 0; i < n; ++i) {
        uint32_t s[8]; memcpy(s, scalars + 32 * i, 32);
        for (int w = 0; w < W; ++w) {
            int d = msm_signed_digit(s, w, c);
            if (d) bucket[(size_t)w * B + (uint32_t)((d < 0 ? -d : d) - 1)].push_back((uint32_t)i | (d < 0 ? 0x80000000u : 0u));
        }
    }
    std::vector<G1Xyzz> win(W);
    for (int w = 0; w < W; ++w) {
        G1Xyzz wsum = G1Xyzz::identity();
        for (uint32_t t = 0; t < T; ++t) {
            uint32_t lo = t * per;
            G1Xyzz run = G1Xyzz::identity(), acc = G1Xyzz::identity();
            for (uint32_t j = lo + per; j-- > lo;) {
                const std::vector<uint32_t>& e = bucket[(size_t)w * B + j];
                for (size_t s0 = 0; s0 < e.size(); s0 += S) {                 // segments
                    G1Xyzz seg = G1Xyzz::identity();
                    for (size_t q = s0; q < e.size() && q < s0 + S; ++q) {
                        G1Affine p = pts[e[q] & 0x7fffffffu];
                        if (e[q] >> 31) p.y = p.y.neg();
                        seg.add_mixed(p);
                    }
                    run.add(seg);
                }
                acc.add(run);
            }
            G1Xyzz m = G1Xyzz::identity();
            for (int bit = 31; bit >= 0; --bit) { m = m.dbl(); if ((lo >> bit) & 1u) m.add(run); }
            acc.add(m);
            wsum.add(acc);
        }
        win[w] = wsum;
    }
    G1Xyzz acc = win[W - 1];
    for (int w = W - 2; w >= 0; --w) { for (int k = 0; k < c; ++k) acc = acc.dbl(); acc.add(win[w]); }
    store(acc.to_affine(), out);
    return 0;
}
}

// ---------------------------------------------------------------------------------------
// Batched-affine tree accumulation (msm_tree.cuh): the kernel bodies run thread by thread on the
// CPU over the same sorted bucket layout the device builds, `levels` tree levels, then each
// bucket's remaining entries are summed and the windows combined.  Checks the output <-> input
// index arithmetic, the prefix-product / back-substitution chain across the depths of the product
// tree and the special cases (copy, identity operand, doubling, cancellation) against the oracle.
#include "msm_tree.cuh"
template <class F, int SZ> static int emu_msm_tree(const uint8_t* bases, const uint8_t* scalars, size_t n, int c, uint32_t K, uint32_t K2, int levels, uint8_t* out) {
    const int W = msm_windows(c);
    const uint32_t B = 1u << (c - 1), nb = (uint32_t)W * B;
    std::vector<Affine<F>> pts(n);
    for (size_t i = 0; i < n; ++i) if (!load(pts[i], bases + SZ * i)) return -2;
    std::vector<std::vector<uint32_t>> bucket(nb);
    for (size_t i = 0; i < n; ++i) {
        uint32_t s[8]; memcpy(s, scalars + 32 * i, 32);
        if (pts[i].is_identity()) continue;
        for (int w = 0; w < W; ++w) {
            int d = msm_signed_digit(s, w, c);
            if (d) bucket[(size_t)w * B + (uint32_t)((d < 0 ? -d : d) - 1)].push_back((uint32_t)i | (d < 0 ? 0x80000000u : 0u));
        }
    }
    std::vector<uint32_t> sorted, off(nb + 1, 0);
    for (uint32_t b = 0; b < nb; ++b) { off[b] = (uint32_t)sorted.size(); sorted.insert(sorted.end(), bucket[b].begin(), bucket[b].end()); }
    off[nb] = (uint32_t)sorted.size();
    TreeIn<F> in{pts.data(), sorted.data(), nullptr, off.data()};
    std::vector<Affine<F>> cur, nxt;
    std::vector<uint32_t> off_in = off, off_out(nb + 1);
    for (int l = 0; l < levels; ++l) {
        for (uint32_t b = 0, run = 0; b <= nb; ++b) { off_out[b] = run; if (b < nb) run += (off_in[b + 1] - off_in[b] + 1) / 2; }
        const uint32_t e_out = off_out[nb];
        nxt.assign(e_out, Affine<F>::identity());
        // product tree: depth 0 = the denominators (implicit), depth d >= 1 = totals
        std::vector<std::vector<F>> val, pp;
        std::vector<F> pp0(e_out);
        val.push_back(std::vector<F>(tree_count_at(e_out, 1, K, K2)));
        in.off = off_in.data();
        for (uint32_t t = 0; t < val[0].size() + 2; ++t) tree_up0_body<F>(in, off_out.data(), nb, K, t, pp0.data(), val[0].data());   // +2: threads past the end must do nothing
        int depth = 1;
        while (val.back().size() > TREE_TOP) {
            const uint32_t cnt = (uint32_t)val.back().size();
            if (cnt != tree_count_at(e_out, depth, K, K2)) return -5;
            std::vector<F> tot((cnt + K2 - 1) / K2), p(cnt);
            for (uint32_t t = 0; t < tot.size() + 2; ++t) tree_up_body<F>(val.back().data(), cnt, K2, t, p.data(), tot.data());
            pp.push_back(p); val.push_back(tot); ++depth;
        }
        { std::vector<F> scratch(val.back().size() + 1); tree_top_body<F>(val.back().data(), (uint32_t)val.back().size(), scratch.data()); }
        for (int d = (int)val.size() - 2; d >= 0; --d) {
            const uint32_t cnt = (uint32_t)val[d].size();
            for (uint32_t t = 0; t < val[d + 1].size() + 2; ++t) tree_down_body<F>(val[d].data(), cnt, K2, t, pp[d].data(), val[d + 1].data());
        }
        for (uint32_t t = 0; t < val[0].size() + 2; ++t) tree_down0_body<F>(in, off_out.data(), nb, K, t, pp0.data(), val[0].data(), nxt.data());
        cur.swap(nxt);
        off_in = off_out;
        in = TreeIn<F>{nullptr, nullptr, cur.data(), off_in.data()};
    }
    in.off = off_in.data();
    std::vector<Xyzz<F>> win(W);
    for (int w = 0; w < W; ++w) {
        Xyzz<F> run = Xyzz<F>::identity(), acc = Xyzz<F>::identity();
        for (uint32_t j = B; j-- > 0;) {
            const uint32_t b = (uint32_t)w * B + j;
            Xyzz<F> bk = Xyzz<F>::identity();
            for (uint32_t e = off_in[b]; e < off_in[b + 1]; ++e) bk.add_mixed(tree_entry(in, e));
            run.add(bk);
            acc.add(run);
        }
        win[w] = acc;
    }
    Xyzz<F> acc = win[W - 1];
    for (int w = W - 2; w >= 0; --w) { for (int k = 0; k < c; ++k) acc = acc.dbl(); acc.add(win[w]); }
    store(acc.to_affine(), out);
    return 0;
}
extern "C" {
int emu_msm_tree_g1(const uint8_t* bases, const uint8_t* scalars, size_t n, int c, uint32_t K, uint32_t K2, int levels, uint8_t* out) {
    return emu_msm_tree<Fq, 96>(bases, scalars, n, c, K, K2, levels, out);
}
int emu_msm_tree_g2(const uint8_t* bases, const uint8_t* scalars, size_t n, int c, uint32_t K, uint32_t K2, int levels, uint8_t* out) {
    return emu_msm_tree<Fq2, 192>(bases, scalars, n, c, K, K2, levels, out);
}
}
