// CPU restatement of the Groth16 proving path (fields, curves, MSM, NTT, prove).
//
// TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library.  The product
// (crypto3-fil-proofs_b200/) never links or calls it.
//
// Parity status: "parity unpinned" at byte level.  The reference tree
// (/root/reference) contains no implementation of this path: the arithmetic lives
// in un-vendored submodules  nilfoundation/zk @ 44a19899 (r1cs_gg_ppzksnark prover,
// r1cs_to_qap witness map), crypto3-algebra @ b2e1a199 (BLS12-381, multiexp),
// crypto3-math @ 16c55f5c (evaluation_domain), crypto3-multiprecision @ 5dbfd51a
// (.SUBMODULES.json:65-148, all "missing").  This file restates their published
// algorithms (Montgomery arithmetic, short-Weierstrass group law, bucket-method
// multiexp, radix-2 Cooley-Tukey FFT, Groth16 prover) and is pinned against
//   (1) oracle/pyref (big-int Python, affine group law, O(n^2) DFT, pairing check),
//   (2) the reference's own KATs: Fr canonical encodings
//       libs/storage/test/core/fr32.cpp:51-85, libs/filecoin/test/api/mod.cpp:35-43,
//       proof length 192 B  libs/filecoin/include/nil/filecoin/proofs/constants.hpp:93,
//       prove->verify round trips  libs/storage/test/**/compound.cpp.
// Reference call sites this stands in for:
//   libs/storage/include/nil/filecoin/storage/proofs/core/proof/compound_proof.hpp:127-137
//   libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:46-66
//
// Build: g++ -O3 -march=native -fopenmp -shared -fPIC oracle.cpp -o ../_build/liboracle.so
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint64_t u64;
typedef unsigned __int128 u128;

// ------------------------------------------------------------------ prime fields
template <int N> struct Modulus {
    u64 p[N];
    u64 inv;        // -p^-1 mod 2^64
    u64 one[N];     // R mod p
    u64 r2[N];      // R^2 mod p
};

template <int N> static bool geq(const u64* a, const u64* b) {
    for (int i = N - 1; i >= 0; --i) {
        if (a[i] != b[i]) return a[i] > b[i];
    }
    return true;
}
template <int N> static u64 sub_n(u64* r, const u64* a, const u64* b) {
    u64 borrow = 0;
    for (int i = 0; i < N; ++i) {
        u128 t = (u128)a[i] - b[i] - borrow;
        r[i] = (u64)t;
        borrow = (u64)(t >> 64) & 1;
    }
    return borrow;
}
template <int N> static u64 add_n(u64* r, const u64* a, const u64* b) {
    u64 carry = 0;
    for (int i = 0; i < N; ++i) {
        u128 t = (u128)a[i] + b[i] + carry;
        r[i] = (u64)t;
        carry = (u64)(t >> 64);
    }
    return carry;
}

template <int N> static void init_modulus(Modulus<N>& m, const u64* p) {
    memcpy(m.p, p, sizeof(m.p));
    u64 x = 1;                       // Newton: x = p^-1 mod 2^64
    for (int i = 0; i < 7; ++i) x *= 2 - p[0] * x;
    m.inv = (u64)0 - x;
    // R mod p and R^2 mod p by repeated doubling of 1
    u64 v[N] = {1};
    for (int i = 0; i < 2 * 64 * N; ++i) {
        u64 c = add_n<N>(v, v, v);
        if (c || geq<N>(v, p)) sub_n<N>(v, v, p);
        if (i == 64 * N - 1) memcpy(m.one, v, sizeof(v));
    }
    memcpy(m.r2, v, sizeof(v));
}

template <int N, const Modulus<N>& M()> struct Fp {
    u64 v[N];
    static Fp zero() { Fp r; memset(r.v, 0, sizeof(r.v)); return r; }
    static Fp one() { Fp r; memcpy(r.v, M().one, sizeof(r.v)); return r; }
    bool is_zero() const { u64 t = 0; for (int i = 0; i < N; ++i) t |= v[i]; return t == 0; }
    bool operator==(const Fp& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
    bool operator!=(const Fp& o) const { return !(*this == o); }
    Fp operator+(const Fp& o) const {
        Fp r; u64 c = add_n<N>(r.v, v, o.v);
        if (c || geq<N>(r.v, M().p)) sub_n<N>(r.v, r.v, M().p);
        return r;
    }
    Fp operator-(const Fp& o) const {
        Fp r; u64 b = sub_n<N>(r.v, v, o.v);
        if (b) add_n<N>(r.v, r.v, M().p);
        return r;
    }
    Fp neg() const { return is_zero() ? *this : (zero() - *this); }
    Fp dbl() const { return *this + *this; }
    Fp operator*(const Fp& o) const {           // CIOS Montgomery product
        u64 t[N + 2] = {0};
        const u64* p = M().p;
        for (int i = 0; i < N; ++i) {
            u64 c = 0;
            for (int j = 0; j < N; ++j) {
                u128 s = (u128)v[j] * o.v[i] + t[j] + c;
                t[j] = (u64)s; c = (u64)(s >> 64);
            }
            u128 s = (u128)t[N] + c; t[N] = (u64)s; t[N + 1] = (u64)(s >> 64);
            u64 m = t[0] * M().inv;
            s = (u128)m * p[0] + t[0]; c = (u64)(s >> 64);
            for (int j = 1; j < N; ++j) {
                s = (u128)m * p[j] + t[j] + c;
                t[j - 1] = (u64)s; c = (u64)(s >> 64);
            }
            s = (u128)t[N] + c; t[N - 1] = (u64)s; t[N] = t[N + 1] + (u64)(s >> 64);
        }
        Fp r;
        if (t[N] || geq<N>(t, p)) sub_n<N>(r.v, t, p); else memcpy(r.v, t, sizeof(r.v));
        return r;
    }
    Fp sqr() const { return *this * *this; }
    Fp pow(const u64* e, int n) const {
        Fp r = one();
        for (int i = n * 64 - 1; i >= 0; --i) {
            r = r.sqr();
            if ((e[i / 64] >> (i % 64)) & 1) r = r * *this;
        }
        return r;
    }
    Fp inv() const {                             // Fermat: a^(p-2)
        u64 e[N]; u64 two[N] = {2};
        sub_n<N>(e, M().p, two);
        return pow(e, N);
    }
    // canonical little-endian bytes <-> Montgomery form
    static bool from_bytes(Fp& out, const uint8_t* b) {
        Fp t; memcpy(t.v, b, sizeof(t.v));
        if (geq<N>(t.v, M().p)) return false;
        Fp r2; memcpy(r2.v, M().r2, sizeof(r2.v));
        out = t * r2;
        return true;
    }
    void to_bytes(uint8_t* b) const {
        Fp o; memset(o.v, 0, sizeof(o.v)); o.v[0] = 1;
        Fp c = *this * o;                        // divide by R
        memcpy(b, c.v, sizeof(c.v));
    }
    static Fp from_u64(u64 x) { uint8_t b[8 * N] = {0}; memcpy(b, &x, 8); Fp r; from_bytes(r, b); return r; }
};

static Modulus<4> g_fr; static Modulus<6> g_fq;
static const Modulus<4>& MR() { return g_fr; }
static const Modulus<6>& MQ() { return g_fq; }
typedef Fp<4, MR> Fr;
typedef Fp<6, MQ> Fq;

struct Fq2 {
    Fq c0, c1;
    static Fq2 zero() { return {Fq::zero(), Fq::zero()}; }
    static Fq2 one() { return {Fq::one(), Fq::zero()}; }
    bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    bool operator==(const Fq2& o) const { return c0 == o.c0 && c1 == o.c1; }
    bool operator!=(const Fq2& o) const { return !(*this == o); }
    Fq2 operator+(const Fq2& o) const { return {c0 + o.c0, c1 + o.c1}; }
    Fq2 operator-(const Fq2& o) const { return {c0 - o.c0, c1 - o.c1}; }
    Fq2 neg() const { return {c0.neg(), c1.neg()}; }
    Fq2 dbl() const { return {c0.dbl(), c1.dbl()}; }
    Fq2 operator*(const Fq2& o) const {          // (a+bu)(c+du), u^2=-1
        Fq ac = c0 * o.c0, bd = c1 * o.c1;
        Fq t = (c0 + c1) * (o.c0 + o.c1);
        return {ac - bd, t - ac - bd};
    }
    Fq2 sqr() const { return *this * *this; }
    Fq2 inv() const {
        Fq n = (c0.sqr() + c1.sqr()).inv();
        return {c0 * n, (c1 * n).neg()};
    }
    static bool from_bytes(Fq2& o, const uint8_t* b) { return Fq::from_bytes(o.c0, b) && Fq::from_bytes(o.c1, b + 48); }
    void to_bytes(uint8_t* b) const { c0.to_bytes(b); c1.to_bytes(b + 48); }
};

// ---------------------------------------------------------------- curve points
// y^2 = x^3 + b (a = 0).  XYZZ coordinates: x = X/ZZ, y = Y/ZZZ; identity: ZZ = 0.
template <class F> struct Affine {
    F x, y; bool inf;
};
template <class F> struct Xyzz {
    F X, Y, ZZ, ZZZ;
    static Xyzz identity() { return {F::zero(), F::zero(), F::zero(), F::zero()}; }
    bool is_identity() const { return ZZ.is_zero(); }
    static Xyzz from_affine(const Affine<F>& a) {
        if (a.inf) return identity();
        return {a.x, a.y, F::one(), F::one()};
    }
    Xyzz dbl() const {
        if (is_identity()) return *this;
        F U = Y.dbl(), V = U.sqr(), W = U * V, S = X * V;
        F X2 = X.sqr(), M = X2.dbl() + X2;
        F X3 = M.sqr() - S.dbl();
        F Y3 = M * (S - X3) - W * Y;
        return {X3, Y3, V * ZZ, W * ZZZ};
    }
    Xyzz add(const Xyzz& o) const {
        if (is_identity()) return o;
        if (o.is_identity()) return *this;
        F U1 = X * o.ZZ, U2 = o.X * ZZ, S1 = Y * o.ZZZ, S2 = o.Y * ZZZ;
        F Pp = U2 - U1, Rr = S2 - S1;
        if (Pp.is_zero()) {
            if (Rr.is_zero()) return dbl();
            return identity();
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = U1 * PP;
        F X3 = Rr.sqr() - PPP - Q.dbl();
        F Y3 = Rr * (Q - X3) - S1 * PPP;
        return {X3, Y3, ZZ * o.ZZ * PP, ZZZ * o.ZZZ * PPP};
    }
    Xyzz add_mixed(const Affine<F>& a) const {
        if (a.inf) return *this;
        if (is_identity()) return from_affine(a);
        F U2 = a.x * ZZ, S2 = a.y * ZZZ;
        F Pp = U2 - X, Rr = S2 - Y;
        if (Pp.is_zero()) {
            if (Rr.is_zero()) return from_affine(a).dbl();
            return identity();
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = X * PP;
        F X3 = Rr.sqr() - PPP - Q.dbl();
        F Y3 = Rr * (Q - X3) - Y * PPP;
        return {X3, Y3, ZZ * PP, ZZZ * PPP};
    }
    Xyzz neg() const { return {X, Y.neg(), ZZ, ZZZ}; }
    Affine<F> to_affine() const {
        if (is_identity()) return {F::zero(), F::zero(), true};
        F I = (ZZ * ZZZ).inv();
        return {X * I * ZZZ, Y * I * ZZ, false};
    }
    Xyzz mul(const Fr& k) const {                // double-and-add on canonical bits
        uint8_t b[32]; k.to_bytes(b);
        Xyzz acc = identity();
        for (int i = 254; i >= 0; --i) {
            acc = acc.dbl();
            if ((b[i / 8] >> (i % 8)) & 1) acc = acc.add(*this);
        }
        return acc;
    }
};
typedef Affine<Fq> G1A; typedef Xyzz<Fq> G1;
typedef Affine<Fq2> G2A; typedef Xyzz<Fq2> G2;

template <class F> static int fsize();
template <> int fsize<Fq>() { return 48; }
template <> int fsize<Fq2>() { return 96; }

// ABI encodings: x||y canonical LE; identity = all-zero.
template <class F> static bool affine_from_bytes(Affine<F>& a, const uint8_t* b) {
    int n = fsize<F>();
    bool allz = true;
    for (int i = 0; i < 2 * n; ++i) if (b[i]) { allz = false; break; }
    if (allz) { a = {F::zero(), F::zero(), true}; return true; }
    a.inf = false;
    return F::from_bytes(a.x, b) && F::from_bytes(a.y, b + n);
}
template <class F> static void affine_to_bytes(const Affine<F>& a, uint8_t* b) {
    int n = fsize<F>();
    if (a.inf) { memset(b, 0, 2 * n); return; }
    a.x.to_bytes(b); a.y.to_bytes(b + n);
}

// batch normalisation (Montgomery's trick)
template <class F> static void batch_to_affine(const std::vector<Xyzz<F>>& in, std::vector<Affine<F>>& out) {
    size_t n = in.size();
    out.resize(n);
    std::vector<F> pre(n);
    F acc = F::one();
    for (size_t i = 0; i < n; ++i) {
        pre[i] = acc;
        if (!in[i].is_identity()) acc = acc * (in[i].ZZ * in[i].ZZZ);
    }
    F inv = acc.inv();
    for (size_t i = n; i-- > 0;) {
        if (in[i].is_identity()) { out[i] = {F::zero(), F::zero(), true}; continue; }
        F I = inv * pre[i];
        inv = inv * (in[i].ZZ * in[i].ZZZ);
        out[i] = {in[i].X * I * in[i].ZZZ, in[i].Y * I * in[i].ZZ, false};
    }
}

// --------------------------------------------------------------- init / consts
static bool g_init = false;
static G1A g_g1; static G2A g_g2;
static void hex_to_limbs(const char* hex, u64* out, int n) {
    memset(out, 0, 8 * n);
    int len = (int)strlen(hex);
    for (int i = 0; i < len; ++i) {
        char ch = hex[len - 1 - i];
        u64 d = (ch >= '0' && ch <= '9') ? ch - '0' : (ch | 32) - 'a' + 10;
        out[i / 16] |= d << (4 * (i % 16));
    }
}
static Fq fq_hex(const char* h) { u64 l[6]; hex_to_limbs(h, l, 6); Fq r; Fq::from_bytes(r, (const uint8_t*)l); return r; }

static void ensure_init() {
    if (g_init) return;
#pragma omp critical(orc_init)
    {
        if (!g_init) {
            u64 r[4], p[6];
            hex_to_limbs("73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001", r, 4);
            hex_to_limbs("1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab", p, 6);
            init_modulus<4>(g_fr, r);
            init_modulus<6>(g_fq, p);
            g_g1 = {fq_hex("17f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb"),
                    fq_hex("08b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1"), false};
            g_g2 = {{fq_hex("024aa2b2f08f0a91260805272dc51051c6e47ad4fa403b02b4510b647ae3d1770bac0326a805bbefd48056c8c121bdb8"),
                     fq_hex("13e02b6052719f607dacd3a088274f65596bd0d09920b61ab5da61bbdc7f5049334cf11213945d57e5ac7d055d042b7e")},
                    {fq_hex("0ce5d527727d6e118cc9cdc6da2e351aadfd9baa8cbdd3a76d429a695160d12c923ac9cc3baca289e193548608b82801"),
                     fq_hex("0606c4a02ea734cc32acd2b02bc28b99cb3e287e85a763af267492ab572e99ab3f370d275cec1da1aaa9075ff05f79be")}, false};
            g_init = true;
        }
    }
}

static Fr fr_pow_u64(Fr b, u64 e) { return b.pow(&e, 1); }
static Fr fr_root_of_unity(unsigned log_n) {     // 7^((r-1)/2^log_n)
    u64 e[4]; u64 one[4] = {1};
    sub_n<4>(e, g_fr.p, one);
    for (unsigned s = 0; s < log_n; ++s) {        // e >>= 1
        for (int i = 0; i < 4; ++i) e[i] = (e[i] >> 1) | (i < 3 ? e[i + 1] << 63 : 0);
    }
    return Fr::from_u64(7).pow(e, 4);
}

// -------------------------------------------------------------------------- NTT
// Natural order in/out, iterative radix-2 (bit-reverse + DIT), OpenMP over
// butterflies.  X[k] = sum x[n] w^(nk); inverse scales by 1/n.
static void bitrev_permute(std::vector<Fr>& a, unsigned log_n) {
    size_t n = a.size();
    if (log_n == 0) return;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        size_t j = 0, t = i;
        for (unsigned b = 0; b < log_n; ++b) { j = (j << 1) | (t & 1); t >>= 1; }
        if (i < j) std::swap(a[i], a[j]);
    }
}
static void ntt_inplace(std::vector<Fr>& a, unsigned log_n, bool inverse) {
    size_t n = (size_t)1 << log_n;
    if (n == 1) return;
    Fr w = fr_root_of_unity(log_n);
    if (inverse) w = w.inv();
    std::vector<Fr> tw(n / 2);
    {   // twiddle table w^i, built in blocks so it parallelises
        size_t B = 1024;
        size_t nb = (n / 2 + B - 1) / B;
        Fr wB = fr_pow_u64(w, B);
        std::vector<Fr> start(nb);
        Fr cur = Fr::one();
        for (size_t b = 0; b < nb; ++b) { start[b] = cur; cur = cur * wB; }
#pragma omp parallel for schedule(static)
        for (size_t b = 0; b < nb; ++b) {
            Fr c = start[b];
            for (size_t i = b * B; i < std::min(n / 2, (b + 1) * B); ++i) { tw[i] = c; c = c * w; }
        }
    }
    bitrev_permute(a, log_n);
    for (unsigned s = 1; s <= log_n; ++s) {
        size_t half = (size_t)1 << (s - 1), step = n >> s;
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n / 2; ++i) {
            size_t grp = i / half, k = i % half;
            size_t lo = grp * 2 * half + k, hi = lo + half;
            Fr u = a[lo], v = a[hi] * tw[k * step];
            a[lo] = u + v; a[hi] = u - v;
        }
    }
    if (inverse) {
        Fr ninv = Fr::from_u64(n).inv();
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) a[i] = a[i] * ninv;
    }
}
static void scale_by_powers(std::vector<Fr>& a, Fr g) {   // a[i] *= g^i
    size_t n = a.size(), B = 4096, nb = (n + B - 1) / B;
    Fr gB = fr_pow_u64(g, B);
    std::vector<Fr> start(nb);
    Fr cur = Fr::one();
    for (size_t b = 0; b < nb; ++b) { start[b] = cur; cur = cur * gB; }
#pragma omp parallel for schedule(static)
    for (size_t b = 0; b < nb; ++b) {
        Fr c = start[b];
        for (size_t i = b * B; i < std::min(n, (b + 1) * B); ++i) { a[i] = a[i] * c; c = c * g; }
    }
}
static void transform(std::vector<Fr>& a, unsigned log_n, bool inverse, bool coset) {
    Fr g = Fr::from_u64(7);
    if (coset && !inverse) scale_by_powers(a, g);
    ntt_inplace(a, log_n, inverse);
    if (coset && inverse) scale_by_powers(a, g.inv());
}

static bool fr_vec_from_bytes(std::vector<Fr>& v, const uint8_t* b, size_t n) {
    v.resize(n);
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (size_t i = 0; i < n; ++i) if (!Fr::from_bytes(v[i], b + 32 * i)) bad |= 1;
    return !bad;
}
static void fr_vec_to_bytes(const std::vector<Fr>& v, uint8_t* b) {
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < v.size(); ++i) v[i].to_bytes(b + 32 * i);
}

// -------------------------------------------------------------------------- MSM
// Bucket method (Pippenger / BDLO12-style), unsigned c-bit windows; points are
// split into one contiguous chunk per thread (crypto3's `chunks` argument), each
// chunk reduced independently, partial results summed.
static unsigned pick_window(size_t n) {
    unsigned c = 1;
    while (((size_t)1 << (c + 3)) < n && c < 16) ++c;   // ~ log2(n) - 3
    return std::max(c, 2u);
}
template <class F>
static Xyzz<F> msm_chunk(const Affine<F>* bases, const uint8_t* scal /*canonical LE*/, size_t n, unsigned c) {
    unsigned windows = (255 + c - 1) / c;
    std::vector<Xyzz<F>> buckets(((size_t)1 << c) - 1);
    Xyzz<F> total = Xyzz<F>::identity();
    for (int w = (int)windows - 1; w >= 0; --w) {
        for (unsigned k = 0; k < c; ++k) total = total.dbl();
        for (auto& b : buckets) b = Xyzz<F>::identity();
        unsigned bit0 = w * c;
        for (size_t i = 0; i < n; ++i) {
            if (bases[i].inf) continue;
            const uint8_t* s = scal + 32 * i;
            // extract c bits starting at bit0
            unsigned byte = bit0 / 8, sh = bit0 % 8;
            u64 v = 0;
            for (unsigned k = 0; k < 5 && byte + k < 32; ++k) v |= (u64)s[byte + k] << (8 * k);
            unsigned d = (unsigned)((v >> sh) & (((u64)1 << c) - 1));
            if (d) buckets[d - 1] = buckets[d - 1].add_mixed(bases[i]);
        }
        Xyzz<F> run = Xyzz<F>::identity(), sum = Xyzz<F>::identity();
        for (size_t k = buckets.size(); k-- > 0;) {
            run = run.add(buckets[k]);
            sum = sum.add(run);
        }
        total = total.add(sum);
    }
    return total;
}
template <class F>
static Xyzz<F> msm(const std::vector<Affine<F>>& bases, const uint8_t* scal, size_t n) {
    if (n == 0) return Xyzz<F>::identity();
    int T = 1;
#ifdef _OPENMP
    T = omp_get_max_threads();
#endif
    size_t chunks = std::min<size_t>(T, (n + 255) / 256);
    size_t per = (n + chunks - 1) / chunks;
    std::vector<Xyzz<F>> part(chunks);
#pragma omp parallel for schedule(dynamic, 1)
    for (size_t t = 0; t < chunks; ++t) {
        size_t lo = t * per, hi = std::min(n, lo + per);
        part[t] = lo < hi ? msm_chunk<F>(bases.data() + lo, scal + 32 * lo, hi - lo, pick_window(hi - lo))
                          : Xyzz<F>::identity();
    }
    Xyzz<F> r = Xyzz<F>::identity();
    for (auto& p : part) r = r.add(p);
    return r;
}
template <class F>
static bool bases_from_bytes(std::vector<Affine<F>>& v, const uint8_t* b, size_t n) {
    v.resize(n);
    int bad = 0; int sz = 2 * fsize<F>();
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (size_t i = 0; i < n; ++i) if (!affine_from_bytes<F>(v[i], b + (size_t)sz * i)) bad |= 1;
    return !bad;
}
static bool scalars_canonical(const uint8_t* s, size_t n) {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (size_t i = 0; i < n; ++i) { u64 t[4]; memcpy(t, s + 32 * i, 32); if (geq<4>(t, g_fr.p)) bad |= 1; }
    return !bad;
}

// fixed-base k*G with an 8-bit window table
template <class F> struct FixedBase {
    std::vector<Affine<F>> table;   // [32][255]
    void build(const Affine<F>& g) {
        std::vector<Xyzz<F>> t(32 * 255);
        Xyzz<F> cur = Xyzz<F>::from_affine(g);
        for (int w = 0; w < 32; ++w) {
            Xyzz<F> acc = Xyzz<F>::identity();
            for (int k = 0; k < 255; ++k) { acc = acc.add(cur); t[w * 255 + k] = acc; }
            cur = acc.add(cur);
        }
        batch_to_affine<F>(t, table);
    }
    Xyzz<F> mul(const uint8_t* k /*canonical LE*/) const {
        Xyzz<F> acc = Xyzz<F>::identity();
        for (int w = 0; w < 32; ++w) if (k[w]) acc = acc.add_mixed(table[w * 255 + k[w] - 1]);
        return acc;
    }
};
static FixedBase<Fq> g_fb1; static FixedBase<Fq2> g_fb2; static bool g_fb_init = false;
static void ensure_fb() {
    ensure_init();
    if (g_fb_init) return;
#pragma omp critical(orc_fb)
    { if (!g_fb_init) { g_fb1.build(g_g1); g_fb2.build(g_g2); g_fb_init = true; } }
}
template <class F>
static void gen_mul_many(const FixedBase<F>& fb, const std::vector<Fr>& k, uint8_t* out) {
    size_t n = k.size(), B = 1024;
    int sz = 2 * fsize<F>();
#pragma omp parallel for schedule(dynamic, 1)
    for (size_t b0 = 0; b0 < n; b0 += B) {
        size_t b1 = std::min(n, b0 + B);
        std::vector<Xyzz<F>> tmp(b1 - b0);
        std::vector<Affine<F>> aff;
        for (size_t i = b0; i < b1; ++i) { uint8_t kb[32]; k[i].to_bytes(kb); tmp[i - b0] = fb.mul(kb); }
        batch_to_affine<F>(tmp, aff);
        for (size_t i = b0; i < b1; ++i) affine_to_bytes<F>(aff[i - b0], out + (size_t)sz * i);
    }
}

// ------------------------------------------------------------------------- R1CS
struct orc_csr { u64 n_rows; u64 nnz; const uint32_t* row_ptr; const uint32_t* col; const uint8_t* val; };
struct orc_r1cs { uint32_t log_m; u64 n_cons; u64 n_in; u64 n_aux; orc_csr A, B, C; };

static bool spmv(const orc_csr& M, const std::vector<Fr>& z, Fr* out) {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (size_t i = 0; i < M.n_rows; ++i) {
        Fr acc = Fr::zero();
        for (uint32_t k = M.row_ptr[i]; k < M.row_ptr[i + 1]; ++k) {
            Fr v; if (!Fr::from_bytes(v, M.val + 32 * (size_t)k)) { bad |= 1; continue; }
            acc = acc + v * z[M.col[k]];
        }
        out[i] = acc;
    }
    return !bad;
}
// a,b,c on the domain, with the n_in+1 input-consistency rows appended to A.
static bool abc_evals(const orc_r1cs& cs, const std::vector<Fr>& z, std::vector<Fr>& a, std::vector<Fr>& b, std::vector<Fr>& c) {
    size_t m = (size_t)1 << cs.log_m;
    if (cs.n_cons + cs.n_in + 1 > m) return false;
    a.assign(m, Fr::zero()); b.assign(m, Fr::zero()); c.assign(m, Fr::zero());
    bool ok = spmv(cs.A, z, a.data()) & spmv(cs.B, z, b.data()) & spmv(cs.C, z, c.data());
    for (size_t j = 0; j <= cs.n_in; ++j) a[cs.n_cons + j] = z[j];
    return ok;
}
static void h_from_abc(std::vector<Fr>& a, std::vector<Fr>& b, std::vector<Fr>& c, unsigned log_m) {
    size_t m = (size_t)1 << log_m;
    std::vector<Fr>* v[3] = {&a, &b, &c};
    for (auto p : v) { transform(*p, log_m, true, false); transform(*p, log_m, false, true); }
    Fr zinv = (fr_pow_u64(Fr::from_u64(7), m) - Fr::one()).inv();
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < m; ++i) a[i] = (a[i] * b[i] - c[i]) * zinv;
    transform(a, log_m, true, true);
}

// ------------------------------------------------------------------ C interface
extern "C" {

int orc_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_threads(int t) {
#ifdef _OPENMP
    omp_set_num_threads(t);
#endif
    (void)t;
}

// 1 if the 32 LE bytes are a canonical Fr (fr32 contract, core/fr32.hpp:36-52)
int orc_fr_is_canonical(const uint8_t* b) { ensure_init(); u64 t[4]; memcpy(t, b, 32); return geq<4>(t, g_fr.p) ? 0 : 1; }

// elementwise field ops for kernel parity tests. op: 0 add 1 sub 2 mul 3 inv(a) 4 sqr(a) 5 neg(a)
int orc_fr_op(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
    ensure_init();
    for (size_t i = 0; i < n; ++i) {
        Fr x, y = Fr::zero(), r;
        if (!Fr::from_bytes(x, a + 32 * i)) return -2;
        if (op <= 2 && !Fr::from_bytes(y, b + 32 * i)) return -2;
        switch (op) { case 0: r = x + y; break; case 1: r = x - y; break; case 2: r = x * y; break;
            case 3: r = x.is_zero() ? x : x.inv(); break; case 4: r = x.sqr(); break; default: r = x.neg(); }
        r.to_bytes(out + 32 * i);
    }
    return 0;
}
int orc_fq_op(int op, const uint8_t* a, const uint8_t* b, size_t n, uint8_t* out) {
    ensure_init();
    for (size_t i = 0; i < n; ++i) {
        Fq x, y = Fq::zero(), r;
        if (!Fq::from_bytes(x, a + 48 * i)) return -2;
        if (op <= 2 && !Fq::from_bytes(y, b + 48 * i)) return -2;
        switch (op) { case 0: r = x + y; break; case 1: r = x - y; break; case 2: r = x * y; break;
            case 3: r = x.is_zero() ? x : x.inv(); break; case 4: r = x.sqr(); break; default: r = x.neg(); }
        r.to_bytes(out + 48 * i);
    }
    return 0;
}

int orc_ntt(uint8_t* data, unsigned log_n, int inverse, int coset) {
    ensure_init();
    std::vector<Fr> a;
    if (!fr_vec_from_bytes(a, data, (size_t)1 << log_n)) return -2;
    transform(a, log_n, inverse != 0, coset != 0);
    fr_vec_to_bytes(a, data);
    return 0;
}

int orc_msm_g1(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t* out96) {
    ensure_init();
    std::vector<G1A> b;
    if (!bases_from_bytes<Fq>(b, bases, n) || !scalars_canonical(scalars, n)) return -2;
    affine_to_bytes<Fq>(msm<Fq>(b, scalars, n).to_affine(), out96);
    return 0;
}
int orc_msm_g2(const uint8_t* bases, const uint8_t* scalars, size_t n, uint8_t* out192) {
    ensure_init();
    std::vector<G2A> b;
    if (!bases_from_bytes<Fq2>(b, bases, n) || !scalars_canonical(scalars, n)) return -2;
    affine_to_bytes<Fq2>(msm<Fq2>(b, scalars, n).to_affine(), out192);
    return 0;
}
// out[i] = k[i] * G  (affine ABI encoding)
int orc_g1_gen_mul(const uint8_t* k, size_t n, uint8_t* out) {
    ensure_fb(); std::vector<Fr> v; if (!fr_vec_from_bytes(v, k, n)) return -2;
    gen_mul_many<Fq>(g_fb1, v, out); return 0;
}
int orc_g2_gen_mul(const uint8_t* k, size_t n, uint8_t* out) {
    ensure_fb(); std::vector<Fr> v; if (!fr_vec_from_bytes(v, k, n)) return -2;
    gen_mul_many<Fq2>(g_fb2, v, out); return 0;
}
// generic point ops for tests: out = a + b ; out = k * a
int orc_g1_add(const uint8_t* a, const uint8_t* b, uint8_t* out) {
    ensure_init(); G1A x, y; if (!affine_from_bytes<Fq>(x, a) || !affine_from_bytes<Fq>(y, b)) return -2;
    affine_to_bytes<Fq>(G1::from_affine(x).add(G1::from_affine(y)).to_affine(), out); return 0;
}
int orc_g2_add(const uint8_t* a, const uint8_t* b, uint8_t* out) {
    ensure_init(); G2A x, y; if (!affine_from_bytes<Fq2>(x, a) || !affine_from_bytes<Fq2>(y, b)) return -2;
    affine_to_bytes<Fq2>(G2::from_affine(x).add(G2::from_affine(y)).to_affine(), out); return 0;
}
int orc_g1_mul(const uint8_t* a, const uint8_t* k, uint8_t* out) {
    ensure_init(); G1A x; Fr s; if (!affine_from_bytes<Fq>(x, a) || !Fr::from_bytes(s, k)) return -2;
    affine_to_bytes<Fq>(G1::from_affine(x).mul(s).to_affine(), out); return 0;
}
int orc_g2_mul(const uint8_t* a, const uint8_t* k, uint8_t* out) {
    ensure_init(); G2A x; Fr s; if (!affine_from_bytes<Fq2>(x, a) || !Fr::from_bytes(s, k)) return -2;
    affine_to_bytes<Fq2>(G2::from_affine(x).mul(s).to_affine(), out); return 0;
}
int orc_g1_on_curve(const uint8_t* a) {
    ensure_init(); G1A x; if (!affine_from_bytes<Fq>(x, a)) return 0; if (x.inf) return 1;
    return x.y.sqr() == x.x.sqr() * x.x + Fq::from_u64(4);
}
int orc_g2_on_curve(const uint8_t* a) {
    ensure_init(); G2A x; if (!affine_from_bytes<Fq2>(x, a)) return 0; if (x.inf) return 1;
    Fq2 b = {Fq::from_u64(4), Fq::from_u64(4)};
    return x.y.sqr() == x.x.sqr() * x.x + b;
}

int orc_spmv(const orc_csr* M, const uint8_t* z, size_t n_vars, uint8_t* out) {
    ensure_init();
    std::vector<Fr> zz; if (!fr_vec_from_bytes(zz, z, n_vars)) return -2;
    std::vector<Fr> o(M->n_rows);
    if (!spmv(*M, zz, o.data())) return -2;
    fr_vec_to_bytes(o, out); return 0;
}

// a,b,c: m*32 bytes each (evaluations on the domain) -> h: m*32 bytes (coefficients)
int orc_h_from_abc(const uint8_t* a, const uint8_t* b, const uint8_t* c, unsigned log_m, uint8_t* h) {
    ensure_init();
    size_t m = (size_t)1 << log_m;
    std::vector<Fr> va, vb, vc;
    if (!fr_vec_from_bytes(va, a, m) || !fr_vec_from_bytes(vb, b, m) || !fr_vec_from_bytes(vc, c, m)) return -2;
    h_from_abc(va, vb, vc, log_m);
    fr_vec_to_bytes(va, h);
    return 0;
}

// a,b,c evaluation vectors (m*32 each) from the constraint system and assignment
int orc_abc_evals(const orc_r1cs* cs, const uint8_t* z, uint8_t* a, uint8_t* b, uint8_t* c) {
    ensure_init();
    std::vector<Fr> zz; if (!fr_vec_from_bytes(zz, z, 1 + cs->n_in + cs->n_aux)) return -2;
    std::vector<Fr> va, vb, vc;
    if (!abc_evals(*cs, zz, va, vb, vc)) return -2;
    fr_vec_to_bytes(va, a); fr_vec_to_bytes(vb, b); fr_vec_to_bytes(vc, c);
    return 0;
}

// Synthetic witness: fills aux variables defined by C's single diagonal entry.
// Layered instance (tests/helpers synth): rows [r0,r1) only read vars < visible.
int orc_synth_witness(const orc_r1cs* cs, uint8_t* z /*in: 1+n_in+n_free set; out: all*/, u64 n_free) {
    ensure_init();
    size_t nv = 1 + cs->n_in + cs->n_aux, base = 1 + cs->n_in + n_free;
    std::vector<Fr> zz(nv, Fr::zero());
    for (size_t i = 0; i < base; ++i) if (!Fr::from_bytes(zz[i], z + 32 * i)) return -2;
    size_t row = 0;
    while (row < cs->n_cons) {
        size_t visible = base + row;
        size_t size = std::min<size_t>(std::max<size_t>(visible / 2, 1), cs->n_cons - row);
#pragma omp parallel for schedule(static)
        for (size_t i = row; i < row + size; ++i) {
            Fr sa = Fr::zero(), sb = Fr::zero();
            for (uint32_t k = cs->A.row_ptr[i]; k < cs->A.row_ptr[i + 1]; ++k) { Fr v; Fr::from_bytes(v, cs->A.val + 32 * (size_t)k); sa = sa + v * zz[cs->A.col[k]]; }
            for (uint32_t k = cs->B.row_ptr[i]; k < cs->B.row_ptr[i + 1]; ++k) { Fr v; Fr::from_bytes(v, cs->B.val + 32 * (size_t)k); sb = sb + v * zz[cs->B.col[k]]; }
            zz[base + i] = sa * sb;
        }
        row += size;
    }
    fr_vec_to_bytes(zz, z);
    return 0;
}

// ---- Groth16 key material, dense crypto3/libsnark style (one base per variable).
struct orc_pk {
    const uint8_t *alpha_g1, *beta_g1, *delta_g1;     // 96 B
    const uint8_t *beta_g2, *delta_g2;                // 192 B
    const uint8_t *a, *b_g1, *b_g2;                   // n_vars entries
    const uint8_t *l;                                 // n_aux entries
    const uint8_t *h;                                 // m-1 entries
};
struct orc_pk_out {
    uint8_t *alpha_g1, *beta_g1, *delta_g1, *beta_g2, *gamma_g2, *delta_g2;
    uint8_t *ic;      // (n_in+1) * 96
    uint8_t *a, *b_g1, *b_g2, *l, *h;
};

// toxic = tau, alpha, beta, gamma, delta (5 * 32 B canonical)
int orc_groth16_setup(const orc_r1cs* cs, const uint8_t* toxic, orc_pk_out* out) {
    ensure_fb();
    Fr tx[5];
    for (int i = 0; i < 5; ++i) if (!Fr::from_bytes(tx[i], toxic + 32 * i)) return -2;
    Fr tau = tx[0], alpha = tx[1], beta = tx[2], gamma = tx[3], delta = tx[4];
    size_t m = (size_t)1 << cs->log_m, nv = 1 + cs->n_in + cs->n_aux;
    if (cs->n_cons + cs->n_in + 1 > m) return -1;
    // Lagrange basis at tau: L_i = (tau^m - 1)/m * w^i / (tau - w^i)
    Fr w = fr_root_of_unity(cs->log_m);
    Fr zt = fr_pow_u64(tau, m) - Fr::one();
    Fr pref = zt * Fr::from_u64(m).inv();
    std::vector<Fr> L(m), wi(m);
    { Fr c = Fr::one(); for (size_t i = 0; i < m; ++i) { wi[i] = c; c = c * w; } }
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < m; ++i) L[i] = pref * wi[i] * (tau - wi[i]).inv();
    std::vector<Fr> u(nv, Fr::zero()), v(nv, Fr::zero()), ww(nv, Fr::zero());
    const orc_csr* Ms[3] = {&cs->A, &cs->B, &cs->C};
    std::vector<Fr>* acc[3] = {&u, &v, &ww};
    for (int t = 0; t < 3; ++t)
        for (size_t i = 0; i < cs->n_cons; ++i)
            for (uint32_t k = Ms[t]->row_ptr[i]; k < Ms[t]->row_ptr[i + 1]; ++k) {
                Fr val; if (!Fr::from_bytes(val, Ms[t]->val + 32 * (size_t)k)) return -2;
                Fr& d = (*acc[t])[Ms[t]->col[k]];
                d = d + val * L[i];
            }
    for (size_t j = 0; j <= cs->n_in; ++j) u[j] = u[j] + L[cs->n_cons + j];
    Fr gi = gamma.inv(), di = delta.inv();
    std::vector<Fr> one1 = {alpha}, k_ic(cs->n_in + 1), k_l(cs->n_aux), k_h(m - 1);
    for (size_t j = 0; j < nv; ++j) {
        Fr k = beta * u[j] + alpha * v[j] + ww[j];
        if (j <= cs->n_in) k_ic[j] = k * gi; else k_l[j - cs->n_in - 1] = k * di;
    }
    { Fr zd = zt * di, c = Fr::one(); for (size_t i = 0; i + 1 < m; ++i) { k_h[i] = c * zd; c = c * tau; } }
    auto one_g1 = [&](Fr k, uint8_t* o) { std::vector<Fr> t = {k}; gen_mul_many<Fq>(g_fb1, t, o); };
    auto one_g2 = [&](Fr k, uint8_t* o) { std::vector<Fr> t = {k}; gen_mul_many<Fq2>(g_fb2, t, o); };
    one_g1(alpha, out->alpha_g1); one_g1(beta, out->beta_g1); one_g1(delta, out->delta_g1);
    one_g2(beta, out->beta_g2); one_g2(gamma, out->gamma_g2); one_g2(delta, out->delta_g2);
    gen_mul_many<Fq>(g_fb1, k_ic, out->ic);
    gen_mul_many<Fq>(g_fb1, u, out->a);
    gen_mul_many<Fq>(g_fb1, v, out->b_g1);
    gen_mul_many<Fq2>(g_fb2, v, out->b_g2);
    gen_mul_many<Fq>(g_fb1, k_l, out->l);
    gen_mul_many<Fq>(g_fb1, k_h, out->h);
    return 0;
}

// Trace of every intermediate (any pointer may be null).
struct orc_trace {
    uint8_t *a, *b, *c;          // m*32 evaluations
    uint8_t *h;                  // m*32 coefficients
    uint8_t *msm_a, *msm_b1, *msm_l, *msm_h;   // 96
    uint8_t *msm_b2;             // 192
    uint8_t *A, *C;              // 96  affine proof points
    uint8_t *B;                  // 192
};

static void put_be48(const Fq& x, uint8_t* o) { uint8_t le[48]; x.to_bytes(le); for (int i = 0; i < 48; ++i) o[i] = le[47 - i]; }
static bool fq_gt_half(const Fq& y) {     // y > (p-1)/2  <=>  y > -y
    uint8_t a[48], b[48]; y.to_bytes(a); y.neg().to_bytes(b);
    for (int i = 47; i >= 0; --i) if (a[i] != b[i]) return a[i] > b[i];
    return false;
}
static void g1_compress(const G1A& p, uint8_t* o) {
    if (p.inf) { memset(o, 0, 48); o[0] = 0xC0; return; }
    put_be48(p.x, o); o[0] |= 0x80; if (fq_gt_half(p.y)) o[0] |= 0x20;
}
static void g2_compress(const G2A& p, uint8_t* o) {
    if (p.inf) { memset(o, 0, 96); o[0] = 0xC0; return; }
    put_be48(p.x.c1, o); put_be48(p.x.c0, o + 48); o[0] |= 0x80;
    bool larger = p.y.c1.is_zero() ? fq_gt_half(p.y.c0) : fq_gt_half(p.y.c1);
    if (larger) o[0] |= 0x20;
}
int orc_g1_compress(const uint8_t* a, uint8_t* o) { ensure_init(); G1A x; if (!affine_from_bytes<Fq>(x, a)) return -2; g1_compress(x, o); return 0; }
int orc_g2_compress(const uint8_t* a, uint8_t* o) { ensure_init(); G2A x; if (!affine_from_bytes<Fq2>(x, a)) return -2; g2_compress(x, o); return 0; }

// Groth16 prove.  If abc != null it holds a||b||c evaluations (3*m*32)
// and the constraint system is not consulted.
int orc_groth16_prove(const orc_pk* pk, const orc_r1cs* cs, const uint8_t* abc, const uint8_t* z,
                      const uint8_t* r32, const uint8_t* s32, uint8_t* proof192, orc_trace* tr) {
    ensure_init();
    size_t m = (size_t)1 << cs->log_m, nv = 1 + cs->n_in + cs->n_aux;
    Fr r, s; if (!Fr::from_bytes(r, r32) || !Fr::from_bytes(s, s32)) return -2;
    std::vector<Fr> zz; if (!fr_vec_from_bytes(zz, z, nv)) return -2;
    std::vector<Fr> a, b, c;
    if (abc) {
        if (!fr_vec_from_bytes(a, abc, m) || !fr_vec_from_bytes(b, abc + 32 * m, m) || !fr_vec_from_bytes(c, abc + 64 * m, m)) return -2;
    } else if (!abc_evals(*cs, zz, a, b, c)) return -2;
    if (tr && tr->a) fr_vec_to_bytes(a, tr->a);
    if (tr && tr->b) fr_vec_to_bytes(b, tr->b);
    if (tr && tr->c) fr_vec_to_bytes(c, tr->c);
    h_from_abc(a, b, c, cs->log_m);
    std::vector<uint8_t> hb(32 * m); fr_vec_to_bytes(a, hb.data());
    if (tr && tr->h) memcpy(tr->h, hb.data(), 32 * m);
    std::vector<G1A> qa, qb1, ql, qh; std::vector<G2A> qb2;
    if (!bases_from_bytes<Fq>(qa, pk->a, nv) || !bases_from_bytes<Fq>(qb1, pk->b_g1, nv) ||
        !bases_from_bytes<Fq2>(qb2, pk->b_g2, nv) || !bases_from_bytes<Fq>(ql, pk->l, cs->n_aux) ||
        !bases_from_bytes<Fq>(qh, pk->h, m - 1)) return -2;
    G1 ma = msm<Fq>(qa, z, nv), mb1 = msm<Fq>(qb1, z, nv);
    G2 mb2 = msm<Fq2>(qb2, z, nv);
    G1 ml = msm<Fq>(ql, z + 32 * (1 + cs->n_in), cs->n_aux);
    G1 mh = msm<Fq>(qh, hb.data(), m - 1);
    if (tr && tr->msm_a) affine_to_bytes<Fq>(ma.to_affine(), tr->msm_a);
    if (tr && tr->msm_b1) affine_to_bytes<Fq>(mb1.to_affine(), tr->msm_b1);
    if (tr && tr->msm_b2) affine_to_bytes<Fq2>(mb2.to_affine(), tr->msm_b2);
    if (tr && tr->msm_l) affine_to_bytes<Fq>(ml.to_affine(), tr->msm_l);
    if (tr && tr->msm_h) affine_to_bytes<Fq>(mh.to_affine(), tr->msm_h);
    G1A al, be1, de1; G2A be2, de2;
    if (!affine_from_bytes<Fq>(al, pk->alpha_g1) || !affine_from_bytes<Fq>(be1, pk->beta_g1) ||
        !affine_from_bytes<Fq>(de1, pk->delta_g1) || !affine_from_bytes<Fq2>(be2, pk->beta_g2) ||
        !affine_from_bytes<Fq2>(de2, pk->delta_g2)) return -2;
    G1 D1 = G1::from_affine(de1); G2 D2 = G2::from_affine(de2);
    G1 A = ma.add_mixed(al).add(D1.mul(r));
    G2 B = mb2.add_mixed(be2).add(D2.mul(s));
    G1 B1 = mb1.add_mixed(be1).add(D1.mul(s));
    G1 C = mh.add(ml).add(A.mul(s)).add(B1.mul(r)).add(D1.mul(r * s).neg());
    G1A Aa = A.to_affine(), Ca = C.to_affine(); G2A Ba = B.to_affine();
    if (tr && tr->A) affine_to_bytes<Fq>(Aa, tr->A);
    if (tr && tr->B) affine_to_bytes<Fq2>(Ba, tr->B);
    if (tr && tr->C) affine_to_bytes<Fq>(Ca, tr->C);
    g1_compress(Aa, proof192); g2_compress(Ba, proof192 + 48); g1_compress(Ca, proof192 + 144);
    return 0;
}

}  // extern "C"
