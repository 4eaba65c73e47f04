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
            case 3: r = x.inv(); break; case 4: r = x.sqr(); break; default: r = x.neg(); }
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
}
