// Short-Weierstrass (a = 0) group law in XYZZ coordinates over Fq (G1) / Fq2 (G2).
//
// Replaces crypto3-algebra `curves::bls12<381>::g1_type<>/g2_type<>` value types
// (un-vendored; used by the reference at
//  libs/storage/include/nil/filecoin/storage/proofs/core/crypto/scheme_params.hpp:40-66).
// x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2; identity <=> ZZ == 0.  Formulas: SURVEY.md App. E
// (EFD madd-2008-s / add-2008-s / dbl-2008-s-1, public-domain math).
// Affine identity is encoded (0,0), which is on neither curve.
#pragma once
#include "fp.cuh"

namespace b2 {

template <class F> struct Affine {
    F x, y;
    B2_HD bool is_identity() const { return x.is_zero() && y.is_zero(); }
    B2_HD static Affine identity() { return {F::zero(), F::zero()}; }
    B2_HD Affine neg() const { return {x, y.neg()}; }
};

template <class F> struct Xyzz {
    F X, Y, ZZ, ZZZ;
    B2_HD static Xyzz identity() { return {F::zero(), F::zero(), F::zero(), F::zero()}; }
    B2_HD bool is_identity() const { return ZZ.is_zero(); }
    B2_HD static Xyzz from_affine(const Affine<F>& a) {
        if (a.is_identity()) return identity();
        return {a.x, a.y, F::one(), F::one()};
    }
    B2_HD Xyzz neg() const { return {X, Y.neg(), ZZ, ZZZ}; }

    B2_HD Xyzz dbl() const {
        if (is_identity()) return *this;
        F U = Y.dbl(), V = U.sqr(), W = U * V, S = X * V;
        F X2 = X.sqr(), M = X2.dbl() + X2;
        F X3 = M.sqr() - S.dbl();
        F Y3 = M * (S - X3) - W * Y;
        return {X3, Y3, V * ZZ, W * ZZZ};
    }
    // 2 * (affine point), 8 products
    B2_HD static Xyzz dbl_affine(const Affine<F>& a) {
        F U = a.y.dbl(), V = U.sqr(), W = U * V, S = a.x * V;
        F X2 = a.x.sqr(), M = X2.dbl() + X2;
        F X3 = M.sqr() - S.dbl();
        F Y3 = M * (S - X3) - W * a.y;
        return {X3, Y3, V, W};
    }
    // acc += affine (8M + 2S).  `a` must not be the identity unless checked by caller;
    // the identity test here keeps the function total.
    B2_HD void add_mixed(const Affine<F>& a) {
        if (a.is_identity()) return;
        if (is_identity()) { X = a.x; Y = a.y; ZZ = F::one(); ZZZ = F::one(); return; }
        F U2 = a.x * ZZ, S2 = a.y * ZZZ;
        F Pp = U2 - X, Rr = S2 - Y;
        if (Pp.is_zero()) {
            if (Rr.is_zero()) *this = dbl_affine(a); else *this = identity();
            return;
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = X * PP;
        F X3 = Rr.sqr() - PPP - Q.dbl();
        Y = Rr * (Q - X3) - Y * PPP;
        X = X3;
        ZZ = ZZ * PP;
        ZZZ = ZZZ * PPP;
    }
    // acc += other (12M + 2S)
    B2_HD void add(const Xyzz& o) {
        if (o.is_identity()) return;
        if (is_identity()) { *this = o; return; }
        F U1 = X * o.ZZ, U2 = o.X * ZZ, S1 = Y * o.ZZZ, S2 = o.Y * ZZZ;
        F Pp = U2 - U1, Rr = S2 - S1;
        if (Pp.is_zero()) {
            if (Rr.is_zero()) *this = dbl(); else *this = identity();
            return;
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = U1 * PP;
        F X3 = Rr.sqr() - PPP - Q.dbl();
        Y = Rr * (Q - X3) - S1 * PPP;
        X = X3;
        ZZ = ZZ * o.ZZ * PP;
        ZZZ = ZZZ * o.ZZZ * PPP;
    }
    B2_HD Affine<F> to_affine() const {
        if (is_identity()) return Affine<F>::identity();
        F I = (ZZ * ZZZ).inv();
        return {X * I * ZZZ, Y * I * ZZ};
    }
    // the same point with the inversion by the binary Euclidean algorithm (fp.cuh inv_gcd): a much shorter serial chain
    // than the Fermat ladder, which is what a lone thread at the end of a multiexp pays for
    B2_HD Affine<F> to_affine_gcd() const {
        if (is_identity()) return Affine<F>::identity();
        F I = (ZZ * ZZZ).inv_gcd();
        return {X * I * ZZZ, Y * I * ZZ};
    }
};

#if defined(__CUDACC__)
// Out-of-line group operations for everything except the bucket-accumulation loop
// (code size / compile time; these run O(buckets) times, not O(points)).
template <class F> static __device__ __noinline__ void xyzz_add_call(Xyzz<F>* a, const Xyzz<F>* b) { a->add(*b); }
template <class F> static __device__ __noinline__ void xyzz_add_mixed_call(Xyzz<F>* a, const Affine<F>* b) { a->add_mixed(*b); }
template <class F> static __device__ __noinline__ void xyzz_dbl_call(Xyzz<F>* a) { *a = a->dbl(); }
template <class F> static __device__ __noinline__ void xyzz_to_affine_call(Affine<F>* r, const Xyzz<F>* a) { *r = a->to_affine(); }
#endif

typedef Affine<Fq> G1Affine;
typedef Affine<Fq2> G2Affine;
typedef Xyzz<Fq> G1Xyzz;
typedef Xyzz<Fq2> G2Xyzz;

}  // namespace b2
