// BN254 G1 (over Fq) and G2 (over Fq2) group law for sm_100a, y^2 = x^3 + b, a = 0.
//
// Replaces the reference's complete Jacobian add/dbl (mcl::EcT::add/dbl, MCL/include/mcl/ec.hpp:310-364,465-533,592-632;
// libff twin FF/algebra/curves/alt_bn128/alt_bn128_g1.cpp:139-358) with extended-Jacobian "XYZZ" coordinates
// (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2): a mixed add is 8M + 2S instead of 7M + 4S, and all the special cases the
// reference handles (P = Q, P = -Q, either operand at infinity) are handled here too.  Results are compared after
// normalisation to affine, where they are unique, so the coordinate system is not observable.
#pragma once
#include "fp.cuh"

namespace b200 {

// Affine point; the pair (0, 0) is not on either curve (b != 0) and encodes the point at infinity.
// Out-of-line group operations: one copy per translation unit instead of one per call site.  Only the mixed addition of
// the accumulate kernel is worth inlining; every other use (reductions, helpers) is latency-bound and the inlined
// bodies (a G2 addition is ~10^4 instructions) only cost compile time and instruction cache.
#ifdef B200_COLD_INLINE
#define B200_DEV_COLD __device__ __forceinline__
#else
#define B200_DEV_COLD __device__ __noinline__
#endif

template <typename F>
struct Affine {
    F x, y;
    B200_DEV bool is_inf() const { return x.is_zero() && y.is_zero(); }
    static B200_DEV Affine inf() { return Affine{F::zero(), F::zero()}; }
};

template <typename F>
struct XYZZ {
    F X, Y, ZZ, ZZZ;

    static B200_DEV XYZZ inf() { return XYZZ{F::zero(), F::one(), F::zero(), F::zero()}; }
    B200_DEV bool is_inf() const { return ZZ.is_zero(); }
    static B200_DEV XYZZ from_affine(const Affine<F>& p) {
        if (p.is_inf()) return inf();
        return XYZZ{p.x, p.y, F::one(), F::one()};
    }

    // 2 * (affine p), p not at infinity   (mdbl-2008-s-1)
    static B200_DEV_COLD XYZZ dbl_affine(const Affine<F>& p) {
        F U = p.y.dbl();
        if (U.is_zero()) return inf();
        F V = U.sqr(), W = U * V, S = p.x * V;
        F xx = p.x.sqr();
        F M = xx.dbl() + xx;
        XYZZ r;
        r.X = M.sqr() - S.dbl();
        r.Y = M * (S - r.X) - W * p.y;
        r.ZZ = V; r.ZZZ = W;
        return r;
    }
    // this = 2 * this   (dbl-2008-s-1)
    B200_DEV_COLD void dbl() {
        if (is_inf()) return;
        F U = Y.dbl();
        if (U.is_zero()) { *this = inf(); return; }
        F V = U.sqr(), W = U * V, S = X * V;
        F xx = X.sqr();
        F M = xx.dbl() + xx;
        F X3 = M.sqr() - S.dbl();
        Y = M * (S - X3) - W * Y;
        X = X3;
        ZZ = V * ZZ; ZZZ = W * ZZZ;
    }
    // this += affine p   (madd-2008-s: 8M + 2S), complete
    B200_DEV void add_mixed(const Affine<F>& p) {
        if (p.is_inf()) return;
        if (is_inf()) { X = p.x; Y = p.y; ZZ = F::one(); ZZZ = F::one(); return; }
        F U2 = p.x * ZZ, S2 = p.y * ZZZ;
        F Pp = U2 - X, R = S2 - Y;
        if (Pp.is_zero()) {
            if (R.is_zero()) *this = dbl_affine(p); else *this = inf();
            return;
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = X * PP;
        F X3 = R.sqr() - PPP - Q.dbl();
        Y = diff_of_products(R, Q - X3, Y, PPP);
        X = X3;
        ZZ = ZZ * PP; ZZZ = ZZZ * PPP;
    }
    // this += o   (add-2008-s: 12M + 2S), complete
    B200_DEV_COLD void add(const XYZZ& o) {
        if (o.is_inf()) return;
        if (is_inf()) { *this = o; return; }
        F U1 = X * o.ZZ, U2 = o.X * ZZ, S1 = Y * o.ZZZ, S2 = o.Y * ZZZ;
        F Pp = U2 - U1, R = S2 - S1;
        if (Pp.is_zero()) {
            if (R.is_zero()) dbl(); else *this = inf();
            return;
        }
        F PP = Pp.sqr(), PPP = Pp * PP, Q = U1 * PP;
        F X3 = R.sqr() - PPP - Q.dbl();
        Y = diff_of_products(R, Q - X3, S1, PPP);
        X = X3;
        ZZ = ZZ * o.ZZ * PP; ZZZ = ZZZ * o.ZZZ * PPP;
    }
    B200_DEV void negate() { Y = Y.neg(); }
    // k * this for a small unsigned k (double-and-add, MSB first)
    B200_DEV XYZZ mul_u32(uint32_t k) const {
        XYZZ r = inf();
        for (int i = 31; i >= 0; --i) {
            r.dbl();
            if ((k >> i) & 1) r.add(*this);
        }
        return r;
    }
    B200_DEV Affine<F> to_affine() const {
        if (is_inf()) return Affine<F>::inf();
        F I = (ZZ * ZZZ).inv();
        return Affine<F>{X * (I * ZZZ), Y * (I * ZZ)};
    }
    // the same with the binary-Euclid inverse: ~3x lower latency than the 380-product Fermat chain when ONE thread
    // normalises one point at the very end of a proof (g_finish_kernel), where nothing else hides it
    B200_DEV Affine<F> to_affine_low_latency() const {
        if (is_inf()) return Affine<F>::inf();
        F I = (ZZ * ZZZ).inv_binary();
        return Affine<F>{X * (I * ZZZ), Y * (I * ZZ)};
    }
};

typedef Affine<Fq> G1Affine;
typedef Affine<Fq2> G2Affine;
typedef XYZZ<Fq> G1XYZZ;
typedef XYZZ<Fq2> G2XYZZ;

}  // namespace b200
