#pragma once
// G2 accumulate with one Fq2 point spread over a PAIR of adjacent lanes: the even lane holds the c0 ("a") component of
// X, Y, ZZ, ZZZ, the odd lane the c1 ("b") component.  Same chunk schedule and same outputs as msm_accumulate<G2T>
// (msm_acc.cuh); the difference is the register footprint: the one-thread kernel needs 254 registers (two resident
// blocks per SM, IMAD pipe 64 % busy, ncu r1), a lane of the pair holds half the state.
//   x * y   in Fq2 = (xa ya - xb yb) + (xa yb + xb ya) u: each lane computes ITS component as one fused two-product
//           Montgomery row (Fq::dot2, 200 multiply-adds) - 400 per Fq2 product against 408 for one-thread Karatsuba;
//   x^2     = (xa + xb)(xa - xb) + (2 xa xb) u: one plain product per lane.
// The partner's component of each operand comes through __shfl_xor_sync(.., 1) (11 exchanges of 8 limbs per mixed add,
// ~3 % of the executed instructions).  Control flow is uniform within a pair (both lanes walk the same chunk and all the
// special-case predicates are combined over the pair), so the pair mask is always converged.
// Measured on the 2^20 B query (B200): 8.24 ms -> 7.54 ms, IMAD pipe 64 % -> 71 % busy (profiles/r1_ncu_g2_pair.txt).
#include "msm_common.cuh"

namespace b200 {

#ifndef B200_PAIR_FUSED_Y
#define B200_PAIR_FUSED_Y 1     // R (Q - X3) - Y PPP as one four-product row set (PairLane::diff)
#endif

struct HalfXYZZ { Fq X, Y, ZZ, ZZZ; };   // one Fq2 component of a G2 XYZZ point

// (fp.cuh: neg_lazy, add_lazy, sub_lazy; the products themselves still return fully reduced values, so everything
// stored or compared is canonical in every mode.)
static __device__ __noinline__ Fq pair_dot2_ool(Fq a0, Fq b0, Fq a1, Fq b1) { return Fq::dot2(a0, b0, a1, b1); }
static __device__ __noinline__ Fq pair_mul_ool(Fq a, Fq b) { return a * b; }

// MODE 1: strict operand preparation.  MODE 2 (LAZY): the operands of the products are prepared without their own
// reductions.  MODE 3: LAZY and the products are calls to ONE out-of-line copy each (the inlined body is ~150 KB of SASS
// and ncu shows 15 % of the stall samples waiting for instruction fetch).
template <int MODE>
struct PairLane {
    static constexpr bool LAZY = MODE >= 2;
    static constexpr bool OOL = MODE == 3;
    bool imag;        // this lane holds the b (u) component
    uint32_t mask;    // the two lanes of the pair
    B200_DEV Fq other(const Fq& x) const {
        Fq r;
        _Pragma("unroll") for (int i = 0; i < 8; ++i) r.v[i] = __shfl_xor_sync(mask, x.v[i], 1);
        return r;
    }
    B200_DEV bool both(bool f) const { const int o = __shfl_xor_sync(mask, (int)f, 1); return f && o; }   // no short circuit: both lanes shuffle
    static B200_DEV Fq sel(bool c, const Fq& t, const Fq& f) {
        Fq r;
        _Pragma("unroll") for (int i = 0; i < 8; ++i) r.v[i] = c ? t.v[i] : f.v[i];
        return r;
    }
    // my component of x * y  (xm, ym mine; xo, yo the partner's)
    B200_DEV Fq mul(const Fq& xm, const Fq& xo, const Fq& ym, const Fq& yo) const {
        const Fq p = sel(imag, xo, xm);             // real: xa ya + (-xb) yb      imag: xa yb + xb ya
        const Fq q = sel(imag, xm, LAZY ? xo.neg_lazy() : xo.neg());
        if (OOL) return pair_dot2_ool(p, ym, q, yo);
        return Fq::dot2(p, ym, q, yo);
    }
    // my component of x^2
    B200_DEV Fq sqr(const Fq& xm, const Fq& xo) const {
        const Fq s = sel(imag, xm, LAZY ? Fq::add_lazy(xm, xo) : xm + xo);
        const Fq t = sel(imag, LAZY ? Fq::add_lazy(xo, xo) : xo.dbl(), LAZY ? Fq::sub_lazy(xm, xo) : xm - xo);
        if (OOL) return pair_mul_ool(s, t);
        return s * t;
    }
    // my component of x * y - z * w under ONE reduction (Fq::dot4: 328 multiply-adds against 400 for two products)
    B200_DEV Fq diff(const Fq& xm, const Fq& xo, const Fq& ym, const Fq& yo, const Fq& zm, const Fq& zo, const Fq& wm, const Fq& wo) const {
#if B200_PAIR_FUSED_Y
        const Fq nzm = zm.neg_lazy(), nzo = zo.neg_lazy();
        const Fq p = sel(imag, xo, xm), q = sel(imag, xm, xo.neg_lazy());          // real: xa ya - xb yb - za wa + zb wb
        const Fq p2 = sel(imag, nzo, nzm), q2 = sel(imag, nzm, zo);                // imag: xa yb + xb ya - za wb - zb wa
        return Fq::dot4(p, ym, q, yo, p2, wm, q2, wo);
#else
        return mul(xm, xo, ym, yo) - mul(zm, zo, wm, wo);
#endif
    }
    B200_DEV Fq one() const { return imag ? Fq::zero() : Fq::one(); }
    B200_DEV HalfXYZZ inf() const { return HalfXYZZ{Fq::zero(), one(), Fq::zero(), Fq::zero()}; }
};

// 2 * (affine p), p not at infinity (mdbl-2008-s-1); rare: only when a bucket meets the same point twice
template <int MODE>
static __device__ __noinline__ HalfXYZZ pair_dbl_affine(PairLane<MODE> ln, Fq px, Fq pxo, Fq py, Fq pyo) {
    const Fq U = py.dbl(), Uo = pyo.dbl();
    if (U.is_zero() && Uo.is_zero()) return ln.inf();
    const Fq V = ln.sqr(U, Uo), Vo = ln.other(V);
    const Fq W = ln.mul(U, Uo, V, Vo), Wo = ln.other(W);
    const Fq S = ln.mul(px, pxo, V, Vo);
    const Fq xx = ln.sqr(px, pxo);
    const Fq M = xx.dbl() + xx, Mo = ln.other(M);
    HalfXYZZ r;
    r.X = ln.sqr(M, Mo) - S.dbl();
    const Fq D = S - r.X, Do = ln.other(D);
    r.Y = ln.mul(M, Mo, D, Do) - ln.mul(W, Wo, py, pyo);
    r.ZZ = V; r.ZZZ = W;
    return r;
}

// a += affine p   (madd-2008-s: 8M + 2S in Fq2), complete.  Every branch condition is the same in both lanes of the pair.
// (A warp-converged formulation - full-mask shuffles, every special case resolved by selects after the generic sum - was
// measured slower: 8.2 ms against 7.5 ms for the 2^20 B query; the partial-mask shuffles are not what limits this kernel.)
template <int MODE>
static B200_DEV void pair_add_mixed(const PairLane<MODE>& ln, HalfXYZZ& a, const Fq& px, const Fq& pxo, const Fq& py, const Fq& pyo) {
    if (px.is_zero() && py.is_zero() && pxo.is_zero() && pyo.is_zero()) return;   // p at infinity
    const Fq zzo = ln.other(a.ZZ);
    if (a.ZZ.is_zero() && zzo.is_zero()) { a.X = px; a.Y = py; a.ZZ = ln.one(); a.ZZZ = ln.one(); return; }
    const Fq zzzo = ln.other(a.ZZZ);
    const Fq Pp = ln.mul(px, pxo, a.ZZ, zzo) - a.X, R = ln.mul(py, pyo, a.ZZZ, zzzo) - a.Y;
    const Fq Ppo = ln.other(Pp), Ro = ln.other(R);
    if (Pp.is_zero() && Ppo.is_zero()) {
        if (R.is_zero() && Ro.is_zero()) a = pair_dbl_affine(ln, px, pxo, py, pyo); else a = ln.inf();
        return;
    }
    const Fq PP = ln.sqr(Pp, Ppo), PPo = ln.other(PP);
    const Fq PPP = ln.mul(Pp, Ppo, PP, PPo), PPPo = ln.other(PPP);
    const Fq Q = ln.mul(a.X, ln.other(a.X), PP, PPo);
    const Fq X3 = ln.sqr(R, Ro) - PPP - Q.dbl();
    const Fq D = Q - X3;
    a.Y = ln.diff(R, Ro, D, ln.other(D), a.Y, ln.other(a.Y), PPP, PPPo);
    a.X = X3;
    a.ZZ = ln.mul(a.ZZ, zzo, PP, PPo);
    a.ZZZ = ln.mul(a.ZZZ, zzzo, PPP, PPPo);
}

// 2 * a for a general XYZZ point (dbl-2008-s-1); rare: only when two partial sums of one bucket coincide
template <int MODE>
static __device__ __noinline__ HalfXYZZ pair_dbl(PairLane<MODE> ln, HalfXYZZ a) {
    const Fq U = a.Y.dbl(), Uo = ln.other(U);
    if (U.is_zero() && Uo.is_zero()) return ln.inf();
    const Fq V = ln.sqr(U, Uo), Vo = ln.other(V);
    const Fq W = ln.mul(U, Uo, V, Vo), Wo = ln.other(W);
    const Fq S = ln.mul(a.X, ln.other(a.X), V, Vo);
    const Fq xx = ln.sqr(a.X, ln.other(a.X));
    const Fq M = xx.dbl() + xx, Mo = ln.other(M);
    HalfXYZZ r;
    r.X = ln.sqr(M, Mo) - S.dbl();
    const Fq D = S - r.X;
    r.Y = ln.mul(M, Mo, D, ln.other(D)) - ln.mul(W, Wo, a.Y, ln.other(a.Y));
    r.ZZ = ln.mul(V, Vo, a.ZZ, ln.other(a.ZZ));
    r.ZZZ = ln.mul(W, Wo, a.ZZZ, ln.other(a.ZZZ));
    return r;
}

// a += b, both XYZZ   (add-2008-s: 12M + 2S in Fq2), complete; every branch condition is the same in both lanes of the pair.
// The one-thread form (XYZZ<Fq2>::add) needs 174-184 registers in the reduction kernels: 1.9 warps per scheduler and the IMAD
// pipe 45-57 % busy (profiles/r1_ncu_other_kernels.txt).
template <int MODE>
static B200_DEV void pair_add_full(const PairLane<MODE>& ln, HalfXYZZ& a, const HalfXYZZ& b) {
    const Fq bzo = ln.other(b.ZZ);
    if (b.ZZ.is_zero() && bzo.is_zero()) return;
    const Fq azo = ln.other(a.ZZ);
    if (a.ZZ.is_zero() && azo.is_zero()) { a = b; return; }
    const Fq azzzo = ln.other(a.ZZZ), bzzzo = ln.other(b.ZZZ);
    const Fq U1 = ln.mul(a.X, ln.other(a.X), b.ZZ, bzo), U2 = ln.mul(b.X, ln.other(b.X), a.ZZ, azo);
    const Fq S1 = ln.mul(a.Y, ln.other(a.Y), b.ZZZ, bzzzo), S2 = ln.mul(b.Y, ln.other(b.Y), a.ZZZ, azzzo);
    const Fq Pp = U2 - U1, R = S2 - S1;
    const Fq Ppo = ln.other(Pp), Ro = ln.other(R);
    if (Pp.is_zero() && Ppo.is_zero()) {
        if (R.is_zero() && Ro.is_zero()) a = pair_dbl(ln, a); else a = ln.inf();
        return;
    }
    const Fq PP = ln.sqr(Pp, Ppo), PPo = ln.other(PP);
    const Fq PPP = ln.mul(Pp, Ppo, PP, PPo), PPPo = ln.other(PPP);
    const Fq Q = ln.mul(U1, ln.other(U1), PP, PPo);
    const Fq X3 = ln.sqr(R, Ro) - PPP - Q.dbl();
    const Fq D = Q - X3;
    a.Y = ln.diff(R, Ro, D, ln.other(D), S1, ln.other(S1), PPP, PPPo);
    a.X = X3;
    const Fq Z2 = ln.mul(a.ZZ, azo, b.ZZ, bzo);
    a.ZZ = ln.mul(Z2, ln.other(Z2), PP, PPo);
    const Fq Z3 = ln.mul(a.ZZZ, azzzo, b.ZZZ, bzzzo);
    a.ZZZ = ln.mul(Z3, ln.other(Z3), PPP, PPPo);
}

static B200_DEV HalfXYZZ pair_load(const G2XYZZ* src, bool imag) {
    const Fq* f = (const Fq*)src + (imag ? 1 : 0);      // X.a X.b Y.a Y.b ZZ.a ZZ.b ZZZ.a ZZZ.b
    return HalfXYZZ{f[0], f[2], f[4], f[6]};
}
static B200_DEV void pair_store(G2XYZZ* dst, bool imag, const HalfXYZZ& a) {
    Fq* f = (Fq*)dst + (imag ? 1 : 0);      // X.a X.b Y.a Y.b ZZ.a ZZ.b ZZZ.a ZZZ.b
    f[0] = a.X; f[2] = a.Y; f[4] = a.ZZ; f[6] = a.ZZZ;
}

#ifndef B200_G2_PAIR_MINB
#define B200_G2_PAIR_MINB 3     // 168 registers, no spills (4 -> 128 registers with spills: 7.8 ms against 7.5 ms)
#endif
#ifndef B200_G2_PAIR_BLOCK
#define B200_G2_PAIR_BLOCK 128  // threads per block (registers are allocated per warp: smaller blocks give finer occupancy steps)
#endif
#ifdef B200_G2_PAIR_MAXREG        // experiment knob: an explicit register cap instead of the launch bounds
#define B200_G2_PAIR_BOUNDS __maxnreg__(B200_G2_PAIR_MAXREG)
#else
#define B200_G2_PAIR_BOUNDS __launch_bounds__(B200_G2_PAIR_BLOCK, B200_G2_PAIR_MINB)
#endif
template <typename G, int MODE>   // G2T only; a template so that only the G2 translation unit emits it
__global__ void B200_G2_PAIR_BOUNDS msm_accumulate_g2_pair(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                                             const uint64_t* __restrict__ offsets, uint32_t nb, uint32_t L,
                                                                             const G2Affine* __restrict__ bases, G2XYZZ* __restrict__ buckets,
                                                                             G2XYZZ* __restrict__ pfirst, G2XYZZ* __restrict__ plast) {
    const uint64_t M = offsets[nb];
    const uint64_t gt = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t t = gt >> 1;
    const uint64_t start = t * L;
    if (start >= M) return;
    PairLane<MODE> ln;
    ln.imag = gt & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    const uint64_t end = min(start + (uint64_t)L, M);
    uint32_t cur = keys[start];
    bool first_run = true;
    HalfXYZZ acc = ln.inf();
    for (uint64_t i = start; i < end; ++i) {
        const uint32_t k = keys[i];
        if (k != cur) {
            pair_store(first_run ? &pfirst[t] : &buckets[cur], ln.imag, acc);
            first_run = false;
            acc = ln.inf();
            cur = k;
        }
        const uint32_t v = vals ? vals[i] : (uint32_t)i;
        const Fq* b = (const Fq*)&bases[v & 0x7fffffffu];       // x.a x.b y.a y.b: a pair reads one 128-byte line
        Fq px = b[ln.imag ? 1 : 0], py = b[ln.imag ? 3 : 2];
        if (v >> 31) py = py.neg();
        const Fq pxo = ln.other(px), pyo = ln.other(py);
        pair_add_mixed(ln, acc, px, pxo, py, pyo);
    }
    pair_store(first_run ? &pfirst[t] : &plast[t], ln.imag, acc);
}

}  // namespace b200
