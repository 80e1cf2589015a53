// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (plain C++17, unsigned __int128, no GMP, no mcl) of the BN254 / alt_bn128
// arithmetic that the reference's Groth16 prove path sits on.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this code.
//
// Reference semantics followed (paths relative to /root/reference, FF = depends/libsnark/depends/libff/libff):
//   Fp Montgomery mul/reduce  : FF/algebra/fields/fp.tcc:161-186 (HAC 14.32), add :405-418, sub :491-508,
//                               as_bigint :228-238, inverse :642-692
//   constants                 : FF/algebra/curves/mcl_bn128/mcl_bn128_init.cpp:41-111,170-174
//   Fq2 = Fq[u]/(u^2+1)       : FF/algebra/fields/fp2.tcc, MCL/include/mcl/fp_tower.hpp:199-225
//   G1/G2 Jacobian add/dbl    : FF/algebra/curves/alt_bn128/alt_bn128_g1.cpp:139-358 (and _g2.cpp)
//
// Parity status: see oracle/README.md ("pinning").
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>
#include <string>

namespace orc {

typedef uint64_t u64;
typedef unsigned __int128 u128;

struct FpParams {
    u64 p[4];
    u64 inv;      // -p^{-1} mod 2^64
    u64 r2[4];    // R^2 mod p
    u64 one[4];   // R mod p
};

inline int cmp4(const u64* a, const u64* b) {
    for (int i = 3; i >= 0; --i) { if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1; }
    return 0;
}
inline u64 add4(u64* r, const u64* a, const u64* b) {
    u128 c = 0;
    for (int i = 0; i < 4; ++i) { c += (u128)a[i] + b[i]; r[i] = (u64)c; c >>= 64; }
    return (u64)c;
}
inline u64 sub4(u64* r, const u64* a, const u64* b) {
    u64 borrow = 0;
    for (int i = 0; i < 4; ++i) {
        u128 d = (u128)a[i] - b[i] - borrow;
        r[i] = (u64)d; borrow = (u64)(d >> 64) & 1;
    }
    return borrow;
}

// TAG 0 = Fr (scalar field), TAG 1 = Fq (base field)
template <int TAG>
struct Fp {
    u64 v[4];
    static FpParams P;

    static Fp zero() { Fp r; memset(r.v, 0, 32); return r; }
    static Fp one() { Fp r; memcpy(r.v, P.one, 32); return r; }
    static Fp from_raw(const u64* m) { Fp r; memcpy(r.v, m, 32); return r; }   // already Montgomery
    static Fp from_std(const u64* s) { Fp t; memcpy(t.v, s, 32); Fp r2; memcpy(r2.v, P.r2, 32); return t * r2; }
    static Fp from_u64(u64 x) { u64 s[4] = {x, 0, 0, 0}; return from_std(s); }
    void to_std(u64* out) const {   // as_bigint: Montgomery multiply by 1 (fp.tcc:228-238)
        Fp o; o.v[0] = 1; o.v[1] = o.v[2] = o.v[3] = 0;
        Fp r = (*this) * o; memcpy(out, r.v, 32);
    }
    bool is_zero() const { return (v[0] | v[1] | v[2] | v[3]) == 0; }
    bool operator==(const Fp& o) const { return memcmp(v, o.v, 32) == 0; }
    bool operator!=(const Fp& o) const { return !(*this == o); }

    Fp operator+(const Fp& o) const {
        Fp r; u64 c = add4(r.v, v, o.v);
        if (c || cmp4(r.v, P.p) >= 0) sub4(r.v, r.v, P.p);
        return r;
    }
    Fp operator-(const Fp& o) const {
        Fp r; u64 b = sub4(r.v, v, o.v);
        if (b) add4(r.v, r.v, P.p);
        return r;
    }
    Fp operator-() const { if (is_zero()) return *this; Fp r; sub4(r.v, P.p, v); return r; }
    Fp dbl() const { return *this + *this; }

    // CIOS Montgomery multiplication, 4x64 limbs, R = 2^256
    Fp operator*(const Fp& o) const {
        u64 t[6] = {0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 4; ++i) {
            u128 c = 0;
            for (int j = 0; j < 4; ++j) {
                c += (u128)t[j] + (u128)v[j] * o.v[i];
                t[j] = (u64)c; c >>= 64;
            }
            c += t[4]; t[4] = (u64)c; t[5] = (u64)(c >> 64);
            u64 m = t[0] * P.inv;
            c = (u128)t[0] + (u128)m * P.p[0]; c >>= 64;
            for (int j = 1; j < 4; ++j) {
                c += (u128)t[j] + (u128)m * P.p[j];
                t[j - 1] = (u64)c; c >>= 64;
            }
            c += t[4]; t[3] = (u64)c; c >>= 64;
            t[4] = t[5] + (u64)c;
        }
        Fp r; memcpy(r.v, t, 32);
        if (t[4] || cmp4(r.v, P.p) >= 0) sub4(r.v, r.v, P.p);
        return r;
    }
    Fp sqr() const { return (*this) * (*this); }
    Fp pow(const u64* e, int nlimbs) const {
        Fp r = one(); bool started = false;
        for (int i = nlimbs * 64 - 1; i >= 0; --i) {
            if (started) r = r.sqr();
            if ((e[i / 64] >> (i % 64)) & 1) { r = started ? r * (*this) : *this; started = true; }
        }
        return r;
    }
    Fp pow_u64(u64 e) const { return pow(&e, 1); }
    Fp inv() const {   // Fermat: a^(p-2); a = 0 -> 0
        u64 e[4]; u64 two[4] = {2, 0, 0, 0}; sub4(e, P.p, two);
        return pow(e, 4);
    }
    Fp& operator+=(const Fp& o) { *this = *this + o; return *this; }
    Fp& operator-=(const Fp& o) { *this = *this - o; return *this; }
    Fp& operator*=(const Fp& o) { *this = *this * o; return *this; }
};
template <int TAG> FpParams Fp<TAG>::P;

typedef Fp<0> Fr;
typedef Fp<1> Fq;

// Fq2 = Fq[u]/(u^2 + 1): element a + b*u  (mcl: Fp2T{a,b}; libff: c0 + c1*u)
struct Fq2 {
    Fq a, b;
    static Fq2 zero() { return Fq2{Fq::zero(), Fq::zero()}; }
    static Fq2 one() { return Fq2{Fq::one(), Fq::zero()}; }
    bool is_zero() const { return a.is_zero() && b.is_zero(); }
    bool operator==(const Fq2& o) const { return a == o.a && b == o.b; }
    bool operator!=(const Fq2& o) const { return !(*this == o); }
    Fq2 operator+(const Fq2& o) const { return Fq2{a + o.a, b + o.b}; }
    Fq2 operator-(const Fq2& o) const { return Fq2{a - o.a, b - o.b}; }
    Fq2 operator-() const { return Fq2{-a, -b}; }
    Fq2 dbl() const { return Fq2{a.dbl(), b.dbl()}; }
    Fq2 operator*(const Fq2& o) const {
        Fq aa = a * o.a, bb = b * o.b;
        Fq m = (a + b) * (o.a + o.b);
        return Fq2{aa - bb, m - aa - bb};
    }
    Fq2 sqr() const {
        Fq ab = a * b;
        return Fq2{(a + b) * (a - b), ab.dbl()};
    }
    Fq2 inv() const {
        Fq n = (a.sqr() + b.sqr()).inv();
        return Fq2{a * n, -(b * n)};
    }
    Fq2& operator+=(const Fq2& o) { *this = *this + o; return *this; }
};

// Short Weierstrass y^2 = x^3 + b, a = 0, Jacobian coordinates (X/Z^2, Y/Z^3); Z = 0 <=> infinity.
template <typename F>
struct Jac {
    F X, Y, Z;
    static Jac zero() { return Jac{F::zero(), F::one(), F::zero()}; }
    static Jac from_affine(const F& x, const F& y) { return Jac{x, y, F::one()}; }
    bool is_zero() const { return Z.is_zero(); }

    Jac dbl() const {   // dbl-2009-l  (alt_bn128_g1.cpp:324-358)
        if (is_zero()) return *this;
        F A = X.sqr(), B = Y.sqr(), C = B.sqr();
        F D = ((X + B).sqr() - A - C).dbl();
        F E = A.dbl() + A, Fv = E.sqr();
        F X3 = Fv - D.dbl();
        F eightC = C.dbl().dbl().dbl();
        F Y3 = E * (D - X3) - eightC;
        F Z3 = (Y * Z).dbl();
        return Jac{X3, Y3, Z3};
    }
    Jac add(const Jac& o) const {   // add-2007-bl with the equal / opposite cases (alt_bn128_g1.cpp:139-254)
        if (is_zero()) return o;
        if (o.is_zero()) return *this;
        F Z1Z1 = Z.sqr(), Z2Z2 = o.Z.sqr();
        F U1 = X * Z2Z2, U2 = o.X * Z1Z1;
        F S1 = Y * o.Z * Z2Z2, S2 = o.Y * Z * Z1Z1;
        if (U1 == U2) {
            if (S1 == S2) return dbl();
            return zero();
        }
        F H = U2 - U1, I = H.dbl().sqr(), J = H * I;
        F r = (S2 - S1).dbl(), V = U1 * I;
        F X3 = r.sqr() - J - V.dbl();
        F Y3 = r * (V - X3) - (S1 * J).dbl();
        F Z3 = ((Z + o.Z).sqr() - Z1Z1 - Z2Z2) * H;
        return Jac{X3, Y3, Z3};
    }
    Jac add_affine(const F& x2, const F& y2) const {   // madd-2007-bl (alt_bn128_g1.cpp:256-322)
        if (is_zero()) return from_affine(x2, y2);
        F Z1Z1 = Z.sqr();
        F U2 = x2 * Z1Z1, S2 = y2 * Z * Z1Z1;
        if (X == U2) {
            if (Y == S2) return dbl();
            return zero();
        }
        F H = U2 - X, HH = H.sqr(), I = HH.dbl().dbl(), J = H * I;
        F r = (S2 - Y).dbl(), V = X * I;
        F X3 = r.sqr() - J - V.dbl();
        F Y3 = r * (V - X3) - (Y * J).dbl();
        F Z3 = (Z + H).sqr() - Z1Z1 - HH;
        return Jac{X3, Y3, Z3};
    }
    Jac neg() const { return Jac{X, -Y, Z}; }
    void to_affine(F& x, F& y, bool& inf) const {
        if (is_zero()) { inf = true; x = F::zero(); y = F::one(); return; }
        F zi = Z.inv(), zi2 = zi.sqr();
        x = X * zi2; y = Y * zi2 * zi; inf = false;
    }
    // scalar in standard (non-Montgomery) little-endian limbs
    Jac mul(const u64* k, int nlimbs = 4) const {
        Jac r = zero(); bool started = false;
        for (int i = nlimbs * 64 - 1; i >= 0; --i) {
            if (started) r = r.dbl();
            if ((k[i / 64] >> (i % 64)) & 1) { r = r.add(*this); started = true; }
        }
        return r;
    }
    bool eq(const Jac& o) const {
        if (is_zero() || o.is_zero()) return is_zero() && o.is_zero();
        F Z1Z1 = Z.sqr(), Z2Z2 = o.Z.sqr();
        if (X * Z2Z2 != o.X * Z1Z1) return false;
        return Y * o.Z * Z2Z2 == o.Y * Z * Z1Z1;
    }
};
typedef Jac<Fq> G1;
typedef Jac<Fq2> G2;

void init();                     // fills Fr::P, Fq::P and generators; idempotent
extern Fr FR_ROOT_OF_UNITY;      // 2^28-th primitive root (mcl_bn128_init.cpp:61)
extern Fr FR_MULT_GEN;           // 7 (mcl_bn128_init.cpp:60)
extern G1 G1_ONE;
extern G2 G2_ONE;
Fr fr_from_dec(const char* s);
Fq fq_from_dec(const char* s);
std::string hex_std(const u64* s);   // lower-case hex without leading zeros ("0" for zero), like mpz_get_str(.,16,.)

}  // namespace orc
