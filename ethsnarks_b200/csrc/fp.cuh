// 254-bit prime-field arithmetic for BN254 (Fr and Fq) on sm_100a: 8 x 32-bit limbs, Montgomery form, R = 2^256.
//
// Replaces (B200-native, not a translation) the reference's libff::Fp_model<4, modulus>:
//   mul_reduce  FF/algebra/fields/fp.tcc:23-186      add :310-420   sub :423-510   as_bigint :228-238   inverse :642-692
// (FF = /root/reference/depends/libsnark/depends/libff/libff).  Values are kept fully reduced in [0, p), so every
// element has the same 32-byte representation as the reference's mont_repr (4 x u64 little-endian == 8 x u32).
//
// Multiplication is a CIOS Montgomery product whose partial products are issued as mad.lo.cc / madc.hi.cc pairs on
// even/odd-aligned accumulators; ptxas fuses each pair into one IMAD.WIDE.U32.X (carry in a predicate), so one
// product costs 128 wide multiply-adds + 8 IMAD.LO for the per-row quotient digit — the fma-heavy (IMAD) pipe is the
// roofline for everything built on this file.
#pragma once
#include <cstdint>

namespace b200 {

#define B200_DEV __device__ __forceinline__

struct FrParams {
    // r = 21888242871839275222246405745257275088548364400416034343698204186575808495617 (mcl_bn128_init.cpp:41)
    static __host__ __device__ constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t one(int i) {   // R mod r
        constexpr uint32_t m[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u, 0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t r2(int i) {    // R^2 mod r
        constexpr uint32_t m[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u, 0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t r3(int i) {    // R^3 mod r
        constexpr uint32_t m[8] = {0xb4bf0040u, 0x5e94d8e1u, 0x1cfbb6b8u, 0x2a489cbeu, 0xa19fcfedu, 0x893cc664u, 0x7fcc657cu, 0x0cf8594bu};
        return m[i];
    }
    static constexpr uint32_t INV = 0xefffffffu;   // -r^{-1} mod 2^32
};

struct FqParams {
    // q = 21888242871839275222246405745257275088696311157297823662689037894645226208583 (mcl_bn128_init.cpp:67)
    static __host__ __device__ constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t one(int i) {
        constexpr uint32_t m[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t r2(int i) {
        constexpr uint32_t m[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u, 0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return m[i];
    }
    static __host__ __device__ constexpr uint32_t r3(int i) {    // R^3 mod q
        constexpr uint32_t m[8] = {0xda1530dfu, 0xb1cd6dafu, 0xa7283db6u, 0x62f210e6u, 0x0ada0afbu, 0xef7f0b0cu, 0x2d592544u, 0x20fd6e90u};
        return m[i];
    }
    static constexpr uint32_t INV = 0xe4866389u;
};

// ---- carry-chain building blocks (each chain is ONE asm statement so the CC flag never crosses a statement) ----

// acc(8 limbs) += {a0,a1,a2,a3} (x) b laid out as four 64-bit products at limb offsets 0,2,4,6; carry-out added to `top`.
B200_DEV void chain_mad_carry(uint32_t* acc, uint32_t& top, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
    asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7]), "+r"(top)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
// same without a carry-out (caller guarantees none)
B200_DEV void chain_mad(uint32_t* acc, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
        "madc.hi.u32 %7, %11, %12, %7;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(acc[6]), "+r"(acc[7])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}
// x0 += y[1];  y = (y >> 64 bits) + {a0..a3} (x) b + carry(x0)      (the one-limb shift of the CIOS row, fused)
B200_DEV void chain_shift_mad(uint32_t* y, uint32_t& x0, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
    asm("add.cc.u32 %8, %8, %1;\n\t"
        "madc.lo.cc.u32 %0, %9, %13, %2;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %3;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %4;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %5;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %6;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %7;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, 0;\n\t"
        "madc.hi.u32 %7, %12, %13, 0;"
        : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7]), "+r"(x0)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
}

// ---- shorter chains and 16-limb passes for the dedicated squaring (Fp::sqr) ----
// acc(2n limbs) += {a0..} (x) b at limb offsets 0, 2, ..; the carry-out is added to `top` (a limb above everything the array
// holds so far, so it cannot overflow: rows are processed bottom-up)
B200_DEV void chain1_carry(uint32_t* acc, uint32_t& top, uint32_t a0, uint32_t b) {
    asm("mad.lo.cc.u32 %0, %3, %4, %0;\n\t"
        "madc.hi.cc.u32 %1, %3, %4, %1;\n\t"
        "addc.u32 %2, %2, 0;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(top)
        : "r"(a0), "r"(b));
}
B200_DEV void chain2_carry(uint32_t* acc, uint32_t& top, uint32_t a0, uint32_t a1, uint32_t b) {
    asm("mad.lo.cc.u32 %0, %5, %7, %0;\n\t"
        "madc.hi.cc.u32 %1, %5, %7, %1;\n\t"
        "madc.lo.cc.u32 %2, %6, %7, %2;\n\t"
        "madc.hi.cc.u32 %3, %6, %7, %3;\n\t"
        "addc.u32 %4, %4, 0;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(top)
        : "r"(a0), "r"(a1), "r"(b));
}
B200_DEV void chain3_carry(uint32_t* acc, uint32_t& top, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t b) {
    asm("mad.lo.cc.u32 %0, %7, %10, %0;\n\t"
        "madc.hi.cc.u32 %1, %7, %10, %1;\n\t"
        "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
        "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
        "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
        "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
        "addc.u32 %6, %6, 0;"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]), "+r"(top)
        : "r"(a0), "r"(a1), "r"(a2), "r"(b));
}
// r(8 limbs) = x + y + cin, carry-out in cout (cin, cout in {0, 1}; the CC flag cannot cross asm statements)
B200_DEV void add8_carry(uint32_t* r, const uint32_t* x, const uint32_t* y, uint32_t cin, uint32_t& cout) {
    uint32_t scratch;
    asm("add.cc.u32 %9, %26, 0xffffffff;\n\t"
        "addc.cc.u32 %0, %10, %18;\n\t"
        "addc.cc.u32 %1, %11, %19;\n\t"
        "addc.cc.u32 %2, %12, %20;\n\t"
        "addc.cc.u32 %3, %13, %21;\n\t"
        "addc.cc.u32 %4, %14, %22;\n\t"
        "addc.cc.u32 %5, %15, %23;\n\t"
        "addc.cc.u32 %6, %16, %24;\n\t"
        "addc.cc.u32 %7, %17, %25;\n\t"
        "addc.u32 %8, 0, 0;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(cout), "=r"(scratch)
        : "r"(x[0]), "r"(x[1]), "r"(x[2]), "r"(x[3]), "r"(x[4]), "r"(x[5]), "r"(x[6]), "r"(x[7]),
          "r"(y[0]), "r"(y[1]), "r"(y[2]), "r"(y[3]), "r"(y[4]), "r"(y[5]), "r"(y[6]), "r"(y[7]), "r"(cin));
}
// t(8 limbs) += {a0^2, a1^2, a2^2, a3^2} at limb offsets 0, 2, 4, 6 (+ cin), carry-out in cout
B200_DEV void diag4_carry(uint32_t* t, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t cin, uint32_t& cout) {
    uint32_t scratch;
    asm("add.cc.u32 %9, %14, 0xffffffff;\n\t"
        "madc.lo.cc.u32 %0, %10, %10, %0;\n\t"
        "madc.hi.cc.u32 %1, %10, %10, %1;\n\t"
        "madc.lo.cc.u32 %2, %11, %11, %2;\n\t"
        "madc.hi.cc.u32 %3, %11, %11, %3;\n\t"
        "madc.lo.cc.u32 %4, %12, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %12, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %13, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %13, %13, %7;\n\t"
        "addc.u32 %8, 0, 0;"
        : "+r"(t[0]), "+r"(t[1]), "+r"(t[2]), "+r"(t[3]), "+r"(t[4]), "+r"(t[5]), "+r"(t[6]), "+r"(t[7]), "=r"(cout), "=r"(scratch)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(cin));
}
// the one-limb shift of a CIOS row with no product to add (the REDC rows of the squaring):
// x0 += y[1];  y = (y >> 64 bits) + carry(x0)
B200_DEV void chain_shift(uint32_t* y, uint32_t& x0) {
    asm("add.cc.u32 %8, %8, %1;\n\t"
        "addc.cc.u32 %0, %2, 0;\n\t"
        "addc.cc.u32 %1, %3, 0;\n\t"
        "addc.cc.u32 %2, %4, 0;\n\t"
        "addc.cc.u32 %3, %5, 0;\n\t"
        "addc.cc.u32 %4, %6, 0;\n\t"
        "addc.cc.u32 %5, %7, 0;\n\t"
        "addc.u32 %6, 0, 0;\n\t"
        "mov.u32 %7, 0;"
        : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7]), "+r"(x0));
}

template <typename P>
struct __align__(16) Fp {
    uint32_t v[8];

    static B200_DEV Fp zero() { Fp r; _Pragma("unroll") for (int i = 0; i < 8; ++i) r.v[i] = 0; return r; }
    static B200_DEV Fp one() { Fp r; _Pragma("unroll") for (int i = 0; i < 8; ++i) r.v[i] = P::one(i); return r; }
    static B200_DEV Fp r2() { Fp r; _Pragma("unroll") for (int i = 0; i < 8; ++i) r.v[i] = P::r2(i); return r; }

    B200_DEV bool is_zero() const {
        uint32_t o = 0;
        _Pragma("unroll") for (int i = 0; i < 8; ++i) o |= v[i];
        return o == 0;
    }
    B200_DEV bool operator==(const Fp& b) const {
        uint32_t o = 0;
        _Pragma("unroll") for (int i = 0; i < 8; ++i) o |= v[i] ^ b.v[i];
        return o == 0;
    }
    B200_DEV bool operator!=(const Fp& b) const { return !(*this == b); }

    // r = (r >= p) ? r - p : r     (r < 2p)
    B200_DEV void reduce_once() {
        uint32_t t[8], borrow;
        asm("sub.cc.u32 %0, %9, %17;\n\t"
            "subc.cc.u32 %1, %10, %18;\n\t"
            "subc.cc.u32 %2, %11, %19;\n\t"
            "subc.cc.u32 %3, %12, %20;\n\t"
            "subc.cc.u32 %4, %13, %21;\n\t"
            "subc.cc.u32 %5, %14, %22;\n\t"
            "subc.cc.u32 %6, %15, %23;\n\t"
            "subc.cc.u32 %7, %16, %24;\n\t"
            "subc.u32 %8, 0, 0;"
            : "=r"(t[0]), "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7]), "=r"(borrow)
            : "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
              "r"(P::mod(0)), "r"(P::mod(1)), "r"(P::mod(2)), "r"(P::mod(3)), "r"(P::mod(4)), "r"(P::mod(5)), "r"(P::mod(6)), "r"(P::mod(7)));
        _Pragma("unroll") for (int i = 0; i < 8; ++i) v[i] = borrow ? v[i] : t[i];
    }

    friend B200_DEV Fp operator+(const Fp& a, const Fp& b) {
        Fp r;
        asm("add.cc.u32 %0, %8, %16;\n\t"
            "addc.cc.u32 %1, %9, %17;\n\t"
            "addc.cc.u32 %2, %10, %18;\n\t"
            "addc.cc.u32 %3, %11, %19;\n\t"
            "addc.cc.u32 %4, %12, %20;\n\t"
            "addc.cc.u32 %5, %13, %21;\n\t"
            "addc.cc.u32 %6, %14, %22;\n\t"
            "addc.u32 %7, %15, %23;"
            : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
            : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
              "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
        r.reduce_once();   // a + b < 2p < 2^255: no carry out of limb 7
        return r;
    }
    friend B200_DEV Fp operator-(const Fp& a, const Fp& b) {
        Fp r;
        uint32_t borrow;
        asm("sub.cc.u32 %0, %9, %17;\n\t"
            "subc.cc.u32 %1, %10, %18;\n\t"
            "subc.cc.u32 %2, %11, %19;\n\t"
            "subc.cc.u32 %3, %12, %20;\n\t"
            "subc.cc.u32 %4, %13, %21;\n\t"
            "subc.cc.u32 %5, %14, %22;\n\t"
            "subc.cc.u32 %6, %15, %23;\n\t"
            "subc.cc.u32 %7, %16, %24;\n\t"
            "subc.u32 %8, 0, 0;"
            : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]), "=r"(borrow)
            : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
              "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
        // borrow is 0 or 0xffffffff: add back p & borrow
        asm("add.cc.u32 %0, %0, %8;\n\t"
            "addc.cc.u32 %1, %1, %9;\n\t"
            "addc.cc.u32 %2, %2, %10;\n\t"
            "addc.cc.u32 %3, %3, %11;\n\t"
            "addc.cc.u32 %4, %4, %12;\n\t"
            "addc.cc.u32 %5, %5, %13;\n\t"
            "addc.cc.u32 %6, %6, %14;\n\t"
            "addc.u32 %7, %7, %15;"
            : "+r"(r.v[0]), "+r"(r.v[1]), "+r"(r.v[2]), "+r"(r.v[3]), "+r"(r.v[4]), "+r"(r.v[5]), "+r"(r.v[6]), "+r"(r.v[7])
            : "r"(P::mod(0) & borrow), "r"(P::mod(1) & borrow), "r"(P::mod(2) & borrow), "r"(P::mod(3) & borrow),
              "r"(P::mod(4) & borrow), "r"(P::mod(5) & borrow), "r"(P::mod(6) & borrow), "r"(P::mod(7) & borrow));
        return r;
    }
    B200_DEV Fp neg() const { return is_zero() ? *this : (zero() - *this); }
    // ---- operand forms for products only (NOT fully reduced; never stored, never compared) -------------------------
    // The CIOS rows of operator* and dot2 stay inside their accumulators for operands below 2p (running total
    // < 3p(2^32 + 1) < 2^288, result < p(4p/R + 1) < 1.76p for one product, one conditional subtraction), so the
    // operand preparation of the lane-pair G2 kernel can skip its own reductions.
    B200_DEV Fp neg_lazy() const {          // p - x, in (0, p]
        Fp r;
        asm("sub.cc.u32 %0, %8, %16;\n\t"
            "subc.cc.u32 %1, %9, %17;\n\t"
            "subc.cc.u32 %2, %10, %18;\n\t"
            "subc.cc.u32 %3, %11, %19;\n\t"
            "subc.cc.u32 %4, %12, %20;\n\t"
            "subc.cc.u32 %5, %13, %21;\n\t"
            "subc.cc.u32 %6, %14, %22;\n\t"
            "subc.u32 %7, %15, %23;"
            : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
            : "r"(P::mod(0)), "r"(P::mod(1)), "r"(P::mod(2)), "r"(P::mod(3)), "r"(P::mod(4)), "r"(P::mod(5)), "r"(P::mod(6)), "r"(P::mod(7)),
              "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]));
        return r;
    }
    static B200_DEV Fp add_lazy(const Fp& a, const Fp& b) {   // a + b, in [0, 2p)
        Fp r;
        asm("add.cc.u32 %0, %8, %16;\n\t"
            "addc.cc.u32 %1, %9, %17;\n\t"
            "addc.cc.u32 %2, %10, %18;\n\t"
            "addc.cc.u32 %3, %11, %19;\n\t"
            "addc.cc.u32 %4, %12, %20;\n\t"
            "addc.cc.u32 %5, %13, %21;\n\t"
            "addc.cc.u32 %6, %14, %22;\n\t"
            "addc.u32 %7, %15, %23;"
            : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
            : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
              "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
        return r;
    }
    static B200_DEV Fp sub_lazy(const Fp& a, const Fp& b) { return add_lazy(a, b.neg_lazy()); }   // a - b + p, in (0, 2p)
    B200_DEV Fp dbl() const { return *this + *this; }

    // Montgomery product a*b*R^-1 mod p, fully reduced: CIOS, one row per limb of b, on two accumulators E / O holding the
    // even- and odd-aligned partial products (so that every mad.lo/mad.hi pair is one IMAD.WIDE).
    static B200_DEV void mont_row0(uint32_t* E, uint32_t* O, const Fp& a, uint32_t b0) {
        // E = a_even (x) b0, O = a_odd (x) b0 (O's limb k has weight 2^(32(k+1)))
        _Pragma("unroll") for (int j = 0; j < 8; j += 2) {
            uint64_t te = (uint64_t)a.v[j] * b0, to = (uint64_t)a.v[j + 1] * b0;
            E[j] = (uint32_t)te; E[j + 1] = (uint32_t)(te >> 32);
            O[j] = (uint32_t)to; O[j + 1] = (uint32_t)(to >> 32);
        }
        uint32_t m = E[0] * P::INV;
        chain_mad(O, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
        chain_mad_carry(E, O[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
    }
    // X: even-aligned role this row, Y: odd-aligned role (holds last row's even array, limb 0 == 0)
    static B200_DEV void mont_row(uint32_t* X, uint32_t* Y, const Fp& a, uint32_t bi) {
        chain_shift_mad(Y, X[0], a.v[1], a.v[3], a.v[5], a.v[7], bi);
        chain_mad_carry(X, Y[7], a.v[0], a.v[2], a.v[4], a.v[6], bi);
        uint32_t m = X[0] * P::INV;
        chain_mad(Y, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
        chain_mad_carry(X, Y[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
    }
    // after row 7: X = O, Y = E; result = Y + (X >> 32)
    static B200_DEV Fp mont_final(const uint32_t* E, const uint32_t* O) {
        Fp r;
        asm("add.cc.u32 %0, %8, %16;\n\t"
            "addc.cc.u32 %1, %9, %17;\n\t"
            "addc.cc.u32 %2, %10, %18;\n\t"
            "addc.cc.u32 %3, %11, %19;\n\t"
            "addc.cc.u32 %4, %12, %20;\n\t"
            "addc.cc.u32 %5, %13, %21;\n\t"
            "addc.cc.u32 %6, %14, %22;\n\t"
            "addc.u32 %7, %15, 0;"
            : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
            : "r"(E[0]), "r"(E[1]), "r"(E[2]), "r"(E[3]), "r"(E[4]), "r"(E[5]), "r"(E[6]), "r"(E[7]),
              "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]));
        r.reduce_once();
        return r;
    }
    friend B200_DEV Fp operator*(const Fp& a, const Fp& b) {
        uint32_t E[8], O[8];
        mont_row0(E, O, a, b.v[0]);
        _Pragma("unroll") for (int i = 1; i < 8; ++i) mont_row((i & 1) ? O : E, (i & 1) ? E : O, a, b.v[i]);
        return mont_final(E, O);
    }
    // two independent products with their rows interleaved: two carry chains in flight per thread, for kernels whose
    // register footprint leaves too few warps per scheduler to hide the IMAD.WIDE issue latency (G2)
    static B200_DEV void mul2(const Fp& a, const Fp& b, const Fp& c, const Fp& d, Fp& ab, Fp& cd) {
        uint32_t E0[8], O0[8], E1[8], O1[8];
        mont_row0(E0, O0, a, b.v[0]);
        mont_row0(E1, O1, c, d.v[0]);
        _Pragma("unroll") for (int i = 1; i < 8; ++i) {
            mont_row((i & 1) ? O0 : E0, (i & 1) ? E0 : O0, a, b.v[i]);
            mont_row((i & 1) ? O1 : E1, (i & 1) ? E1 : O1, c, d.v[i]);
        }
        ab = mont_final(E0, O0);
        cd = mont_final(E1, O1);
    }
    // (a0*b0 + a1*b1) * R^-1 mod p, fully reduced, with ONE reduction: each CIOS row adds both partial products before its
    // m*p step (200 multiply-adds against 272 for two products).  Bound: every running total stays below 3p(2^32 + 1) < 2^288
    // (p < 2^254), so the odd-aligned accumulator never carries out, and the result is below p(2p/R + 1) < 2p.
    static B200_DEV Fp dot2(const Fp& a0, const Fp& b0, const Fp& a1, const Fp& b1) {
        uint32_t E[8], O[8];
        _Pragma("unroll") for (int j = 0; j < 8; j += 2) {
            uint64_t te = (uint64_t)a0.v[j] * b0.v[0], to = (uint64_t)a0.v[j + 1] * b0.v[0];
            E[j] = (uint32_t)te; E[j + 1] = (uint32_t)(te >> 32);
            O[j] = (uint32_t)to; O[j + 1] = (uint32_t)(to >> 32);
        }
        chain_mad(O, a1.v[1], a1.v[3], a1.v[5], a1.v[7], b1.v[0]);
        chain_mad_carry(E, O[7], a1.v[0], a1.v[2], a1.v[4], a1.v[6], b1.v[0]);
        {
            uint32_t m = E[0] * P::INV;
            chain_mad(O, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(E, O[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        _Pragma("unroll") for (int i = 1; i < 8; ++i) {
            uint32_t* X = (i & 1) ? O : E;
            uint32_t* Y = (i & 1) ? E : O;
            chain_shift_mad(Y, X[0], a0.v[1], a0.v[3], a0.v[5], a0.v[7], b0.v[i]);
            chain_mad_carry(X, Y[7], a0.v[0], a0.v[2], a0.v[4], a0.v[6], b0.v[i]);
            chain_mad(Y, a1.v[1], a1.v[3], a1.v[5], a1.v[7], b1.v[i]);
            chain_mad_carry(X, Y[7], a1.v[0], a1.v[2], a1.v[4], a1.v[6], b1.v[i]);
            uint32_t m = X[0] * P::INV;
            chain_mad(Y, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(X, Y[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        return mont_final(E, O);
    }
    // (a0*b0 + a1*b1 + a2*b2 + a3*b3) * R^-1 mod p, fully reduced, with ONE reduction (328 multiply-adds against 400 for two
    // dot2): the imaginary/real component of x*y - z*w in Fq2 as seen by one lane of the lane-pair G2 kernel.  Operands
    // at most p (canonical values or neg_lazy of them).  Bound: a row adds at most 4 p 2^32 + p 2^32 to a running total that
    // the previous shift left below p(5 + 2^-29), so every total stays below 5p(2^32 + 1.1) < 0.946 * 2^288 (p < 0.18904 * 2^256):
    // the odd-aligned accumulator never carries out; the result is below p(4p/R + 1) < 1.76p, one conditional subtraction.
    static B200_DEV Fp dot4(const Fp& a0, const Fp& b0, const Fp& a1, const Fp& b1, const Fp& a2, const Fp& b2, const Fp& a3, const Fp& b3) {
        uint32_t E[8], O[8];
        _Pragma("unroll") for (int j = 0; j < 8; j += 2) {
            uint64_t te = (uint64_t)a0.v[j] * b0.v[0], to = (uint64_t)a0.v[j + 1] * b0.v[0];
            E[j] = (uint32_t)te; E[j + 1] = (uint32_t)(te >> 32);
            O[j] = (uint32_t)to; O[j + 1] = (uint32_t)(to >> 32);
        }
        chain_mad(O, a1.v[1], a1.v[3], a1.v[5], a1.v[7], b1.v[0]);
        chain_mad_carry(E, O[7], a1.v[0], a1.v[2], a1.v[4], a1.v[6], b1.v[0]);
        chain_mad(O, a2.v[1], a2.v[3], a2.v[5], a2.v[7], b2.v[0]);
        chain_mad_carry(E, O[7], a2.v[0], a2.v[2], a2.v[4], a2.v[6], b2.v[0]);
        chain_mad(O, a3.v[1], a3.v[3], a3.v[5], a3.v[7], b3.v[0]);
        chain_mad_carry(E, O[7], a3.v[0], a3.v[2], a3.v[4], a3.v[6], b3.v[0]);
        {
            uint32_t m = E[0] * P::INV;
            chain_mad(O, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(E, O[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        _Pragma("unroll") for (int i = 1; i < 8; ++i) {
            uint32_t* X = (i & 1) ? O : E;
            uint32_t* Y = (i & 1) ? E : O;
            chain_shift_mad(Y, X[0], a0.v[1], a0.v[3], a0.v[5], a0.v[7], b0.v[i]);
            chain_mad_carry(X, Y[7], a0.v[0], a0.v[2], a0.v[4], a0.v[6], b0.v[i]);
            chain_mad(Y, a1.v[1], a1.v[3], a1.v[5], a1.v[7], b1.v[i]);
            chain_mad_carry(X, Y[7], a1.v[0], a1.v[2], a1.v[4], a1.v[6], b1.v[i]);
            chain_mad(Y, a2.v[1], a2.v[3], a2.v[5], a2.v[7], b2.v[i]);
            chain_mad_carry(X, Y[7], a2.v[0], a2.v[2], a2.v[4], a2.v[6], b2.v[i]);
            chain_mad(Y, a3.v[1], a3.v[3], a3.v[5], a3.v[7], b3.v[i]);
            chain_mad_carry(X, Y[7], a3.v[0], a3.v[2], a3.v[4], a3.v[6], b3.v[i]);
            uint32_t m = X[0] * P::INV;
            chain_mad(Y, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(X, Y[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        return mont_final(E, O);
    }
#ifndef B200_FP_SQR_DEDICATED
#define B200_FP_SQR_DEDICATED 0     // measured on B200 (round 2): correct (all GPU parity tests green with 1) but NOT faster - see below
#endif
    // Montgomery square, fully reduced.  Dedicated form (round 2): the 28 cross products a_i a_j (i < j) are accumulated once on
    // an even- and an odd-aligned 16-limb array (rows bottom-up, so a row's carry lands on a limb nothing has touched yet),
    // doubled, the 8 squares a_i^2 added, and the 512-bit value T reduced as (T_lo + M p)/R + T_hi with eight product-free CIOS
    // rows: 36 + 64 wide multiply-adds + 8 quotient digits = 108 against 136 for operator*.  (T_lo + M p)/R <= p and
    // T_hi < p^2/R < p/4, so one conditional subtraction finishes it.  Checked limb for limb against a big-integer model for
    // both moduli (edge values and 2*10^4 random ones) before it was written; tests: OP_SQR against the oracle's product.
    // Result: the G1 accumulate kernel has 7 % fewer IMAD.WIDE (2044 -> 1894 static) and takes 2.236 ms against 2.22 ms - the
    // extra carry passes, moves and the product-free shift rows cost what the 28 saved multiply-adds gain.  Off by default.
    B200_DEV Fp sqr() const {
#if B200_FP_SQR_DEDICATED
        uint32_t Ev[16], Od[16];
        _Pragma("unroll") for (int i = 0; i < 16; ++i) { Ev[i] = 0; Od[i] = 0; }
        const uint32_t* a = v;
        chain3_carry(&Ev[2], Ev[8], a[2], a[4], a[6], a[0]);               // row 0: a0 x (a2 a4 a6 | a1 a3 a5 a7)
        chain_mad_carry(&Od[1], Od[9], a[1], a[3], a[5], a[7], a[0]);
        chain3_carry(&Ev[4], Ev[10], a[3], a[5], a[7], a[1]);              // row 1
        chain3_carry(&Od[3], Od[9], a[2], a[4], a[6], a[1]);
        chain2_carry(&Ev[6], Ev[10], a[4], a[6], a[2]);                    // row 2
        chain3_carry(&Od[5], Od[11], a[3], a[5], a[7], a[2]);
        chain2_carry(&Ev[8], Ev[12], a[5], a[7], a[3]);                    // row 3
        chain2_carry(&Od[7], Od[11], a[4], a[6], a[3]);
        chain1_carry(&Ev[10], Ev[12], a[6], a[4]);                         // row 4
        chain2_carry(&Od[9], Od[13], a[5], a[7], a[4]);
        chain1_carry(&Ev[12], Ev[14], a[7], a[5]);                         // row 5
        chain1_carry(&Od[11], Od[13], a[6], a[5]);
        chain1_carry(&Od[13], Od[15], a[7], a[6]);                         // row 6
        uint32_t T[16], c, c2;
        add8_carry(T, Ev, Od, 0u, c);                                      // C = Ev + Od
        add8_carry(T + 8, Ev + 8, Od + 8, c, c2);
        add8_carry(T, T, T, 0u, c);                                        // 2 C
        add8_carry(T + 8, T + 8, T + 8, c, c2);
        diag4_carry(T, a[0], a[1], a[2], a[3], 0u, c);                     // + a_i^2 at limb 2 i
        diag4_carry(T + 8, a[4], a[5], a[6], a[7], c, c2);
        // REDC of T_lo: the rows of operator* with nothing left to multiply
        uint32_t E[8], O[8];
        _Pragma("unroll") for (int i = 0; i < 8; ++i) { E[i] = T[i]; O[i] = 0; }
        {
            uint32_t m = E[0] * P::INV;
            chain_mad(O, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(E, O[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        _Pragma("unroll") for (int i = 1; i < 8; ++i) {
            uint32_t* X = (i & 1) ? O : E;
            uint32_t* Y = (i & 1) ? E : O;
            chain_shift(Y, X[0]);
            uint32_t m = X[0] * P::INV;
            chain_mad(Y, P::mod(1), P::mod(3), P::mod(5), P::mod(7), m);
            chain_mad_carry(X, Y[7], P::mod(0), P::mod(2), P::mod(4), P::mod(6), m);
        }
        // after row 7: X = O, Y = E; (T_lo + M p)/R = E + (O >> 32); then + T_hi
        uint32_t lo[8], sh[8];
        _Pragma("unroll") for (int i = 0; i < 7; ++i) sh[i] = O[i + 1];
        sh[7] = 0;
        add8_carry(lo, E, sh, 0u, c);
        Fp r;
        add8_carry(r.v, lo, T + 8, 0u, c);
        r.reduce_once();
        return r;
#else
        return (*this) * (*this);
#endif
    }

    B200_DEV Fp to_mont() const { return (*this) * r2(); }
    B200_DEV Fp from_mont() const {   // as_bigint: Montgomery-multiply by 1
        Fp o = zero(); o.v[0] = 1;
        return (*this) * o;
    }
    // this^e for a 32-bit exponent (left-to-right)
    B200_DEV Fp pow_u32(uint32_t e) const {
        Fp r = one();
        for (int i = 31; i >= 0; --i) {
            r = r.sqr();
            if ((e >> i) & 1) r = r * (*this);
        }
        return r;
    }
    // Binary extended Euclid (the reference's Fp_model::inverse is the same algorithm family, fp.tcc:642-692): integer
    // adds, subs and shifts only - no multiplier - so one thread can invert while the other warps keep the IMAD pipe.
    // Input and output in Montgomery form; 0 -> 0.  (x R)^-1 = x^-1 R^-1, one Montgomery product with R^3 gives x^-1 R.
    B200_DEV Fp inv_binary() const {
        if (is_zero()) return *this;
        uint32_t u[8], w[8], b[8], c[8];
        _Pragma("unroll") for (int i = 0; i < 8; ++i) { u[i] = v[i]; w[i] = P::mod(i); b[i] = 0; c[i] = 0; }
        b[0] = 1;
        auto is_one = [](const uint32_t* x) { uint32_t o = x[0] ^ 1u; _Pragma("unroll") for (int i = 1; i < 8; ++i) o |= x[i]; return o == 0; };
        auto shr1 = [](uint32_t* x) { _Pragma("unroll") for (int i = 0; i < 7; ++i) x[i] = (x[i] >> 1) | (x[i + 1] << 31); x[7] >>= 1; };
        auto add_mod_p = [](uint32_t* x) { uint64_t cy = 0; _Pragma("unroll") for (int i = 0; i < 8; ++i) { cy += (uint64_t)x[i] + P::mod(i); x[i] = (uint32_t)cy; cy >>= 32; } };
        auto geq = [](const uint32_t* x, const uint32_t* y) {   // x >= y
            _Pragma("unroll") for (int i = 7; i >= 0; --i) { if (x[i] != y[i]) return x[i] > y[i]; }
            return true;
        };
        auto sub = [](uint32_t* x, const uint32_t* y) {          // x -= y, returns borrow
            int64_t bw = 0;
            _Pragma("unroll") for (int i = 0; i < 8; ++i) { bw += (int64_t)x[i] - (int64_t)y[i]; x[i] = (uint32_t)bw; bw >>= 32; }
            return bw != 0;
        };
        auto halve = [&](uint32_t* x) { if (x[0] & 1u) add_mod_p(x); shr1(x); };   // x/2 mod p (x + p < 2^255)
        while (!is_one(u) && !is_one(w)) {
            while (!(u[0] & 1u)) { shr1(u); halve(b); }
            while (!(w[0] & 1u)) { shr1(w); halve(c); }
            if (geq(u, w)) { sub(u, w); if (sub(b, c)) add_mod_p(b); }
            else { sub(w, u); if (sub(c, b)) add_mod_p(c); }
        }
        Fp y, r3;
        const uint32_t* res = is_one(u) ? b : c;
        _Pragma("unroll") for (int i = 0; i < 8; ++i) { y.v[i] = res[i]; r3.v[i] = P::r3(i); }
        return y * r3;
    }
    // Fermat inverse this^(p-2); 0 -> 0.  (The reference uses binary extended GCD, fp.tcc:642-692; same value.)
    B200_DEV Fp inv() const {
        Fp r = one();
        for (int i = 7; i >= 0; --i) {
            uint32_t w = P::mod(i) - (i == 0 ? 2u : 0u);   // low limb of p is > 2 for both fields
            for (int bit = 31; bit >= 0; --bit) {
                r = r.sqr();
                if ((w >> bit) & 1) r = r * (*this);
            }
        }
        return r;
    }
};

typedef Fp<FrParams> Fr;
typedef Fp<FqParams> Fq;

// Fq2 = Fq[u]/(u^2 + 1): a + b*u   (mcl Fp2T{a,b}, MCL/include/mcl/fp_tower.hpp:199-225; libff c0 + c1*u, fp2.tcc)
struct Fq2 {
    Fq a, b;
    static B200_DEV Fq2 zero() { return Fq2{Fq::zero(), Fq::zero()}; }
    static B200_DEV Fq2 one() { return Fq2{Fq::one(), Fq::zero()}; }
    B200_DEV bool is_zero() const { return a.is_zero() && b.is_zero(); }
    B200_DEV bool operator==(const Fq2& o) const { return a == o.a && b == o.b; }
    B200_DEV bool operator!=(const Fq2& o) const { return !(*this == o); }
    friend B200_DEV Fq2 operator+(const Fq2& x, const Fq2& y) { return Fq2{x.a + y.a, x.b + y.b}; }
    friend B200_DEV Fq2 operator-(const Fq2& x, const Fq2& y) { return Fq2{x.a - y.a, x.b - y.b}; }
    B200_DEV Fq2 neg() const { return Fq2{a.neg(), b.neg()}; }
    B200_DEV Fq2 dbl() const { return Fq2{a.dbl(), b.dbl()}; }
#ifndef B200_FQ2_OOL
#define B200_FQ2_OOL 0      // 1: the Fq products inside Fq2 are calls to one out-of-line copy (smaller code)
#endif
    static __device__ __noinline__ Fq mul_ool(Fq x, Fq y) { return x * y; }
    static B200_DEV Fq fqmul(const Fq& x, const Fq& y) {
#if B200_FQ2_OOL
        return mul_ool(x, y);
#else
        return x * y;
#endif
    }
#ifndef B200_FQ2_MUL2
#define B200_FQ2_MUL2 0     // 1: pairs of Fq products inside Fq2 go through Fq::mul2 (rows interleaved)
#endif
    friend B200_DEV Fq2 operator*(const Fq2& x, const Fq2& y) {   // Karatsuba, 3 Fq products
#if B200_FQ2_MUL2
        Fq aa, bb;
        Fq::mul2(x.a, y.a, x.b, y.b, aa, bb);
#else
        Fq aa = fqmul(x.a, y.a), bb = fqmul(x.b, y.b);
#endif
        Fq m = fqmul(x.a + x.b, y.a + y.b);
        return Fq2{aa - bb, m - aa - bb};
    }
    B200_DEV Fq2 sqr() const {   // complex squaring, 2 Fq products
#if B200_FQ2_MUL2
        Fq ab, d;
        Fq::mul2(a, b, a + b, a - b, ab, d);
        return Fq2{d, ab.dbl()};
#else
        Fq ab = fqmul(a, b);
        return Fq2{fqmul(a + b, a - b), ab.dbl()};
#endif
    }
    B200_DEV Fq2 inv() const {
        Fq n = (a.sqr() + b.sqr()).inv();
        return Fq2{a * n, (b * n).neg()};
    }
    B200_DEV Fq2 inv_binary() const {   // conjugate over the norm, the norm inverted without the multiplier
        Fq n = (a.sqr() + b.sqr()).inv_binary();
        return Fq2{a * n, (b * n).neg()};
    }
};

// x*y - z*w, the Y coordinate of every XYZZ addition (R (Q - X3) - Y1 PPP).  In the prime field it is one two-product row
// set under ONE reduction (Fp::dot2 with z entering as p - z: 200 multiply-adds against 272 and a modular subtraction).
#ifndef B200_FUSED_Y
#define B200_FUSED_Y 1
#endif
template <typename P>
B200_DEV Fp<P> diff_of_products(const Fp<P>& x, const Fp<P>& y, const Fp<P>& z, const Fp<P>& w) {
#if B200_FUSED_Y
    return Fp<P>::dot2(x, y, z.neg_lazy(), w);
#else
    return x * y - z * w;
#endif
}
B200_DEV Fq2 diff_of_products(const Fq2& x, const Fq2& y, const Fq2& z, const Fq2& w) { return x * y - z * w; }

}  // namespace b200
