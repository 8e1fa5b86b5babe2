// field.cuh -- 254-bit Montgomery field arithmetic for BN254 Fr / Fq on the sm_100a integer-multiply pipe.
//
// Elements are 8 x 32-bit little-endian limbs in Montgomery form (R = 2^256), bit-identical in memory to the
// reference's `[u64; 4]` limbs (halo2curves bn256::Fr / Fq; SURVEY.md §8 preamble), always fully reduced.
//
// The multiplier is a fused multiply/Montgomery-reduce that keeps two interleaved accumulators:
// products a[j]*b[i] with even j land on 64-bit aligned limb pairs of one accumulator, odd j on the
// other (offset by one limb), so every partial product is a single mad.lo.cc/madc.hi.cc pair on a
// carry chain (ptxas fuses each pair into one IMAD.WIDE.U32 with carry in/out).  After each row the
// value is shifted down one limb, which swaps the roles of the two accumulators.
//
// Every primitive below is written twice: inline PTX for the device, and a bit-exact C emulation of the
// same PTX instruction (explicit carry flag) for the host, so the *identical* limb-level algorithm is
// unit-tested on the CPU (tests/test_field_host.py via csrc/host_check.cpp) before it ever runs on a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define H2B_HD __host__ __device__ __forceinline__
#define H2B_D __device__ __forceinline__
#else
#define H2B_HD inline
#define H2B_D inline
#endif

namespace h2b {

// ------------------------------------------------------------------------------------------------
// PTX carry-chain primitives (+ host emulation)
// ------------------------------------------------------------------------------------------------
namespace ptx {
#if defined(__CUDA_ARCH__)
H2B_D uint32_t add_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t addc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t addc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t sub_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t subc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t subc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t mul_lo(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t mul_hi(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
H2B_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
H2B_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
H2B_D uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
H2B_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
H2B_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
#else
// Host emulation: one carry flag per thread, semantics as in the PTX ISA (CC.CF).
static thread_local uint32_t CF = 0;
inline uint32_t add_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b + CF; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc(uint32_t a, uint32_t b) { return a + b + CF; }
inline uint32_t sub_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b; CF = (uint32_t)((t >> 32) & 1); return (uint32_t)t; }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b - CF; CF = (uint32_t)((t >> 32) & 1); return (uint32_t)t; }
inline uint32_t subc(uint32_t a, uint32_t b) { return a - b - CF; }
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return (uint32_t)((uint64_t)a * b); }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)mul_lo(a, b) + c; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)mul_lo(a, b) + c + CF; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)mul_hi(a, b) + c; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)mul_hi(a, b) + c + CF; CF = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return mul_hi(a, b) + c + CF; }
#endif
}  // namespace ptx

// ------------------------------------------------------------------------------------------------
// Field parameter packs (values checked numerically in tests/test_oracle.py::test_constants)
// ------------------------------------------------------------------------------------------------
struct FrParams {   // BN254 scalar field r
    static constexpr uint32_t M0 = 0xefffffffu;   // -r^{-1} mod 2^32
    H2B_HD static uint32_t p(int i) {
        constexpr uint32_t P[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return P[i];
    }
    H2B_HD static uint32_t one(int i) {   // R mod r
        constexpr uint32_t V[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u, 0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return V[i];
    }
    H2B_HD static uint32_t r2(int i) {    // R^2 mod r
        constexpr uint32_t V[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u, 0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return V[i];
    }
};
struct FqParams {   // BN254 base field q
    static constexpr uint32_t M0 = 0xe4866389u;   // -q^{-1} mod 2^32
    H2B_HD static uint32_t p(int i) {
        constexpr uint32_t P[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return P[i];
    }
    H2B_HD static uint32_t one(int i) {
        constexpr uint32_t V[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return V[i];
    }
    H2B_HD static uint32_t r2(int i) {
        constexpr uint32_t V[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u, 0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return V[i];
    }
};

// ------------------------------------------------------------------------------------------------
// Field element
// ------------------------------------------------------------------------------------------------
template <class P>
struct alignas(16) Fe {
    uint32_t l[8];

    H2B_HD static Fe zero() { Fe r; for (int i = 0; i < 8; i++) r.l[i] = 0; return r; }
    H2B_HD static Fe one() { Fe r; for (int i = 0; i < 8; i++) r.l[i] = P::one(i); return r; }
    H2B_HD static Fe r2() { Fe r; for (int i = 0; i < 8; i++) r.l[i] = P::r2(i); return r; }
    H2B_HD bool is_zero() const { uint32_t t = 0; for (int i = 0; i < 8; i++) t |= l[i]; return t == 0; }
    H2B_HD bool operator==(const Fe& o) const { uint32_t t = 0; for (int i = 0; i < 8; i++) t |= l[i] ^ o.l[i]; return t == 0; }
    H2B_HD bool operator!=(const Fe& o) const { return !(*this == o); }
};

// r = a - p if a >= p else a        (a < 2p)
template <class P>
H2B_HD void final_sub(uint32_t r[8], const uint32_t a[8]) {
    uint32_t t[8];
    t[0] = ptx::sub_cc(a[0], P::p(0));
#pragma unroll
    for (int i = 1; i < 8; i++) t[i] = ptx::subc_cc(a[i], P::p(i));
    uint32_t borrow = ptx::subc(0, 0);      // 0 if a >= p, 0xffffffff otherwise
#pragma unroll
    for (int i = 0; i < 8; i++) r[i] = borrow ? a[i] : t[i];
}

template <class P>
H2B_HD Fe<P> fe_add(const Fe<P>& a, const Fe<P>& b) {
    uint32_t t[8];
    t[0] = ptx::add_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < 7; i++) t[i] = ptx::addc_cc(a.l[i], b.l[i]);
    t[7] = ptx::addc(a.l[7], b.l[7]);       // a + b < 2^255: no carry out
    Fe<P> r;
    final_sub<P>(r.l, t);
    return r;
}

template <class P>
H2B_HD Fe<P> fe_sub(const Fe<P>& a, const Fe<P>& b) {
    uint32_t t[8];
    t[0] = ptx::sub_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < 8; i++) t[i] = ptx::subc_cc(a.l[i], b.l[i]);
    uint32_t borrow = ptx::subc(0, 0);      // all-ones if a < b
    Fe<P> r;
    r.l[0] = ptx::add_cc(t[0], P::p(0) & borrow);
#pragma unroll
    for (int i = 1; i < 7; i++) r.l[i] = ptx::addc_cc(t[i], P::p(i) & borrow);
    r.l[7] = ptx::addc(t[7], P::p(7) & borrow);
    return r;
}

template <class P>
H2B_HD Fe<P> fe_neg(const Fe<P>& a) {
    // p - a, with 0 -> 0
    uint32_t nz = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) nz |= a.l[i];
    uint32_t mask = nz ? 0xffffffffu : 0u;
    Fe<P> r;
    r.l[0] = ptx::sub_cc(P::p(0) & mask, a.l[0]);
#pragma unroll
    for (int i = 1; i < 7; i++) r.l[i] = ptx::subc_cc(P::p(i) & mask, a.l[i]);
    r.l[7] = ptx::subc(P::p(7) & mask, a.l[7]);
    return r;
}

template <class P>
H2B_HD Fe<P> fe_dbl(const Fe<P>& a) { return fe_add(a, a); }

// ---- the fused Montgomery multiplier --------------------------------------------------------------
namespace detail {
// acc[0..7] += a[OFF], a[OFF+2], a[OFF+4], a[OFF+6] (times bi) on one carry chain; carry left in CC.
template <int OFF>
H2B_HD void cmad4(uint32_t acc[8], const uint32_t a[8], uint32_t bi) {
    acc[0] = ptx::mad_lo_cc(a[OFF + 0], bi, acc[0]);
    acc[1] = ptx::madc_hi_cc(a[OFF + 0], bi, acc[1]);
    acc[2] = ptx::madc_lo_cc(a[OFF + 2], bi, acc[2]);
    acc[3] = ptx::madc_hi_cc(a[OFF + 2], bi, acc[3]);
    acc[4] = ptx::madc_lo_cc(a[OFF + 4], bi, acc[4]);
    acc[5] = ptx::madc_hi_cc(a[OFF + 4], bi, acc[5]);
    acc[6] = ptx::madc_lo_cc(a[OFF + 6], bi, acc[6]);
    acc[7] = ptx::madc_hi_cc(a[OFF + 6], bi, acc[7]);
}
// same with the modulus limbs (compile-time immediates) as the multiplicand
template <class P, int OFF>
H2B_HD void cmad4_mod(uint32_t acc[8], uint32_t m) {
    acc[0] = ptx::mad_lo_cc(P::p(OFF + 0), m, acc[0]);
    acc[1] = ptx::madc_hi_cc(P::p(OFF + 0), m, acc[1]);
    acc[2] = ptx::madc_lo_cc(P::p(OFF + 2), m, acc[2]);
    acc[3] = ptx::madc_hi_cc(P::p(OFF + 2), m, acc[3]);
    acc[4] = ptx::madc_lo_cc(P::p(OFF + 4), m, acc[4]);
    acc[5] = ptx::madc_hi_cc(P::p(OFF + 4), m, acc[5]);
    acc[6] = ptx::madc_lo_cc(P::p(OFF + 6), m, acc[6]);
    acc[7] = ptx::madc_hi_cc(P::p(OFF + 6), m, acc[7]);
}
// One row of the multiplier.  `ev` holds limb positions 0..7, `od` positions 1..8 of the running value.
template <class P, bool FIRST>
H2B_HD void mont_row(uint32_t ev[8], uint32_t od[8], const uint32_t a[8], uint32_t bi) {
    if (FIRST) {
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            ev[j] = ptx::mul_lo(a[j], bi);
            ev[j + 1] = ptx::mul_hi(a[j], bi);
            od[j] = ptx::mul_lo(a[j + 1], bi);
            od[j + 1] = ptx::mul_hi(a[j + 1], bi);
        }
    } else {
        // The previous row left od[0] == 0 (it was that row's `ev`).  Dropping that limb shifts the value down
        // by one position: old od[1] lands on ev[0]; old od[2..7] become positions 1..6 of the new `od`.
        ev[0] = ptx::add_cc(ev[0], od[1]);
        od[0] = ptx::madc_lo_cc(a[1], bi, od[2]);
        od[1] = ptx::madc_hi_cc(a[1], bi, od[3]);
        od[2] = ptx::madc_lo_cc(a[3], bi, od[4]);
        od[3] = ptx::madc_hi_cc(a[3], bi, od[5]);
        od[4] = ptx::madc_lo_cc(a[5], bi, od[6]);
        od[5] = ptx::madc_hi_cc(a[5], bi, od[7]);
        od[6] = ptx::madc_lo_cc(a[7], bi, 0);
        od[7] = ptx::madc_hi(a[7], bi, 0);
        cmad4<0>(ev, a, bi);
        od[7] = ptx::addc(od[7], 0);
    }
    uint32_t m = ptx::mul_lo(ev[0], P::M0);
    cmad4_mod<P, 1>(od, m);               // cannot carry out: the running value stays below 2^288
    cmad4_mod<P, 0>(ev, m);
    od[7] = ptx::addc(od[7], 0);
}
}  // namespace detail

template <class P>
H2B_HD Fe<P> fe_mul(const Fe<P>& a, const Fe<P>& b) {
#if !defined(__CUDA_ARCH__) && !defined(H2B_HOST_EMULATE_PTX)
    // Host-side library code (table setup, window combine, normalisation): 4 x 64-bit CIOS.
    typedef unsigned __int128 u128;
    uint64_t A[4], B[4], M[4], t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        A[i] = (uint64_t)a.l[2 * i] | ((uint64_t)a.l[2 * i + 1] << 32);
        B[i] = (uint64_t)b.l[2 * i] | ((uint64_t)b.l[2 * i + 1] << 32);
        M[i] = (uint64_t)P::p(2 * i) | ((uint64_t)P::p(2 * i + 1) << 32);
    }
    uint64_t inv = 1;                                  // -p^{-1} mod 2^64 by Newton iteration from M0
    for (int i = 0; i < 6; i++) inv *= 2 - M[0] * inv;
    inv = (uint64_t)0 - inv;
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) { c += (u128)A[j] * B[i] + t[j]; t[j] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[4] = (uint64_t)c; t[5] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * inv;
        c = (u128)m * M[0] + t[0]; c >>= 64;
        for (int j = 1; j < 4; j++) { c += (u128)m * M[j] + t[j]; t[j - 1] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[3] = (uint64_t)c; t[4] = t[5] + (uint64_t)(c >> 64);
    }
    bool ge = t[4] != 0;
    if (!ge) { ge = true; for (int i = 3; i >= 0; i--) { if (t[i] != M[i]) { ge = t[i] > M[i]; break; } } }
    if (ge) { u128 br = 0; for (int i = 0; i < 4; i++) { u128 d = (u128)t[i] - M[i] - (uint64_t)br; t[i] = (uint64_t)d; br = (d >> 64) & 1; } }
    Fe<P> r;
    for (int i = 0; i < 4; i++) { r.l[2 * i] = (uint32_t)t[i]; r.l[2 * i + 1] = (uint32_t)(t[i] >> 32); }
    return r;
#else
    uint32_t even[8], odd[8];
    detail::mont_row<P, true>(even, odd, a.l, b.l[0]);
    detail::mont_row<P, false>(odd, even, a.l, b.l[1]);
    detail::mont_row<P, false>(even, odd, a.l, b.l[2]);
    detail::mont_row<P, false>(odd, even, a.l, b.l[3]);
    detail::mont_row<P, false>(even, odd, a.l, b.l[4]);
    detail::mont_row<P, false>(odd, even, a.l, b.l[5]);
    detail::mont_row<P, false>(even, odd, a.l, b.l[6]);
    detail::mont_row<P, false>(odd, even, a.l, b.l[7]);
    // last row used `odd` as ev (odd[0] == 0) and `even` as od: result limb j = odd[j+1] + even[j]
    uint32_t t[8];
    t[0] = ptx::add_cc(even[0], odd[1]);
#pragma unroll
    for (int i = 1; i < 7; i++) t[i] = ptx::addc_cc(even[i], odd[i + 1]);
    t[7] = ptx::addc(even[7], 0);
    Fe<P> r;
    final_sub<P>(r.l, t);
    return r;
#endif
}

template <class P>
H2B_HD Fe<P> fe_sqr(const Fe<P>& a) { return fe_mul(a, a); }

// Out-of-line multiplier for kernels whose fully inlined body would not fit the instruction cache (the NTT
// levels hold ~30 multiplications each).  Operands and result travel in registers (no stack); the ~20 extra
// MOVs per call go to the ALU pipe, which is idle next to the saturated IMAD.WIDE pipe.
template <class P>
struct FePair { Fe<P> a, b; };
#if defined(__CUDA_ARCH__)
template <class P>
__device__ __noinline__ Fe<P> fe_mul_call(Fe<P> a, Fe<P> b) { return fe_mul(a, b); }
// two independent products per call: ptxas interleaves the four carry chains (2x the ILP of one product)
template <class P>
__device__ __noinline__ FePair<P> fe_mul2_call(Fe<P> a, Fe<P> b, Fe<P> c, Fe<P> d) {
    FePair<P> r;
    r.a = fe_mul(a, b);
    r.b = fe_mul(c, d);
    return r;
}
#else
template <class P>
inline Fe<P> fe_mul_call(const Fe<P>& a, const Fe<P>& b) { return fe_mul(a, b); }
template <class P>
inline FePair<P> fe_mul2_call(const Fe<P>& a, const Fe<P>& b, const Fe<P>& c, const Fe<P>& d) {
    FePair<P> r;
    r.a = fe_mul(a, b);
    r.b = fe_mul(c, d);
    return r;
}
#endif

// Montgomery -> canonical integer (halo2curves to_repr): multiply by 1.
template <class P>
H2B_HD Fe<P> fe_from_mont(const Fe<P>& a) {
    Fe<P> o = Fe<P>::zero();
    o.l[0] = 1;
    return fe_mul(a, o);
}
template <class P>
H2B_HD Fe<P> fe_to_mont(const Fe<P>& a) { return fe_mul(a, Fe<P>::r2()); }

template <class P>
H2B_HD Fe<P> fe_pow(const Fe<P>& a, const uint32_t e[8]) {
    Fe<P> acc = Fe<P>::one();
    for (int i = 255; i >= 0; i--) {
        acc = fe_sqr(acc);
        if ((e[i >> 5] >> (i & 31)) & 1) acc = fe_mul(acc, a);
    }
    return acc;
}
template <class P>
H2B_HD Fe<P> fe_pow_u64(const Fe<P>& a, uint64_t e) {
    Fe<P> acc = Fe<P>::one();
    for (int i = 63; i >= 0; i--) {
        acc = fe_sqr(acc);
        if ((e >> i) & 1) acc = fe_mul(acc, a);
    }
    return acc;
}
template <class P>
H2B_HD Fe<P> fe_inv(const Fe<P>& a) {     // Fermat; 0 -> 0
    uint32_t e[8];
    for (int i = 0; i < 8; i++) e[i] = P::p(i);
    e[0] -= 2;                             // p is odd and p(0) >= 2
    return fe_pow(a, e);
}
template <class P>
H2B_HD Fe<P> fe_from_u64(uint64_t v) {
    Fe<P> t = Fe<P>::zero();
    t.l[0] = (uint32_t)v;
    t.l[1] = (uint32_t)(v >> 32);
    return fe_to_mont(t);
}

using Fr = Fe<FrParams>;
using Fq = Fe<FqParams>;

// 128-bit vectorised load / store of one element (two 16-byte halves; pointers are 16-byte aligned).
struct alignas(16) U4 { uint32_t x, y, z, w; };
template <class P>
H2B_HD Fe<P> fe_load(const Fe<P>* p) {
    const U4* q = reinterpret_cast<const U4*>(p);
    U4 a = q[0], b = q[1];
    Fe<P> r;
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <class P>
H2B_HD void fe_store(Fe<P>* p, const Fe<P>& v) {
    U4* q = reinterpret_cast<U4*>(p);
    U4 a, b;
    a.x = v.l[0]; a.y = v.l[1]; a.z = v.l[2]; a.w = v.l[3];
    b.x = v.l[4]; b.y = v.l[5]; b.z = v.l[6]; b.w = v.l[7];
    q[0] = a; q[1] = b;
}
template <class P>
H2B_HD void fe_split(const Fe<P>& v, U4& lo, U4& hi) {
    lo.x = v.l[0]; lo.y = v.l[1]; lo.z = v.l[2]; lo.w = v.l[3];
    hi.x = v.l[4]; hi.y = v.l[5]; hi.z = v.l[6]; hi.w = v.l[7];
}
template <class P>
H2B_HD Fe<P> fe_join(const U4& a, const U4& b) {
    Fe<P> r;
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}

}  // namespace h2b
