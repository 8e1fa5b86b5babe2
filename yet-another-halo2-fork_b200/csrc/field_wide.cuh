// field_wide.cuh -- split multiplier: 8x8 -> 16-limb product, dedicated squaring, and a 16 -> 8 limb Montgomery
// reduction, so that (a) a square costs 36 + 64 IMAD.WIDE instead of 128 and (b) a sum / difference of products is
// reduced ONCE (x*y - z*w costs 64 + 64 + 64 instead of 256).  IMAD.WIDE issues at half rate on sm_100a
// (profiles/r1_imad_micro.txt), so the IMAD.WIDE count is the cost model; the extra carry handling runs on the
// otherwise idle ALU pipe.  Same even/odd accumulator idea as field.cuh: products landing on even limb positions go
// to one array, odd positions to another (offset by one limb), each as mad.lo.cc/madc.hi.cc pairs on a carry chain.
#pragma once
#include "field.cuh"

namespace h2b {

struct Wide { uint32_t l[16]; };   // a 512-bit integer, little-endian limbs

namespace detail {
// acc[0..7] (+ carry out in CC) += a[OFF], a[OFF+2], a[OFF+4], a[OFF+6] times bi; acc[8] receives the carry.
template <int OFF>
H2B_HD void wide_row_chain(uint32_t* acc, const uint32_t a[8], uint32_t bi, bool top_fresh) {
    cmad4<OFF>(acc, a, bi);
    // acc[8] is either untouched so far (fresh: set it to the carry) or already holds a carry limb (add to it)
    (void)top_fresh;
    acc[8] = ptx::addc(acc[8], 0);
}
}  // namespace detail

// 16-limb product of two 8-limb integers (64 IMAD.WIDE).
H2B_HD Wide mul_wide(const uint32_t a[8], const uint32_t b[8]) {
    // E[k] = position k, O[k] = position k + 1
    uint32_t E[18], O[18];
#pragma unroll
    for (int k = 0; k < 18; k++) { E[k] = 0; O[k] = 0; }
    // row 0: fresh products
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        E[j] = ptx::mul_lo(a[j], b[0]);
        E[j + 1] = ptx::mul_hi(a[j], b[0]);
        O[j] = ptx::mul_lo(a[j + 1], b[0]);
        O[j + 1] = ptx::mul_hi(a[j + 1], b[0]);
    }
#pragma unroll
    for (int i = 1; i < 8; i++) {
        if (i & 1) {
            // a_even * b_i lands on odd positions i + j  -> O index i - 1 + j ; a_odd * b_i on even positions i + 1 + (j-1) -> E
            detail::wide_row_chain<0>(O + (i - 1), a, b[i], false);
            detail::wide_row_chain<1>(E + (i + 1), a, b[i], true);
        } else {
            detail::wide_row_chain<0>(E + i, a, b[i], false);
            detail::wide_row_chain<1>(O + i, a, b[i], true);
        }
    }
    // result = E + (O << 32)
    Wide r;
    r.l[0] = E[0];
    r.l[1] = ptx::add_cc(E[1], O[0]);
#pragma unroll
    for (int k = 2; k < 15; k++) r.l[k] = ptx::addc_cc(E[k], O[k - 1]);
    r.l[15] = ptx::addc(E[15], O[14]);
    return r;
}

// 16-limb square (36 IMAD.WIDE): off-diagonal products once, doubled, plus the diagonal.
H2B_HD Wide sqr_wide(const uint32_t a[8]) {
    uint32_t E[18], O[18];
#pragma unroll
    for (int k = 0; k < 18; k++) { E[k] = 0; O[k] = 0; }
    // off-diagonal a_i * a_j, i < j, accumulated per row i over j = i+1..7 (position i + j)
#pragma unroll
    for (int i = 0; i < 7; i++) {
        // split the row's products by parity of (i + j)
        bool started_e = false, started_o = false;
        int last_e = -1, last_o = -1;
#pragma unroll
        for (int j = i + 1; j < 8; j++) {
            int pos = i + j;
            if ((pos & 1) == 0) {
                if (!started_e) { E[pos] = ptx::mad_lo_cc(a[i], a[j], E[pos]); started_e = true; }
                else E[pos] = ptx::madc_lo_cc(a[i], a[j], E[pos]);
                E[pos + 1] = ptx::madc_hi_cc(a[i], a[j], E[pos + 1]);
                last_e = pos + 2;
                // the even chain of this row is contiguous in pos (step 2), but interleaved in program order with the
                // odd chain: finish each chain separately below instead
            }
        }
        if (started_e) E[last_e] = ptx::addc(E[last_e], 0);
#pragma unroll
        for (int j = i + 1; j < 8; j++) {
            int pos = i + j;
            if ((pos & 1) == 1) {
                if (!started_o) { O[pos - 1] = ptx::mad_lo_cc(a[i], a[j], O[pos - 1]); started_o = true; }
                else O[pos - 1] = ptx::madc_lo_cc(a[i], a[j], O[pos - 1]);
                O[pos] = ptx::madc_hi_cc(a[i], a[j], O[pos]);
                last_o = pos + 1;
            }
        }
        if (started_o) O[last_o] = ptx::addc(O[last_o], 0);
    }
    // merge: T = E + (O << 32)   (off-diagonal sum < 2^511)
    uint32_t t[16];
    t[0] = E[0];
    t[1] = ptx::add_cc(E[1], O[0]);
#pragma unroll
    for (int k = 2; k < 15; k++) t[k] = ptx::addc_cc(E[k], O[k - 1]);
    t[15] = ptx::addc(E[15], O[14]);
    // double it
    uint32_t d[16];
    d[0] = ptx::add_cc(t[0], t[0]);
#pragma unroll
    for (int k = 1; k < 15; k++) d[k] = ptx::addc_cc(t[k], t[k]);
    d[15] = ptx::addc(t[15], t[15]);
    // add the diagonal a_i^2 at position 2i (aligned pairs: one carry chain)
    Wide r;
    r.l[0] = ptx::mad_lo_cc(a[0], a[0], d[0]);
    r.l[1] = ptx::madc_hi_cc(a[0], a[0], d[1]);
#pragma unroll
    for (int i = 1; i < 8; i++) {
        r.l[2 * i] = ptx::madc_lo_cc(a[i], a[i], d[2 * i]);
        if (i < 7) r.l[2 * i + 1] = ptx::madc_hi_cc(a[i], a[i], d[2 * i + 1]);
        else r.l[2 * i + 1] = ptx::madc_hi(a[i], a[i], d[2 * i + 1]);
    }
    return r;
}

namespace detail {
// One reduction row.  `e` holds limb positions 0..7 of the running value, `o` positions 1..8.  For rows after the
// first, `o` is the array the previous row used as `e` (its limb 0 is now zero): the value is shifted down one limb
// and the next input limb `next` enters at position 7.
template <class P, bool FIRST>
H2B_HD void redc_row(uint32_t e[8], uint32_t o[8], uint32_t next) {
    if (!FIRST) {
        e[0] = ptx::add_cc(e[0], o[1]);
        o[0] = ptx::addc_cc(o[2], 0);
        o[1] = ptx::addc_cc(o[3], 0);
        o[2] = ptx::addc_cc(o[4], 0);
        o[3] = ptx::addc_cc(o[5], 0);
        o[4] = ptx::addc_cc(o[6], 0);
        o[5] = ptx::addc_cc(o[7], 0);
        o[6] = ptx::addc_cc(next, 0);
        o[7] = ptx::addc(0, 0);
    }
    uint32_t m = ptx::mul_lo(e[0], P::M0);
    cmad4_mod<P, 1>(o, m);
    cmad4_mod<P, 0>(e, m);
    o[7] = ptx::addc(o[7], 0);
}
}  // namespace detail

// Montgomery reduction of a 16-limb value t < p * 2^256: returns t / 2^256 mod p, fully reduced (64 IMAD.WIDE).
template <class P>
H2B_HD Fe<P> redc_wide(const Wide& t) {
    uint32_t ev[8], od[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { ev[k] = t.l[k]; od[k] = 0; }
    detail::redc_row<P, true>(ev, od, 0);
    detail::redc_row<P, false>(od, ev, t.l[8]);
    detail::redc_row<P, false>(ev, od, t.l[9]);
    detail::redc_row<P, false>(od, ev, t.l[10]);
    detail::redc_row<P, false>(ev, od, t.l[11]);
    detail::redc_row<P, false>(od, ev, t.l[12]);
    detail::redc_row<P, false>(ev, od, t.l[13]);
    detail::redc_row<P, false>(od, ev, t.l[14]);
    // last row used `od` as e (od[0] == 0) and `ev` as o: limb j = od[j+1] + ev[j]; the top input limb joins limb 7
    uint32_t r[8];
    r[0] = ptx::add_cc(ev[0], od[1]);
#pragma unroll
    for (int k = 1; k < 7; k++) r[k] = ptx::addc_cc(ev[k], od[k + 1]);
    r[7] = ptx::addc(ev[7], t.l[15]);       // result < 2p < 2^255: no carry out
    Fe<P> out;
    final_sub<P>(out.l, r);
    return out;
}

template <class P>
H2B_HD Fe<P> fe_mul_w(const Fe<P>& a, const Fe<P>& b) { return redc_wide<P>(mul_wide(a.l, b.l)); }
template <class P>
H2B_HD Fe<P> fe_sqr_w(const Fe<P>& a) {
#if !defined(__CUDA_ARCH__) && !defined(H2B_HOST_EMULATE_PTX)
    return fe_mul(a, a);            // host library code: 64-bit limb path
#else
    return redc_wide<P>(sqr_wide(a.l));
#endif
}

// a*b - c*d with ONE reduction.  The difference is made non-negative by adding p * 2^256 (as a 16-limb constant
// whose low 8 limbs are zero), which the reduction maps to + p == 0 (mod p).  Requires a,b,c,d < p.
template <class P>
H2B_HD Fe<P> fe_mul_sub_mul(const Fe<P>& a, const Fe<P>& b, const Fe<P>& c, const Fe<P>& d) {
#if !defined(__CUDA_ARCH__) && !defined(H2B_HOST_EMULATE_PTX)
    return fe_sub(fe_mul(a, b), fe_mul(c, d));
#endif
    Wide x = mul_wide(a.l, b.l), y = mul_wide(c.l, d.l);
    Wide t;
    t.l[0] = ptx::sub_cc(x.l[0], y.l[0]);
#pragma unroll
    for (int k = 1; k < 16; k++) t.l[k] = ptx::subc_cc(x.l[k], y.l[k]);
    uint32_t borrow = ptx::subc(0, 0);          // all-ones if x < y
    // add p << 256 when negative: x - y + p*2^256 in [0, p*2^256)
    t.l[8] = ptx::add_cc(t.l[8], P::p(0) & borrow);
#pragma unroll
    for (int k = 1; k < 7; k++) t.l[8 + k] = ptx::addc_cc(t.l[8 + k], P::p(k) & borrow);
    t.l[15] = ptx::addc(t.l[15], P::p(7) & borrow);
    return redc_wide<P>(t);
}

}  // namespace h2b
