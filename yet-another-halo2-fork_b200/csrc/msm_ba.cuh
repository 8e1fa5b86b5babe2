// msm_ba.cuh -- bucket accumulation with batched AFFINE additions (the alternative to msm_accumulate_kernel's chain of XYZZ
// mixed additions; same place in the pipeline behind MsmAccel::msm, reference halo2_middleware/src/zal.rs:57-58,136-138).
//
// Cost model.  The multiplier (IMAD.WIDE) is the binding pipe.  A mixed XYZZ addition executes 6 mul + 2 sqr +
// one fused product difference = ~1250 multiplier-pipe instructions.  An affine addition P1 + P2 needs
//     lambda = (y2 - y1) / (x2 - x1),  x3 = lambda^2 - x1 - x2,  y3 = lambda (x1 - x3) - y1            (2 mul + 1 sqr + 1 inversion)
// and Montgomery's trick turns the K inversions of a thread's K independent additions into 3 multiplications each plus ONE
// inversion: 5 mul + 1 sqr = ~756 multiplier-pipe instructions per addition.
//
// Where the independent additions come from: a bucket's m entries are paired (entry 2k, entry 2k+1), k < K = m / 2.  One thread
// owns one bucket (population order, as in msm_accumulate_kernel, so the 32 lanes of a warp have equal K):
//     phase A   walk the pairs, d_k = x2 - x1, store the running product of the earlier d's (32 B per pair, [k][thread] layout so
//               a warp writes 1 KiB runs; it comes back out of L2 a moment later) -- 1 mul per pair;
//     invert    the 128 threads of the block multiply their totals up a binary tree in shared memory (1.5 mul per thread), thread
//               0 inverts the root with a binary extended Euclid -- on the ALU pipe, which idles next to the multiplier -- and the
//               tree hands every thread the inverse of its own total on the way down (3 mul per thread);
//     phase B   walk the pairs backwards: 1/d_k = inv * prefix_k, inv *= d_k, the affine sum (x3, y3), and ONE mixed addition of
//               that sum into the bucket's XYZZ accumulator.
// Per pair of entries: 756 + 1250 multiplier instructions instead of 2 x 1250 (-20 %).  With `rounds` = 2 the sums of round 0 are
// written out ([k][thread], 64 B) and paired again before they meet the accumulator: 3 x 756 + 1250 per four entries (-30 %).
//
// Pairs that the affine formula cannot take -- equal x (a doubling or P + (-P)), or an identity operand (x = 0 is treated as one:
// the check has to be the same in both phases, and phase A reads x only) -- contribute 1 to the running product and are added to
// the accumulator entry by entry with the complete mixed addition instead.  Results are group elements: any evaluation order
// gives the same normalised point, so parity with the oracle is bit for bit as before.
#pragma once
#include "curve.cuh"

namespace h2b {

constexpr uint32_t MSM_BA_KCAP = 64;          // pairs per bucket and launch that go through the affine path (scratch rows)
constexpr int MSM_BA_THREADS = 128;

// a^-1 (Montgomery in, Montgomery out; a != 0 -- 0 returns 0) by the binary extended Euclid: shifts, additions and
// subtractions on 8 x 32-bit limbs, no multiplier use except the two final conversions.
template <class P>
H2B_HD Fe<P> fe_inv_bgcd(const Fe<P>& a) {
    if (a.is_zero()) return a;
    uint32_t u[8], v[8];
    Fe<P> x1 = Fe<P>::zero(), x2 = Fe<P>::zero();
    x1.l[0] = 1;
#pragma unroll
    for (int i = 0; i < 8; i++) { u[i] = a.l[i]; v[i] = P::p(i); }
    // invariants: x1 * A == u, x2 * A == v (mod p), A = the integer held in a.l
    auto halve = [](uint32_t w[8], Fe<P>& x) {
#pragma unroll
        for (int i = 0; i < 7; i++) w[i] = (w[i] >> 1) | (w[i + 1] << 31);
        w[7] >>= 1;
        const uint32_t odd = 0u - (x.l[0] & 1u);
        uint32_t t[8];
        t[0] = ptx::add_cc(x.l[0], P::p(0) & odd);
#pragma unroll
        for (int i = 1; i < 7; i++) t[i] = ptx::addc_cc(x.l[i], P::p(i) & odd);
        t[7] = ptx::addc(x.l[7], P::p(7) & odd);        // x + p < 2^255: no carry out
#pragma unroll
        for (int i = 0; i < 7; i++) x.l[i] = (t[i] >> 1) | (t[i + 1] << 31);
        x.l[7] = t[7] >> 1;
    };
    for (;;) {
        while ((u[0] & 1u) == 0) halve(u, x1);
        while ((v[0] & 1u) == 0) halve(v, x2);
        uint32_t t[8];
        t[0] = ptx::sub_cc(u[0], v[0]);
#pragma unroll
        for (int i = 1; i < 8; i++) t[i] = ptx::subc_cc(u[i], v[i]);
        const uint32_t borrow = ptx::subc(0, 0);
        if (!borrow) {                                   // u >= v
            uint32_t nz = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) { u[i] = t[i]; nz |= t[i]; }
            x1 = fe_sub(x1, x2);
            if (!nz) break;                              // u == v == gcd == 1: x2 * A == 1
        } else {
            v[0] = ptx::sub_cc(v[0], u[0]);
#pragma unroll
            for (int i = 1; i < 7; i++) v[i] = ptx::subc_cc(v[i], u[i]);
            v[7] = ptx::subc(v[7], u[7]);
            x2 = fe_sub(x2, x1);
        }
    }
    // x2 = A^-1 = (a R)^-1; the Montgomery form of a^-1 is a^-1 R = A^-1 R^2
    return fe_to_mont(fe_to_mont(x2));
}

// The hot loops below hold ~15 multiplications per iteration; inlined they are ~60 KB of code, twice the instruction cache next
// to the SM (32 KB), and with four warps per scheduler at different places of the loop the fetch stalls cost more than the
// multiplier pipe saves (ncu: stall_no_inst on every line, pipe 64 % busy).  So the multiplier routines are CALLED here, as in the
// NTT passes: operands and result travel in registers, the ~20 extra MOVs per call go to the ALU pipe.
#if defined(__CUDA_ARCH__)
__device__ __noinline__ Fq ba_mul(Fq a, Fq b) { return fe_mul(a, b); }
__device__ __noinline__ Fq ba_sqr(Fq a) { return fe_sqr_w(a); }
__device__ __noinline__ Fq ba_mul_sub_mul(Fq a, Fq b, Fq c, Fq d) { return fe_mul_sub_mul(a, b, c, d); }
__device__ __forceinline__ void ba_prefetch(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#else
inline Fq ba_mul(const Fq& a, const Fq& b) { return fe_mul(a, b); }
inline Fq ba_sqr(const Fq& a) { return fe_sqr_w(a); }
inline Fq ba_mul_sub_mul(const Fq& a, const Fq& b, const Fq& c, const Fq& d) { return fe_mul_sub_mul(a, b, c, d); }
inline void ba_prefetch(const void*) {}
#endif

// xyzz_madd (curve.cuh: madd-2008-s) over the called routines; (x2, y2) is never the identity
H2B_HD void ba_madd(Xyzz& acc, const Fq& x2, const Fq& y2) {
    if (acc.is_identity()) {
        acc.x = x2; acc.y = y2; acc.zz = Fq::one(); acc.zzz = Fq::one();
        return;
    }
    const Fq p = fe_sub(ba_mul(x2, acc.zz), acc.x);
    const Fq r = fe_sub(ba_mul(y2, acc.zzz), acc.y);
    if (p.is_zero()) {
        if (r.is_zero()) {                                   // the same point: 2 * (x2, y2), mdbl-2008-s-1 with a = 0
            const Fq u = fe_dbl(y2), v = ba_sqr(u), w = ba_mul(u, v), sx = ba_mul(x2, v), xx = ba_sqr(x2);
            const Fq m3 = fe_add(fe_dbl(xx), xx);
            acc.x = fe_sub(ba_sqr(m3), fe_dbl(sx));
            acc.y = ba_mul_sub_mul(m3, fe_sub(sx, acc.x), w, y2);
            acc.zz = v; acc.zzz = w;
        } else acc = Xyzz::identity();
        return;
    }
    const Fq pp = ba_sqr(p);
    const Fq ppp = ba_mul(p, pp);
    const Fq q = ba_mul(acc.x, pp);
    const Fq x3 = fe_sub(fe_sub(ba_sqr(r), ppp), fe_dbl(q));
    acc.y = ba_mul_sub_mul(r, fe_sub(q, x3), acc.y, ppp);
    acc.x = x3;
    acc.zz = ba_mul(acc.zz, pp);
    acc.zzz = ba_mul(acc.zzz, ppp);
}

// the affine formula cannot take this pair (see the header): decided from the two x coordinates alone
H2B_HD bool msm_ba_special(const Fq& x1, const Fq& x2, const Fq& d) { return x1.is_zero() || x2.is_zero() || d.is_zero(); }

// Where a round's operands come from: the bucket's entries (indices into the bases, sign in bit 31), or the affine sums of the
// previous round held as two planes [k][thread].
struct MsmBaSource {
    const Affine* bases;
    const uint32_t* entries;     // this bucket's first entry
    const Fq* rx;                // previous round: x of sum k at rx[k * stride], y at ry[k * stride]
    const Fq* ry;
    size_t stride;
    bool from_entries;
};
H2B_HD Fq msm_ba_x(const MsmBaSource& s, uint32_t i) {
    return s.from_entries ? fe_load(&s.bases[s.entries[i] & 0x7fffffffu].x) : fe_load(s.rx + (size_t)i * s.stride);
}
H2B_HD Affine msm_ba_point(const MsmBaSource& s, uint32_t i) {       // sign applied
    Affine p;
    if (s.from_entries) {
        const uint32_t ent = s.entries[i];
        p = affine_load(s.bases + (ent & 0x7fffffffu));
        if ((ent >> 31) && !p.is_identity()) p.y = fe_neg(p.y);
    } else {
        p.x = fe_load(s.rx + (size_t)i * s.stride);
        p.y = fe_load(s.ry + (size_t)i * s.stride);
    }
    return p;
}

H2B_HD void msm_ba_prefetch(const MsmBaSource& s, uint32_t i) {
    if (s.from_entries) ba_prefetch(s.bases + (s.entries[i] & 0x7fffffffu));
    else { ba_prefetch(s.rx + (size_t)i * s.stride); ba_prefetch(s.ry + (size_t)i * s.stride); }
}

// phase A of one round: K pairs (operand 2k, operand 2k+1); prefix[k * stride] = product of the earlier usable differences.
// Returns the product of all usable differences (never zero).
H2B_HD Fq msm_ba_phase_a(const MsmBaSource& s, uint32_t K, Fq* prefix, size_t stride) {
    Fq run = Fq::one();
    if (K == 0) return run;
    Fq x1 = msm_ba_x(s, 0), x2 = msm_ba_x(s, 1);
    for (uint32_t k = 0; k < K; k++) {
        Fq x1n = x1, x2n = x2;
        if (k + 1 < K) { x1n = msm_ba_x(s, 2 * k + 2); x2n = msm_ba_x(s, 2 * k + 3); }      // in flight under the multiplication
        if (k + 3 < K) { msm_ba_prefetch(s, 2 * k + 6); msm_ba_prefetch(s, 2 * k + 7); }
        const Fq d = fe_sub(x2, x1);
        fe_store(prefix + (size_t)k * stride, run);
        if (!msm_ba_special(x1, x2, d)) run = ba_mul(run, d);
        x1 = x1n; x2 = x2n;
    }
    return run;
}

// phase B of one round, pair k (walked from K - 1 down to 0): `inv` holds 1 / (d_0 ... d_k) on entry and 1 / (d_0 ... d_(k-1))
// on return.  Returns true with the affine sum in (x3, y3); false for a pair the affine formula cannot take (inv untouched).
H2B_HD bool msm_ba_pair(const Affine& p1, const Affine& p2, const Fq& pre, Fq& inv, Fq& x3, Fq& y3) {
    const Fq d = fe_sub(p2.x, p1.x);
    if (msm_ba_special(p1.x, p2.x, d)) return false;
    const Fq inv_d = ba_mul(inv, pre);
    inv = ba_mul(inv, d);
    const Fq lam = ba_mul(fe_sub(p2.y, p1.y), inv_d);
    x3 = fe_sub(fe_sub(ba_sqr(lam), p1.x), p2.x);
    y3 = fe_sub(ba_mul(lam, fe_sub(p1.x, x3)), p1.y);
    return true;
}

// Phase B of one round for one bucket: the K pairs walked backwards, then the operands no pair took (index >= 2K, < m).
// ONE site for the mixed addition: whatever has to meet the accumulator -- an affine sum, the two halves of a pair the affine
// formula cannot take, a leftover operand -- is produced by the small state machine and added at the bottom of the loop.
// PARK: the affine sums are not added but written to (out_x, out_y)[slot * stride], slots dense from 0; returns their number.
template <bool PARK>
H2B_HD uint32_t msm_ba_phase_b(const MsmBaSource& s, uint32_t K, uint32_t m, const Fq* prefix, size_t stride, Fq inv, Xyzz& acc,
                               Fq* out_x, Fq* out_y) {
    uint32_t k = K, rest = 2 * K, pend = 0, parked = 0;
    bool has_pend = false;
    for (;;) {
        Affine p;
        if (has_pend) { p = msm_ba_point(s, pend); has_pend = false; }
        else if (k > 0) {
            k--;
            if (k > 0) {                                          // the next pair's operands and prefix: into L2 under this pair
                msm_ba_prefetch(s, 2 * k - 2); msm_ba_prefetch(s, 2 * k - 1);
                ba_prefetch(prefix + (size_t)(k - 1) * stride);
            }
            const Affine p1 = msm_ba_point(s, 2 * k), p2 = msm_ba_point(s, 2 * k + 1);
            Fq x3, y3;
            if (msm_ba_pair(p1, p2, fe_load(prefix + (size_t)k * stride), inv, x3, y3)) {
                if (PARK) {
                    fe_store(out_x + (size_t)parked * stride, x3);
                    fe_store(out_y + (size_t)parked * stride, y3);
                    parked++;
                    continue;
                }
                p.x = x3; p.y = y3;
            } else { p = p1; pend = 2 * k + 1; has_pend = true; }
        } else if (rest < m) { p = msm_ba_point(s, rest); rest++; }
        else break;
        if (!p.is_identity()) ba_madd(acc, p.x, p.y);
    }
    return parked;
}

#if defined(__CUDACC__) && defined(H2B_BA_KERNELS)
// Every thread of the block brings the product `run` of its differences and leaves with its inverse (see the header).
// tree: 2 * MSM_BA_THREADS elements of shared memory.
__device__ __forceinline__ Fq msm_ba_block_invert(Fq* tree, const Fq& run) {
    const uint32_t t = threadIdx.x;
    constexpr uint32_t T = MSM_BA_THREADS;
    fe_store(tree + T + t, run);                                   // leaves at [T, 2T), node i has children 2i, 2i + 1
    __syncthreads();
#pragma unroll 1
    for (uint32_t s = T / 2; s >= 1; s >>= 1) {
        if (t < s) fe_store(tree + s + t, ba_mul(fe_load(tree + 2 * (s + t)), fe_load(tree + 2 * (s + t) + 1)));
        __syncthreads();
    }
    if (t == 0) fe_store(tree + 1, fe_inv_bgcd(fe_load(tree + 1)));
    __syncthreads();
#pragma unroll 1
    for (uint32_t s = 1; s < T; s <<= 1) {                         // nodes [s, 2s) hold inverses, their children still products
        Fq l, r, iv;
        if (t < s) { l = fe_load(tree + 2 * (s + t)); r = fe_load(tree + 2 * (s + t) + 1); iv = fe_load(tree + s + t); }
        __syncthreads();
        if (t < s) { fe_store(tree + 2 * (s + t), ba_mul(iv, r)); fe_store(tree + 2 * (s + t) + 1, ba_mul(iv, l)); }
        __syncthreads();
    }
    return fe_load(tree + T + t);
}

// One launch = msm_accumulate_kernel's job (every bucket of the sort set summed into the shared bucket array), `ROUNDS` (1 or 2)
// rounds of batched affine additions in front of the mixed additions.
//   prefix : [MSM_BA_KCAP][stride] Fq            running products of phase A
//   rx, ry : [MSM_BA_KCAP][stride] Fq each       affine sums of round 0 (ROUNDS == 2 only)
template <int ROUNDS, int MINB>
__global__ void __launch_bounds__(MSM_BA_THREADS, MINB)
msm_accumulate_ba_kernel(const Affine* __restrict__ bases, const uint32_t* __restrict__ entries,
                         const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends,
                         const uint32_t* __restrict__ order, uint32_t nkeys, uint32_t long_threshold,
                         Xyzz* __restrict__ buckets, uint32_t add_to, Fq* prefix, Fq* rx, Fq* ry, uint32_t stride) {
    __shared__ Fq tree[2 * MSM_BA_THREADS];
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t key = 0, e = 0, end = 0;
    bool is_long = true;                                           // (threads past the last bucket: nothing to store either)
    if (tid < nkeys) {
        key = order[tid];
        e = offsets[key]; end = ends[key];
        is_long = end - e > long_threshold;                        // left to msm_accumulate_long_kernel
        if (is_long) end = e;
    }
    const uint32_t m = end - e;
    prefix += tid; rx += tid; ry += tid;

    // ---- round 0: pairs of entries
    MsmBaSource s0;
    s0.bases = bases; s0.entries = entries + e; s0.rx = nullptr; s0.ry = nullptr; s0.stride = stride; s0.from_entries = true;
    uint32_t K0 = m >> 1;
    if (K0 > MSM_BA_KCAP) K0 = MSM_BA_KCAP;
    Fq inv = msm_ba_block_invert(tree, msm_ba_phase_a(s0, K0, prefix, stride));

    Xyzz acc = Xyzz::identity();
    if (m != 0 && add_to) acc = xyzz_load(buckets + key);
    uint32_t K1 = msm_ba_phase_b<ROUNDS == 2>(s0, K0, m, prefix, stride, inv, acc, rx, ry);
    if (ROUNDS == 2) {
        // ---- round 1: pairs of round-0 sums (this thread's own stores: visible to it without a fence)
        MsmBaSource s1;
        s1.bases = nullptr; s1.entries = nullptr; s1.rx = rx; s1.ry = ry; s1.stride = stride; s1.from_entries = false;
        const uint32_t Kp = K1 >> 1;
        inv = msm_ba_block_invert(tree, msm_ba_phase_a(s1, Kp, prefix, stride));
        msm_ba_phase_b<false>(s1, Kp, K1, prefix, stride, inv, acc, nullptr, nullptr);
    }
    // as msm_accumulate_kernel: oversized buckets belong to the long path, a later chunk leaves untouched buckets alone
    if (!is_long && (m != 0 || !add_to)) xyzz_store(buckets + key, acc);
}
#endif  // __CUDACC__ && H2B_BA_KERNELS

}  // namespace h2b
