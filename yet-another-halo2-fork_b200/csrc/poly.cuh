// poly.cuh -- the O(n) polynomial helpers that sit between the accelerated MSM / NTT steps of create_proof
// (SURVEY.md 8f row N3): serial CPU loops in the reference, scan-shaped on the GPU.
//
//   grand product     z[0] = init, z[i] = z[i-1] * v[i-1]   permutation/prover.rs:160-166, lookup/prover.rs:254-259,
//                                                           shuffle/prover.rs:161-166 (the Z polynomials)
//   batch inversion   a[i] <- 1 / a[i], zeros untouched     ff::BatchInvert, called at permutation/prover.rs:124,
//                                                           lookup/prover.rs:225, shuffle/prover.rs:150
//   eval_polynomial   sum a_i x^i                            arithmetic.rs:56-82
//   kate_division     a(X) / (X - b) without remainder      arithmetic.rs:99-120
//
// All four are prefix computations over an associative operator, done in three small phases over tiles of
// POLY_T threads x POLY_K consecutive elements: (1) per-tile summary, (2) one block scans the tile summaries,
// (3) per-tile scan seeded with the tile's carry-in.  Every element crosses HBM twice each way at most (64-96 B per
// element); the field multiplications (5-7 per element) are what the time goes to.
//   * product scan: operator = multiplication.
//   * Horner scan (kate_division / eval_polynomial): reading the coefficients from the top, s_j = b * s_{j-1} + r_j.
//     A run of `len` elements acts on its carry-in as  x -> b^len * x + S;  runs of equal length compose with the
//     fixed multipliers b^len, b^(2 len), b^(4 len), ... (a Hillis-Steele scan with one multiplication per step).
//   * batch inversion: Montgomery's trick on two levels (tile products, then inside each tile) with ONE Fermat inversion per call.
#pragma once
#include "field.cuh"
#include "field_wide.cuh"

namespace h2b {

constexpr int POLY_T = 256;                     // threads per tile
constexpr int POLY_K = 8;                       // consecutive elements per thread
constexpr int POLY_TILE = POLY_T * POLY_K;      // 2048 elements (64 KiB) per tile
constexpr int POLY_LOG_T = 8;

// a^(p-2) with 4-bit fixed windows and the dedicated squaring (about 256 squarings + 78 multiplications); 0 -> 0
template <class P>
H2B_HD Fe<P> fe_inv_window(const Fe<P>& a) {
    Fe<P> tab[16];
    tab[0] = Fe<P>::one();
    tab[1] = a;
    for (int i = 2; i < 16; i++) tab[i] = fe_mul(tab[i - 1], a);
    uint32_t e[8];
    for (int i = 0; i < 8; i++) e[i] = P::p(i);
    e[0] -= 2;
    Fe<P> acc = Fe<P>::one();
    for (int nib = 63; nib >= 0; nib--) {
        if (nib != 63) { acc = fe_sqr_w(acc); acc = fe_sqr_w(acc); acc = fe_sqr_w(acc); acc = fe_sqr_w(acc); }
        uint32_t d = (e[nib >> 3] >> ((nib & 7) * 4)) & 15u;
        if (d) acc = fe_mul(acc, tab[d]);
    }
    return acc;
}

// one (polynomial, point) pair of a batched evaluation (poly_horner_batch_* kernels)
struct alignas(16) HornerJob {
    const Fr* a;
    uint64_t reserved;
    Fr b, bk, bt;          // the point, b^POLY_K, b^POLY_TILE
};

#if defined(__CUDACC__) && !defined(H2B_NO_KERNELS)

struct PolyShared { Fr s[2][POLY_T]; };

// inclusive scan over the block under multiplication; returns prod_{u <= t} v_u.  `cur` tells which half holds the result.
H2B_D Fr block_scan_mul(Fr v, PolyShared& sh, uint32_t t, int& cur) {
    cur = 0;
    fe_store(&sh.s[0][t], v);
    __syncthreads();
#pragma unroll 1
    for (uint32_t d = 1; d < (uint32_t)POLY_T; d <<= 1) {
        Fr x = fe_load(&sh.s[cur][t]);
        if (t >= d) x = fe_mul(fe_load(&sh.s[cur][t - d]), x);
        fe_store(&sh.s[cur ^ 1][t], x);
        cur ^= 1;
        __syncthreads();
    }
    return fe_load(&sh.s[cur][t]);
}
// inclusive SUFFIX scan: returns prod_{u >= t} v_u
H2B_D Fr block_suffix_scan_mul(Fr v, PolyShared& sh, uint32_t t, int& cur) {
    cur = 0;
    __syncthreads();
    fe_store(&sh.s[0][t], v);
    __syncthreads();
#pragma unroll 1
    for (uint32_t d = 1; d < (uint32_t)POLY_T; d <<= 1) {
        Fr x = fe_load(&sh.s[cur][t]);
        if (t + d < (uint32_t)POLY_T) x = fe_mul(x, fe_load(&sh.s[cur][t + d]));
        fe_store(&sh.s[cur ^ 1][t], x);
        cur ^= 1;
        __syncthreads();
    }
    return fe_load(&sh.s[cur][t]);
}
// inclusive Horner scan over the block: S_t <- sum_{u <= t} S_u * m^(t-u), where pw[l] = m^(2^l)
H2B_D Fr block_scan_horner(Fr v, const Fr* pw, PolyShared& sh, uint32_t t, int& cur) {
    cur = 0;
    __syncthreads();
    fe_store(&sh.s[0][t], v);
    __syncthreads();
#pragma unroll
    for (int l = 0; l < POLY_LOG_T; l++) {
        const uint32_t d = 1u << l;
        Fr x = fe_load(&sh.s[cur][t]);
        if (t >= d) x = fe_add(x, fe_mul(fe_load(&sh.s[cur][t - d]), pw[l]));
        fe_store(&sh.s[cur ^ 1][t], x);
        cur ^= 1;
        __syncthreads();
    }
    return fe_load(&sh.s[cur][t]);
}

// ---- product scan ------------------------------------------------------------------------------------------------
// phase 1: tile_tot[tile] = product of the tile's elements (indices >= n count as 1)
// zero_as_one: zeros are skipped (batch inversion leaves them untouched, ff::BatchInvert)
__global__ void __launch_bounds__(POLY_T)
poly_prod_tile_reduce_kernel(const Fr* __restrict__ v, uint64_t n, Fr* __restrict__ tile_tot, uint32_t zero_as_one) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr p = Fr::one();
#pragma unroll
    for (int k = 0; k < POLY_K; k++)
        if (base + k < n) {
            Fr e = fe_load(v + base + k);
            if (zero_as_one && e.is_zero()) e = Fr::one();
            p = k == 0 ? e : fe_mul(p, e);
        }
    int cur;
    Fr incl = block_scan_mul(p, sh, t, cur);
    if (t == POLY_T - 1) fe_store(tile_tot + blockIdx.x, incl);
}
// phase 2 (one block): tile_tot[i] <- init * prod_{j < i} tile_tot[j]
__global__ void __launch_bounds__(POLY_T)
poly_prod_scan_totals_kernel(Fr* tile_tot, uint32_t ntiles, Fr init) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    Fr carry = init;
    for (uint32_t base = 0; base < ntiles; base += POLY_T) {
        Fr v = base + t < ntiles ? fe_load(tile_tot + base + t) : Fr::one();
        int cur;
        Fr incl = block_scan_mul(v, sh, t, cur);
        Fr excl = t ? fe_load(&sh.s[cur][t - 1]) : Fr::one();
        Fr total = fe_load(&sh.s[cur][POLY_T - 1]);
        if (base + t < ntiles) fe_store(tile_tot + base + t, fe_mul(carry, excl));
        carry = fe_mul(carry, total);
        (void)incl;
        __syncthreads();
    }
}
// phase 3: z[i] = tile_prefix[tile] * prod_{tile start <= j < i} v[j]   for i < out_len   (z may alias v)
__global__ void __launch_bounds__(POLY_T)
poly_prod_tile_scan_kernel(const Fr* v, uint64_t n, const Fr* __restrict__ tile_prefix, Fr* z, uint64_t out_len) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr e[POLY_K];
    Fr p = Fr::one();
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        e[k] = base + k < n ? fe_load(v + base + k) : Fr::one();
        p = k == 0 ? e[0] : fe_mul(p, e[k]);
    }
    int cur;
    block_scan_mul(p, sh, t, cur);
    Fr c = fe_load(tile_prefix + blockIdx.x);
    if (t) c = fe_mul(c, fe_load(&sh.s[cur][t - 1]));
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        if (base + k < out_len) fe_store(z + base + k, c);
        if (k + 1 < POLY_K) c = fe_mul(c, e[k]);
    }
}

// ---- batch inversion (Montgomery's trick on two levels; zeros stay zero as in ff::BatchInvert) ---------------------------
// level 1: tile products (poly_prod_tile_reduce_kernel, zeros skipped); level 2: ONE block inverts all tile products with a
// single Fermat inversion (below); level 3: each tile inverts its elements from 1 / (tile product).
// tot[i] <- 1 / tot[i] for i < ntiles; pre = scratch of ntiles elements.  Tile products are never zero.
__global__ void __launch_bounds__(POLY_T)
poly_invert_totals_kernel(Fr* tot, Fr* pre, uint32_t ntiles) {
    __shared__ PolyShared sh;
    __shared__ Fr inv_total;
    const uint32_t t = threadIdx.x;
    const uint32_t m = (ntiles + POLY_T - 1) / POLY_T;          // consecutive tile products per thread
    const uint32_t lo = t * m, hi = lo + m < ntiles ? lo + m : ntiles;
    Fr p = Fr::one();
    for (uint32_t i = lo; i < hi; i++) {                        // pre[i] = product of this thread's entries before i
        fe_store(pre + i, p);
        p = fe_mul(p, fe_load(tot + i));
    }
    int cur;
    block_scan_mul(p, sh, t, cur);
    Fr excl_pre = t ? fe_load(&sh.s[cur][t - 1]) : Fr::one();
    if (t == 0) inv_total = fe_inv_window(fe_load(&sh.s[cur][POLY_T - 1]));
    block_suffix_scan_mul(p, sh, t, cur);                      // (its leading barrier also publishes inv_total)
    Fr excl_suf = t + 1 < (uint32_t)POLY_T ? fe_load(&sh.s[cur][t + 1]) : Fr::one();
    Fr inv = fe_mul(fe_mul(inv_total, excl_pre), excl_suf);     // 1 / (this thread's product)
    for (uint32_t i = hi; i-- > lo;) {
        Fr ti = fe_load(tot + i);
        fe_store(tot + i, fe_mul(inv, fe_load(pre + i)));
        inv = fe_mul(inv, ti);
    }
}
__global__ void __launch_bounds__(POLY_T)
poly_batch_invert_kernel(Fr* a, uint64_t n, const Fr* __restrict__ tile_inv) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr e[POLY_K], pre[POLY_K];          // pre[k] = e[0] * ... * e[k] with zeros replaced by one
    uint32_t zero_mask = 0;
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        e[k] = base + k < n ? fe_load(a + base + k) : Fr::one();
        if (e[k].is_zero()) { zero_mask |= 1u << k; e[k] = Fr::one(); }
        pre[k] = k == 0 ? e[0] : fe_mul(pre[k - 1], e[k]);
    }
    int cur;
    block_scan_mul(pre[POLY_K - 1], sh, t, cur);
    Fr excl_pre = t ? fe_load(&sh.s[cur][t - 1]) : Fr::one();
    block_suffix_scan_mul(pre[POLY_K - 1], sh, t, cur);
    Fr excl_suf = t + 1 < (uint32_t)POLY_T ? fe_load(&sh.s[cur][t + 1]) : Fr::one();
    // 1 / (this thread's product) = 1/(tile product) * (product before) * (product after)
    Fr inv = fe_mul(fe_mul(fe_load(tile_inv + blockIdx.x), excl_pre), excl_suf);
#pragma unroll
    for (int k = POLY_K - 1; k >= 0; k--) {
        Fr r = k ? fe_mul(inv, pre[k - 1]) : inv;
        if (k) inv = fe_mul(inv, e[k]);
        if (base + k < n && !((zero_mask >> k) & 1u)) fe_store(a + base + k, r);
    }
}

// ---- Horner scan: kate_division and eval_polynomial ------------------------------------------------------------------
// Reversed view, left-padded with `pad` zeros to a whole number of tiles: r_j = a[n-1-(j-pad)] for j >= pad, else 0 (leading
// zeros leave the Horner value untouched).  s_j = b * s_{j-1} + r_j, s_{-1} = 0.
// kate: q[n-2-(j-pad)] = s_j for pad <= j <= pad+n-2.  eval: s of the last element.
// pw: b^(K * 2^l) for l < 8 (thread runs inside a tile) are derived in-kernel from bK = b^K.
H2B_D Fr horner_view(const Fr* __restrict__ a, uint64_t n, uint64_t pad, uint64_t j) {
    return j >= pad ? fe_load(a + (n - 1 - (j - pad))) : Fr::zero();
}
H2B_D void horner_powers(Fr bk, Fr* pw) {
    pw[0] = bk;
#pragma unroll
    for (int l = 1; l < POLY_LOG_T; l++) pw[l] = fe_sqr_w(pw[l - 1]);
}
// phase 1: tile_val[tile] = Horner value of the tile's run with zero carry-in
__global__ void __launch_bounds__(POLY_T)
poly_horner_tile_reduce_kernel(const Fr* __restrict__ a, uint64_t n, uint64_t pad, Fr b, Fr bk, Fr* __restrict__ tile_val) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    const uint64_t j0 = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr pw[POLY_LOG_T];
    horner_powers(bk, pw);
    Fr s = Fr::zero();
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        Fr r = horner_view(a, n, pad, j0 + k);
        s = k == 0 ? r : fe_add(fe_mul(s, b), r);
    }
    int cur;
    Fr incl = block_scan_horner(s, pw, sh, t, cur);
    if (t == POLY_T - 1) fe_store(tile_val + blockIdx.x, incl);
}
// phase 2 (one block): tile_val[i] <- carry-in of tile i = sum_{j < i} tile_val[j] * bT^(i-1-j), bT = b^TILE; returns in
// *final_out the carry OUT of the last tile (the Horner value of the whole reversed view)
__global__ void __launch_bounds__(POLY_T)
poly_horner_scan_totals_kernel(Fr* tile_val, uint32_t ntiles, Fr bt, Fr* final_out) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    Fr pw[POLY_LOG_T];
    horner_powers(bt, pw);
    Fr carry = Fr::zero();
    for (uint32_t base = 0; base < ntiles; base += POLY_T) {
        Fr v = base + t < ntiles ? fe_load(tile_val + base + t) : Fr::zero();
        // seed thread 0 with the running carry: S_0 <- carry * bT + v_0
        if (t == 0) v = fe_add(fe_mul(carry, bt), v);
        int cur;
        block_scan_horner(v, pw, sh, t, cur);
        // carry-in of tile base+t = inclusive value of t-1 (or the running carry for t = 0)
        Fr cin = t ? fe_load(&sh.s[cur][t - 1]) : carry;
        if (base + t < ntiles) fe_store(tile_val + base + t, cin);
        uint32_t last = ntiles - base < (uint32_t)POLY_T ? ntiles - base - 1 : POLY_T - 1;
        Fr out = fe_load(&sh.s[cur][last]);
        __syncthreads();
        carry = out;
    }
    if (t == 0 && final_out) fe_store(final_out, carry);
}
// ---- batched evaluation: count (polynomial, point) pairs of one length in two launches ---------------------------------
// The evaluations a proof writes after the challenge x (prover.rs:742-811) and the ones the multi-open argument takes
// (shplonk/prover.rs:117-131) are dozens of independent eval_polynomial calls: one job per pair, grid.y = job.
__global__ void __launch_bounds__(POLY_T)
poly_horner_batch_tile_reduce_kernel(const HornerJob* __restrict__ jobs, uint64_t n, uint64_t pad, uint32_t tiles,
                                     Fr* __restrict__ tile_val) {
    __shared__ PolyShared sh;
    const HornerJob* job = jobs + blockIdx.y;
    const Fr* __restrict__ a = job->a;
    const Fr b = fe_load(&job->b);
    const uint32_t t = threadIdx.x;
    const uint64_t j0 = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr pw[POLY_LOG_T];
    horner_powers(fe_load(&job->bk), pw);
    Fr s = Fr::zero();
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        Fr r = horner_view(a, n, pad, j0 + k);
        s = k == 0 ? r : fe_add(fe_mul(s, b), r);
    }
    int cur;
    Fr incl = block_scan_horner(s, pw, sh, t, cur);
    if (t == POLY_T - 1) fe_store(tile_val + (size_t)blockIdx.y * tiles + blockIdx.x, incl);
}
// one block per job folds that job's tile values: out[job] = sum_i tile_val[i] * bT^(tiles-1-i)
__global__ void __launch_bounds__(POLY_T)
poly_horner_batch_totals_kernel(const HornerJob* __restrict__ jobs, const Fr* __restrict__ tile_val, uint32_t tiles,
                                Fr* __restrict__ out) {
    __shared__ PolyShared sh;
    const HornerJob* job = jobs + blockIdx.x;
    const Fr* tv = tile_val + (size_t)blockIdx.x * tiles;
    const Fr bt = fe_load(&job->bt);
    const uint32_t t = threadIdx.x;
    Fr pw[POLY_LOG_T];
    horner_powers(bt, pw);
    Fr carry = Fr::zero();
    for (uint32_t base = 0; base < tiles; base += POLY_T) {
        Fr v = base + t < tiles ? fe_load(tv + base + t) : Fr::zero();
        if (t == 0) v = fe_add(fe_mul(carry, bt), v);
        int cur;
        block_scan_horner(v, pw, sh, t, cur);
        uint32_t last = tiles - base < (uint32_t)POLY_T ? tiles - base - 1 : POLY_T - 1;
        Fr o = fe_load(&sh.s[cur][last]);
        __syncthreads();
        carry = o;
    }
    if (t == 0) fe_store(out + blockIdx.x, carry);
}
// phase 3 (kate): q[n-2-j] = s_j for j <= n-2, seeded with the tile carry-in
__global__ void __launch_bounds__(POLY_T)
poly_horner_tile_scan_kernel(const Fr* __restrict__ a, uint64_t n, uint64_t pad, Fr b, Fr bk, const Fr* __restrict__ tile_cin,
                             Fr* __restrict__ q) {
    __shared__ PolyShared sh;
    const uint32_t t = threadIdx.x;
    const uint64_t j0 = (uint64_t)blockIdx.x * POLY_TILE + (uint64_t)t * POLY_K;
    Fr pw[POLY_LOG_T];
    horner_powers(bk, pw);
    Fr r[POLY_K];
    Fr s = t == 0 ? fe_load(tile_cin + blockIdx.x) : Fr::zero();
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        r[k] = horner_view(a, n, pad, j0 + k);
        s = (k == 0 && t != 0) ? r[0] : fe_add(fe_mul(s, b), r[k]);
    }
    int cur;
    block_scan_horner(s, pw, sh, t, cur);
    Fr c = t ? fe_load(&sh.s[cur][t - 1]) : fe_load(tile_cin + blockIdx.x);
#pragma unroll
    for (int k = 0; k < POLY_K; k++) {
        uint64_t j = j0 + k;
        c = fe_add(fe_mul(c, b), r[k]);
        if (j >= pad && j - pad + 1 < n) fe_store(q + (n - 2 - (j - pad)), c);
    }
}
// out[i] = a[i] * b[i] (* scale): the coefficient scaling a_t * s^t in front of a coset transform (distribute_powers_zeta,
// domain.rs:325-341, for an arbitrary shift s given as its power table) and its inverse afterwards.  out may alias a.
__global__ void __launch_bounds__(256)
poly_mul_kernel(const Fr* __restrict__ a, const Fr* __restrict__ b, Fr scale, uint32_t has_scale, Fr* out, uint64_t n) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Fr v = fe_mul(fe_load(a + i), fe_load(b + i));
        if (has_scale) v = fe_mul(v, scale);
        fe_store(out + i, v);
    }
}
#endif  // __CUDACC__ && !H2B_NO_KERNELS

}  // namespace h2b
