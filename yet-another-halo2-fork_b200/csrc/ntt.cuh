// ntt.cuh -- multi-pass shared-memory Fr NTT (the GPU side of best_fft / EvaluationDomain transforms;
// reference call sites halo2_backend/src/poly/domain.rs:216-351).
//
// Contract (same as best_fft): natural order in, natural order out, X[k] = sum_n x[n] * omega^(n*k).
//
// Decomposition.  N = 2^L is split into m <= 4 "passes" of sub-size S_j = 2^{s_j} (s_j <= 9).  With
// n = M*n1 + n' and k = k1 + N1*k' (M = N/N1):
//     X[k1 + N1*k'] = sum_{n'} [ ( sum_{n1} x[M*n1 + n'] w_{N1}^{n1 k1} ) * w_N^{n' k1} ] * (w^{N1})^{n' k'}
// so a COLUMN pass does size-S sub-NTTs over stride-M columns, multiplies by the boundary twiddle
// w^{n' k1} and stores in place (row k1, column n'); the remaining size-M problem on each contiguous row is
// solved the same way; the final ROW pass reads contiguous rows and writes the transposed (natural-order)
// result.  Every pass moves each element through HBM exactly once in, once out.
//
// Tile.  A CTA owns S x 8 elements: 8 adjacent columns (column pass: 256-byte coalesced runs both ways) or 8
// rows whose outputs are adjacent (row pass).  One thread per 8 elements; a sub-NTT is <= 3 register
// radix-8 / 4 / 2 levels (5 + 7 multiplications per radix-8 level: butterflies + level twiddles), data
// exchanged between levels through shared memory held as two 16-byte planes with a 9-element row pitch so
// that every quarter-warp LDS.128/STS.128 is bank-conflict free.  The level-twiddle table (S entries, <= 16
// KB) is staged into shared memory with one TMA bulk copy (cp.async.bulk + mbarrier) overlapped with the
// first global loads.  The coset scaling, zero extension, 1/n scaling, coset un-scaling and truncation of
// the EvaluationDomain wrappers (domain.rs:230-293, 325-351) are fused into the first load / last store.
//
// All per-thread logic lives in __host__ __device__ functions (`ntt_level`) with no register state carried
// across a barrier, so the host test shim (host_check.cpp) runs the very same index logic thread by thread.
#pragma once
#include "field.cuh"

namespace h2b {

constexpr int NTT_LANES = 8;
constexpr int NTT_PITCH = 9;          // shared-memory pitch (elements) per sub-NTT position
constexpr int NTT_MAX_LOG_S = 9;

enum : uint32_t { NTT_IN_PATTERN = 1u, NTT_OUT_PATTERN = 2u, NTT_OUT_SCALE = 4u, NTT_IN_TABLE = 8u };

struct NttPass {
    uint32_t log_s;        // sub-NTT size S = 2^log_s
    uint32_t log_inner;    // column pass: log2 of the stride between consecutive sub-NTT inputs
    uint32_t log_outer;    // column pass: log2(#independent outer blocks); row pass: log2(#rows)
    uint32_t n_prev;       // row pass: number of earlier passes
    uint32_t prev_log[3];  // row pass: their log sizes, first pass first
    uint32_t tw_h;         // boundary twiddle = tw_lo[e & (2^h-1)] * tw_hi[e >> h]
    uint32_t flags;
    uint32_t in_tab_mask;  // NTT_IN_TABLE: input[i] *= in_tab[i & in_tab_mask]  (divide_by_vanishing_poly, domain.rs:305-310)
    uint64_t in_len;       // inputs at natural index >= in_len read as zero (first pass only)
    uint64_t out_len;      // outputs at natural index >= out_len are not written (row pass only)
    const Fr* tw_s;        // S entries: (omega^(N/S))^e
    const Fr* tw_lo;
    const Fr* tw_hi;
    const Fr* tw_full;     // optional full boundary table: tw_full[k * inner + i] = scale * omega^(outer * i * k)
    const Fr* in_tab;
    Fr in_pat[2];          // multipliers for input index % 3 == 1, 2   (distribute_powers_zeta, domain.rs:325)
    Fr out_pat[2];         // multipliers for output index % 3 == 1, 2
    Fr out_scale;          // single-pass transforms only (multi-pass folds the scale into tw_hi)
};

namespace detail {

H2B_HD constexpr int lev_count(int s) { return (s + 2) / 3; }
H2B_HD constexpr int lev_log_r(int s, int lev) { return lev < s / 3 ? 3 : s % 3; }
H2B_HD constexpr int lev_log_b(int s, int lev) { return s - 3 * lev; }

// size-R NTT in registers, decimation in frequency; x[j] ends up holding X[bitrev(j)].
template <int LOG_R>
H2B_HD void butterflies(Fr x[8], const Fr& w1, const Fr& w2, const Fr& w3) {
    if (LOG_R == 3) {
        Fr t[4];
#pragma unroll
        for (int a = 0; a < 4; a++) {
            t[a] = fe_sub(x[a], x[a + 4]);
            x[a] = fe_add(x[a], x[a + 4]);
        }
        x[4] = t[0];
        FePair<FrParams> pr = fe_mul2_call(t[1], w1, t[2], w2);
        x[5] = pr.a; x[6] = pr.b;
        x[7] = fe_mul_call(t[3], w3);
        Fr u[2];
#pragma unroll
        for (int h = 0; h < 8; h += 4) {
            Fr t0 = fe_sub(x[h], x[h + 2]);
            x[h] = fe_add(x[h], x[h + 2]);
            x[h + 2] = t0;
            u[h >> 2] = fe_sub(x[h + 1], x[h + 3]);
            x[h + 1] = fe_add(x[h + 1], x[h + 3]);
        }
        pr = fe_mul2_call(u[0], w2, u[1], w2);
        x[3] = pr.a; x[7] = pr.b;
#pragma unroll
        for (int h = 0; h < 8; h += 2) {
            Fr tt = fe_sub(x[h], x[h + 1]);
            x[h] = fe_add(x[h], x[h + 1]);
            x[h + 1] = tt;
        }
    } else if (LOG_R == 2) {
        Fr t0 = fe_sub(x[0], x[2]);
        x[0] = fe_add(x[0], x[2]);
        x[2] = t0;
        Fr t1 = fe_sub(x[1], x[3]);
        x[1] = fe_add(x[1], x[3]);
        x[3] = fe_mul_call(t1, w2);       // w2 = primitive 4th root
        Fr t = fe_sub(x[0], x[1]);
        x[0] = fe_add(x[0], x[1]);
        x[1] = t;
        t = fe_sub(x[2], x[3]);
        x[2] = fe_add(x[2], x[3]);
        x[3] = t;
    } else {
        Fr t = fe_sub(x[0], x[1]);
        x[0] = fe_add(x[0], x[1]);
        x[1] = t;
    }
}
H2B_HD constexpr int bitrev(int j, int log_r) {
    return log_r == 3 ? (((j & 1) << 2) | (j & 2) | ((j >> 2) & 1)) : (log_r == 2 ? (((j & 1) << 1) | ((j >> 1) & 1)) : j);
}

// sub-NTT output index of the first element of last-level block `blk` (all earlier levels are radix 8)
template <int LOG_S>
H2B_HD uint32_t block_to_kbase(uint32_t blk) {
    constexpr int NLEV = lev_count(LOG_S);
    // blk = ka_0 * 8^(NLEV-2) + ka_1 * 8^(NLEV-3) + ... ; k_base = ka_0 + 8*ka_1 + ...
    uint32_t k = 0;
#pragma unroll
    for (int l = 0; l < NLEV - 1; l++) {
        uint32_t ka = (blk >> (3 * (NLEV - 2 - l))) & 7u;
        k |= ka << (3 * l);
    }
    return k;
}

// row pass: storage row index o (= k1*N2*N3.. + k2*N3.. + ..) of the row whose results land at
// natural positions rev + outer*k, rev = k1 + N1*k2 + N1*N2*k3
H2B_HD uint64_t rev_to_row(const NttPass& P, uint64_t rev) {
    uint64_t o = 0;
    uint32_t shift_out = P.log_outer;
    for (uint32_t j = 0; j < P.n_prev; j++) {
        uint64_t kj = rev & ((1ull << P.prev_log[j]) - 1);
        rev >>= P.prev_log[j];
        shift_out -= P.prev_log[j];
        o |= kj << shift_out;
    }
    return o;
}

H2B_HD uint32_t mod3(uint64_t v) { return (uint32_t)(v % 3ull); }

#if defined(__CUDA_ARCH__)
H2B_D void mbar_wait(uint32_t mbar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(mbar), "r"(phase) : "memory");
}
#endif

}  // namespace detail

// One level of one pass for thread `t` of tile `tile`.
//   LEV == 0 reads global memory, LEV == last writes global memory, otherwise shared memory planes.
//   `mbar` (device only): shared address of the mbarrier guarding the TMA-staged twiddle table `stw`.
template <int LOG_S, int LEV, bool ROW>
H2B_HD void ntt_level(const NttPass& P, uint32_t t, uint64_t tile, const Fr* in, Fr* out,
                      U4* slo, U4* shi, const Fr* stw, uint32_t mbar) {
    using namespace detail;
    constexpr int S = 1 << LOG_S;
    constexpr int NLEV = lev_count(LOG_S);
    constexpr int LOG_R = lev_log_r(LOG_S, LEV);
    constexpr int R = 1 << LOG_R;
    constexpr int LOG_B = lev_log_b(LOG_S, LEV);
    constexpr int LOG_Q = LOG_B - LOG_R;           // log2(block size / radix) = stride between a-values
    constexpr bool FIRST = LEV == 0, LAST = LEV == NLEV - 1;
    constexpr int GROUPS = 8 / R;                  // radix-R groups handled per thread
    if (t >= (uint32_t)S) return;

    // tile geometry
    uint64_t col_base = 0;                         // column pass: address of element (o, 0, i0)
    uint64_t i0 = 0;
    uint32_t nlanes = NTT_LANES;
    if (!ROW) {
        uint64_t tiles_per_outer = 1ull << (P.log_inner - 3);
        uint64_t o = tile >> (P.log_inner - 3);
        i0 = (tile & (tiles_per_outer - 1)) << 3;
        col_base = (o << (LOG_S + P.log_inner)) + i0;
    } else {
        if (P.log_outer < 3) nlanes = 1u << P.log_outer;
    }

#pragma unroll
    for (int q = 0; q < GROUPS; q++) {
        uint32_t u = t * GROUPS + q;               // unit in [0, (S/R)*8)
        uint32_t lane, sg;
        if (ROW && FIRST) { sg = u & ((S >> LOG_R) - 1); lane = u >> (LOG_S - LOG_R); }   // positions fastest: coalesced row reads
        else { lane = u & 7u; sg = u >> 3; }
        bool active = lane < nlanes;
        uint32_t blk = sg >> LOG_Q, b = sg & ((1u << LOG_Q) - 1);
        uint32_t p0 = (blk << LOG_B) | b;

        uint64_t row_base = 0, rev = 0;
        if (ROW && (FIRST || LAST)) {
            rev = tile * NTT_LANES + lane;
            if (active) row_base = rev_to_row(P, rev) << LOG_S;
        }

        Fr x[8];
        // ---------------- load
#pragma unroll
        for (int a = 0; a < R; a++) {
            uint32_t p = p0 + ((uint32_t)a << LOG_Q);
            if (FIRST) {
                uint64_t addr = ROW ? row_base + p : col_base + ((uint64_t)p << P.log_inner) + lane;
                if (active && addr < P.in_len) {
                    x[a] = fe_load(in + addr);
                    if (P.flags & NTT_IN_TABLE) x[a] = fe_mul_call(x[a], fe_load(P.in_tab + (addr & P.in_tab_mask)));
                    if (P.flags & NTT_IN_PATTERN) {
                        uint32_t m = mod3(addr);
                        if (m) x[a] = fe_mul_call(x[a], P.in_pat[m - 1]);
                    }
                } else {
                    x[a] = Fr::zero();
                }
            } else {
                uint32_t si = p * NTT_PITCH + lane;
                x[a] = fe_join<FrParams>(slo[si], shi[si]);
            }
        }
#if defined(__CUDA_ARCH__)
        if (FIRST && q == 0) mbar_wait(mbar, 0);
#endif
        // ---------------- butterflies
        {
            Fr w1, w2, w3;
            if (LOG_R == 3) { w1 = fe_load(stw + (S >> 3)); w2 = fe_load(stw + 2 * (S >> 3)); w3 = fe_load(stw + 3 * (S >> 3)); }
            else if (LOG_R == 2) { w2 = fe_load(stw + (S >> 2)); w1 = w2; w3 = w2; }
            else { w1 = Fr::zero(); w2 = w1; w3 = w1; }
            butterflies<LOG_R>(x, w1, w2, w3);
        }
        // ---------------- level twiddles + store
        uint32_t kbase = LAST ? block_to_kbase<LOG_S>(blk) : 0;
        if (!LAST) {
            // level twiddles, two per out-of-line call
#pragma unroll
            for (int ka = 1; ka + 1 < R; ka += 2) {
                FePair<FrParams> pr = fe_mul2_call(x[bitrev(ka, LOG_R)], fe_load(stw + ((b * (uint32_t)ka) << (LOG_S - LOG_B))),
                                                   x[bitrev(ka + 1, LOG_R)], fe_load(stw + ((b * (uint32_t)(ka + 1)) << (LOG_S - LOG_B))));
                x[bitrev(ka, LOG_R)] = pr.a;
                x[bitrev(ka + 1, LOG_R)] = pr.b;
            }
            if (R >= 2) {
                constexpr int kl = R - 1;
                x[bitrev(kl, LOG_R)] = fe_mul_call(x[bitrev(kl, LOG_R)], fe_load(stw + ((b * (uint32_t)kl) << (LOG_S - LOG_B))));
            }
        }
        if (LAST && !ROW) {
            // column pass: boundary twiddle w^(outer*i*k), then store in place (row k, column i).
            // Twiddles are fetched four at a time ahead of their multiplications (two products per call).
            const uint64_t i = i0 + lane;
            constexpr int HALF = R >= 4 ? 4 : R;
#pragma unroll
            for (int base = 0; base < R; base += HALF) {
                Fr tw[HALF];
#pragma unroll
                for (int q2 = 0; q2 < HALF; q2++) {
                    uint32_t k = kbase + ((uint32_t)(base + q2) << (LOG_S - LOG_R));
                    if (P.tw_full) {
                        tw[q2] = fe_load(P.tw_full + ((uint64_t)k << P.log_inner) + i);
                    } else {
                        uint64_t e = (i * k) << P.log_outer;
                        tw[q2] = fe_mul_call(fe_load(P.tw_lo + (e & ((1ull << P.tw_h) - 1))), fe_load(P.tw_hi + (e >> P.tw_h)));
                    }
                }
#pragma unroll
                for (int q2 = 0; q2 < HALF; q2 += 2) {
                    uint32_t k0 = kbase + ((uint32_t)(base + q2) << (LOG_S - LOG_R));
                    if (q2 + 1 < HALF) {
                        uint32_t k1 = kbase + ((uint32_t)(base + q2 + 1) << (LOG_S - LOG_R));
                        FePair<FrParams> pr = fe_mul2_call(x[bitrev(base + q2, LOG_R)], tw[q2], x[bitrev(base + q2 + 1, LOG_R)], tw[q2 + 1 < HALF ? q2 + 1 : q2]);
                        fe_store(out + col_base + ((uint64_t)k0 << P.log_inner) + lane, pr.a);
                        fe_store(out + col_base + ((uint64_t)k1 << P.log_inner) + lane, pr.b);
                    } else {
                        fe_store(out + col_base + ((uint64_t)k0 << P.log_inner) + lane, fe_mul_call(x[bitrev(base + q2, LOG_R)], tw[q2]));
                    }
                }
            }
        }
#pragma unroll
        for (int ka = 0; ka < R; ka++) {
            Fr v = x[bitrev(ka, LOG_R)];
            if (!LAST) {
                uint32_t p = p0 + ((uint32_t)ka << LOG_Q);
                uint32_t si = p * NTT_PITCH + lane;
                fe_split(v, slo[si], shi[si]);
            } else {
                uint32_t k = kbase + ((uint32_t)ka << (LOG_S - LOG_R));      // sub-NTT output index
                if (!ROW) {
                    // handled above
                } else if (active) {
                    uint64_t idx = rev + ((uint64_t)k << P.log_outer);
                    if (idx < P.out_len) {
                        if (P.flags & NTT_OUT_SCALE) v = fe_mul_call(v, P.out_scale);
                        if (P.flags & NTT_OUT_PATTERN) {
                            uint32_t m = mod3(idx);
                            if (m) v = fe_mul_call(v, P.out_pat[m - 1]);
                        }
                        fe_store(out + idx, v);
                    }
                }
            }
        }
    }
}

#if defined(__CUDACC__)
// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
template <int LOG_S>
struct NttSmem {
    static constexpr int S = 1 << LOG_S;
    static constexpr size_t tw_bytes = (size_t)S * 32;
    static constexpr size_t plane_elems = (size_t)S * NTT_PITCH;
    static constexpr size_t bytes = tw_bytes + 2 * plane_elems * 16 + 16;
};

template <int LOG_S, bool ROW>
__global__ void __launch_bounds__((1 << LOG_S) < 32 ? 32 : (1 << LOG_S), (1 << LOG_S) < 32 ? 16 : 512 / (1 << LOG_S))
ntt_pass_kernel(const __grid_constant__ NttPass P, const Fr* __restrict__ in, Fr* __restrict__ out,
                uint64_t in_stride, uint64_t out_stride) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int NLEV = detail::lev_count(LOG_S);
    Fr* stw = reinterpret_cast<Fr*>(smem_raw);
    U4* slo = reinterpret_cast<U4*>(smem_raw + NttSmem<LOG_S>::tw_bytes);
    U4* shi = slo + NttSmem<LOG_S>::plane_elems;
    uint64_t* mbar_p = reinterpret_cast<uint64_t*>(shi + NttSmem<LOG_S>::plane_elems);
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(mbar_p);
    const uint32_t t = threadIdx.x;
    const uint64_t tile = blockIdx.x;
    in += (uint64_t)blockIdx.y * in_stride;
    out += (uint64_t)blockIdx.y * out_stride;

    // stage the level-twiddle table with one TMA bulk copy; completion is signalled on the mbarrier
    if (t == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (t == 0) {
        const uint32_t bytes = (uint32_t)NttSmem<LOG_S>::tw_bytes;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(stw);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(P.tw_s), "r"(bytes), "r"(mbar) : "memory");
    }
    ntt_level<LOG_S, 0, ROW>(P, t, tile, in, out, slo, shi, stw, mbar);
    if constexpr (NLEV > 1) {
        __syncthreads();
        ntt_level<LOG_S, 1, ROW>(P, t, tile, in, out, slo, shi, stw, mbar);
    }
    if constexpr (NLEV > 2) {
        __syncthreads();
        ntt_level<LOG_S, 2, ROW>(P, t, tile, in, out, slo, shi, stw, mbar);
    }
}

// out[i] = first * root^i, i < count   (table generation; square-and-multiply on the bits of i)
static __global__ void ntt_gen_powers_kernel(Fr* out, Fr root, Fr first, uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    Fr acc = first, base = root;
    for (uint32_t e = i; e; e >>= 1) {
        if (e & 1) acc = fe_mul(acc, base);
        base = fe_sqr(base);
    }
    fe_store(out + i, acc);
}
// full boundary table of a column pass: out[k * inner + i] = tw_lo[e & mask] * tw_hi[e >> h], e = (i * k) << log_outer
static __global__ void ntt_gen_boundary_kernel(Fr* out, const Fr* tw_lo, const Fr* tw_hi, uint32_t tw_h, uint32_t log_inner,
                                               uint32_t log_outer, uint64_t count) {
    uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= count) return;
    uint64_t k = idx >> log_inner, i = idx & ((1ull << log_inner) - 1);
    uint64_t e = (i * k) << log_outer;
    fe_store(out + idx, fe_mul(fe_load(tw_lo + (e & ((1ull << tw_h) - 1))), fe_load(tw_hi + (e >> tw_h))));
}
#endif  // __CUDACC__

}  // namespace h2b
