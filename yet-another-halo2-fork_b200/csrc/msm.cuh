// msm.cuh -- bucket (Pippenger) multi-scalar multiplication over BN254 G1: the GPU side of
// MsmAccel::msm / best_multiexp (reference halo2_middleware/src/zal.rs:57-58,136-138) behind
// ParamsKZG::commit / commit_lagrange (halo2_backend/src/poly/kzg/commitment.rs:305-317,354-366).
//
// Unlike the reference's chunk-per-thread CPU Pippenger (which replicates the bucket set per thread) the
// GPU pipeline keeps ONE bucket set per window and bins all n*W signed digits with a counting sort:
//   1. digits+count   : scalar Montgomery -> canonical, signed base-2^c digits, histogram per (window,bucket)
//   2. scan           : exclusive prefix sum -> bucket segment offsets
//   3. scatter        : point index (+ sign bit) written into its bucket segment
//   4. order          : buckets ordered by population so that the 32 lanes of a warp run equal trip counts
//   5. accumulate     : one thread per bucket, XYZZ mixed additions (8M+2S) over its segment;
//                       oversized buckets are split into block-cooperative chunks (tree-reduced)
//   6. reduce         : per window sum_b b*B_b by chunked running sums + block tree reduction
//   7. window combine : Horner over the W window sums (host, 64-bit limbs) -> affine
// All curve arithmetic runs on the integer-multiply pipe (IMAD.WIDE carry chains, field.cuh).
#pragma once
#include "curve.cuh"

namespace h2b {

struct MsmGeom {
    uint32_t c;          // window bits
    uint32_t windows;    // W = ceil(255 / c)
    uint32_t nb;         // buckets per window = 2^(c-1)
    uint32_t n;          // number of (scalar, base) pairs
    uint32_t pre_stride; // 0: classic (one bucket set per window, bases[i]).  > 0: the bases buffer holds the
                         // precomputed multiples pre[j * pre_stride + i] = 2^(c*j) * P_i, so all windows share ONE
                         // bucket set: digit j of scalar i goes to bucket |d| with entry index j * pre_stride + i.
    uint32_t idx_offset; // index of scalars[0] within the bases (an MSM streamed in chunks keeps global indices)
};

H2B_HD constexpr uint32_t msm_windows(uint32_t c) { return (255u + c - 1u) / c; }

// Signed window digits of a canonical (non-Montgomery) scalar: value = sum_j d_j * 2^(c*j),
// d_j in [-2^(c-1)+1, 2^(c-1)].  Calls f(j, d_j) for each window.
template <class F>
H2B_HD void msm_signed_digits(const Fr& canon, uint32_t c, uint32_t windows, F f) {
    uint32_t carry = 0;
    const uint32_t mask = (1u << c) - 1u, half = 1u << (c - 1);
    for (uint32_t j = 0; j < windows; j++) {
        uint32_t bit = j * c, word = bit >> 5, off = bit & 31u;
        uint32_t v = 0;
        if (word < 8) {
            v = canon.l[word] >> off;
            if (off && word + 1 < 8) v |= canon.l[word + 1] << (32u - off);
        }
        v = (v & mask) + carry;
        int32_t d;
        if (v > half) { d = (int32_t)v - (int32_t)(1u << c); carry = 1; }
        else { d = (int32_t)v; carry = 0; }
        f(j, d);
    }
}

// Running-sum reduction of `len` consecutive buckets holding values v0, v0+1, ..., v0+len-1:
// returns sum_v v * B_v  (chunked summation by parts + one small scalar multiplication).
H2B_HD Xyzz msm_reduce_chunk(const Xyzz* buckets, uint32_t len, uint32_t v0) {
    Xyzz run = Xyzz::identity(), acc = Xyzz::identity();
    for (uint32_t i = len; i-- > 0;) {
        run = xyzz_add(run, xyzz_load(buckets + i));
        acc = xyzz_add(acc, run);
    }
    // acc = sum (v - v0 + 1) B_v ;  add (v0 - 1) * run
    uint32_t k = v0 - 1;
    if (k) {
        Xyzz m = Xyzz::identity();
        for (int bit = 31; bit >= 0; bit--) {
            m = xyzz_double(m);
            if ((k >> bit) & 1u) m = xyzz_add(m, run);
        }
        acc = xyzz_add(acc, m);
    }
    return acc;
}

// Summation by parts over `len` consecutive buckets: returns sum_{i<len} (i+1) * B_i and, in `run`, sum_i B_i.
H2B_HD Xyzz msm_chunk_sums(const Xyzz* buckets, uint32_t len, Xyzz& run) {
    run = Xyzz::identity();
    Xyzz acc = Xyzz::identity();
    for (uint32_t i = len; i-- > 0;) {
        run = xyzz_add(run, xyzz_load(buckets + i));
        acc = xyzz_add(acc, run);
    }
    return acc;
}
H2B_HD Xyzz xyzz_mul_pow2(Xyzz p, uint32_t log2k) {
    for (uint32_t i = 0; i < log2k; i++) p = xyzz_double(p);
    return p;
}

// host: sum_t 2^(exps[t]) * terms[t] by one Horner pass over the distinct exponents (largest first).  The bucket reduction
// leaves its last, strictly serial step -- ~20 dependent doublings and additions of single points -- to the host: a lone GPU
// thread needs ~9 us per dependent point operation, the host ~0.5 us, and the terms travel with the result copy anyway.
inline Xyzz msm_horner_terms(const Xyzz* terms, const uint32_t* exps, uint32_t n) {
    uint32_t idx[64];
    if (n > 64) n = 64;
    for (uint32_t i = 0; i < n; i++) idx[i] = i;
    for (uint32_t i = 1; i < n; i++) {                      // insertion sort by exponent, descending
        uint32_t v = idx[i], j = i;
        while (j > 0 && exps[idx[j - 1]] < exps[v]) { idx[j] = idx[j - 1]; j--; }
        idx[j] = v;
    }
    Xyzz acc = Xyzz::identity();
    uint32_t cur = n ? exps[idx[0]] : 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t e = exps[idx[i]];
        for (; cur > e; cur--) acc = xyzz_double(acc);
        acc = xyzz_add(acc, terms[idx[i]]);
    }
    for (; cur > 0; cur--) acc = xyzz_double(acc);
    return acc;
}

// host: Horner over the window sums (most significant last), normalise, and write the result as the
// reference's G1 Jacobian layout: (x, y, 1), or (0, 1, 0) for the identity.
inline void msm_finish(const Xyzz* windows, uint32_t nw, uint32_t c, void* out96) {
    Xyzz acc = Xyzz::identity();
    for (uint32_t j = nw; j-- > 0;) {
        for (uint32_t i = 0; i < c; i++) acc = xyzz_double(acc);
        acc = xyzz_add(acc, windows[j]);
    }
    Affine a = xyzz_to_affine(acc);
    Fq out[3];
    if (acc.is_identity()) { out[0] = Fq::zero(); out[1] = Fq::one(); out[2] = Fq::zero(); }
    else { out[0] = a.x; out[1] = a.y; out[2] = Fq::one(); }
    for (int i = 0; i < 3; i++)
        for (int l = 0; l < 8; l++) reinterpret_cast<uint32_t*>(out96)[i * 8 + l] = out[i].l[l];
}

// launch geometry constants and small records shared by the kernels (msm_kernels.cu) and the host orchestration
constexpr int SCAN_BLOCK = 1024;
constexpr uint32_t MSM_SIZE_BINS = 4096;
constexpr int MSM_LONG_THREADS = 128;
constexpr uint32_t MSM_LONG_CHUNK = 4096;
struct MsmLongInfo { uint32_t key, task_base, ntasks, pad; };
struct MsmLongTask { uint32_t key, chunk; };
constexpr int MSM_REDUCE_THREADS = 128;
constexpr int MSM_REDUCE_MAX_SLICES = 256;
constexpr int MSM_MASKED_THREADS = 256;
constexpr uint32_t MSM_MAX_TERMS = 64;           // terms one bucket set is reduced to on the device (<= 2 * 24 index bits + 1)
// Peer exchange of the reduced terms between the ranks of a point-range sharded MSM (one process per GPU): every rank owns a
// mailbox of 2 parities x MSM_PEER_MAX_RANKS slots; a slot holds up to MSM_PEER_SLOT_POINTS points, and one 64-bit epoch flag.
constexpr uint32_t MSM_PEER_MAX_RANKS = 16;
constexpr uint32_t MSM_PEER_SLOT_POINTS = 64;
struct MsmPeerTable { void* mailbox[MSM_PEER_MAX_RANKS]; };
H2B_HD constexpr size_t msm_peer_slot_bytes() { return (size_t)MSM_PEER_SLOT_POINTS * 128; }
H2B_HD constexpr size_t msm_peer_flags_offset() { return 2 * (size_t)MSM_PEER_MAX_RANKS * msm_peer_slot_bytes(); }
H2B_HD constexpr size_t msm_peer_tags_offset() { return msm_peer_flags_offset() + 2 * (size_t)MSM_PEER_MAX_RANKS * 8; }
H2B_HD constexpr size_t msm_peer_mailbox_bytes() { return msm_peer_tags_offset() + 2 * (size_t)MSM_PEER_MAX_RANKS * 8; }

#if defined(__CUDACC__) && !defined(H2B_NO_KERNELS)
// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
// 1. histogram of (window, |digit|) keys.  counts has windows*nb entries.
// canon_out (optional): the canonical scalars, kept for msm_scatter_kernel so that the Montgomery -> canonical multiplication is
// done once per scalar.
__global__ void msm_count_kernel(MsmGeom g, const Fr* __restrict__ scalars, uint32_t* __restrict__ counts, Fr* __restrict__ canon_out) {
    // grid-stride: the launch may be narrower than the chunk (msm_sort_blocks) so that the kernel, which is bound by the L2 and
    // not by the SMs, takes fewer block slots away from the accumulation running beside it
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += gridDim.x * blockDim.x) {
    Fr canon = fe_from_mont(fe_load(scalars + i));
    if (canon_out) fe_store(canon_out + i, canon);
    msm_signed_digits(canon, g.c, g.windows, [&](uint32_t j, int32_t d) {
        if (d != 0) {
            uint32_t mag = d < 0 ? (uint32_t)(-d) : (uint32_t)d;
            atomicAdd(counts + (g.pre_stride ? 0 : (size_t)j * g.nb) + (mag - 1), 1u);
        }
    });
    }
}

// 3. scatter point indices into bucket segments; cursor starts as a copy of the segment offsets.
// canon (optional): what msm_count_kernel left; otherwise the scalar is converted again.
__global__ void msm_scatter_kernel(MsmGeom g, const Fr* __restrict__ scalars, const Fr* __restrict__ canon_in,
                                   uint32_t* __restrict__ cursor, uint32_t* __restrict__ entries) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += gridDim.x * blockDim.x) {
    Fr canon = canon_in ? fe_load(canon_in + i) : fe_from_mont(fe_load(scalars + i));
    msm_signed_digits(canon, g.c, g.windows, [&](uint32_t j, int32_t d) {
        if (d != 0) {
            uint32_t mag = d < 0 ? (uint32_t)(-d) : (uint32_t)d;
            uint32_t pos = atomicAdd(cursor + (g.pre_stride ? 0 : (size_t)j * g.nb) + (mag - 1), 1u);
            uint32_t idx = i + g.idx_offset;
            entries[pos] = (g.pre_stride ? j * g.pre_stride + idx : idx) | (d < 0 ? 0x80000000u : 0u);
        }
    });
    }
}

// ---- exclusive scan over uint32 (three small kernels; sizes up to a few million) ----------------
__global__ void scan_block_sums_kernel(const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t sh[32];
    uint32_t i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    uint32_t v = i < n ? in[i] : 0u;
    for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = sh[threadIdx.x];
        for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
    }
}
// in-place exclusive scan of `m` values by ONE block (m up to ~1e5); total written to *total
__global__ void scan_single_block_kernel(uint32_t* data, uint32_t m, uint32_t* total) {
    __shared__ uint32_t sh[SCAN_BLOCK];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < m; base += SCAN_BLOCK) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = i < m ? data[i] : 0u;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_BLOCK; o <<= 1) {
            uint32_t t = threadIdx.x >= (uint32_t)o ? sh[threadIdx.x - o] : 0u;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        uint32_t incl = sh[threadIdx.x];
        if (i < m) data[i] = carry + incl - v;
        __syncthreads();
        if (threadIdx.x == SCAN_BLOCK - 1) carry += incl;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total) *total = carry;
}
// out[i] = exclusive prefix of in within the block + block_offsets[block]; out2 (optional) gets a copy
__global__ void scan_apply_kernel(const uint32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ block_offsets,
                                  uint32_t* __restrict__ out, uint32_t* __restrict__ out2) {
    __shared__ uint32_t sh[SCAN_BLOCK];
    uint32_t i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    uint32_t v = i < n ? in[i] : 0u;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < SCAN_BLOCK; o <<= 1) {
        uint32_t t = threadIdx.x >= (uint32_t)o ? sh[threadIdx.x - o] : 0u;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    if (i < n) {
        uint32_t e = block_offsets[blockIdx.x] + sh[threadIdx.x] - v;
        out[i] = e;
        if (out2) out2[i] = e;
    }
}

// ---- registration-time precomputation: out[j * stride + i] = 2^(c*j) * in[i], j < windows (affine) ------------
// One thread per point walks the whole doubling chain in XYZZ and normalises the multiples MSM_PRE_BATCH at a time with ONE
// inversion (Montgomery's trick over the ZZ*ZZZ of the batch; the unnormalised X, Y wait in `out`, the denominators in local
// memory): 8 multiplications per multiple instead of a ~300-multiplication Fermat inversion each.  A one-off cost per registered SRS.
constexpr uint32_t MSM_PRE_BATCH = 16;
__global__ void __launch_bounds__(128)
msm_precompute_kernel(const Affine* __restrict__ in, Affine* __restrict__ out, uint32_t n, uint32_t stride, uint32_t c,
                      uint32_t windows) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Affine p = affine_load(in + i);
    fe_store(&out[i].x, p.x);
    fe_store(&out[i].y, p.y);
    if (p.is_identity()) {
        for (uint32_t j = 1; j < windows; j++) {
            fe_store(&out[(size_t)j * stride + i].x, p.x);
            fe_store(&out[(size_t)j * stride + i].y, p.y);
        }
        return;
    }
    Xyzz t = Xyzz::from_affine(p);
    Fq den[MSM_PRE_BATCH], pref[MSM_PRE_BATCH], zz[MSM_PRE_BATCH];
    for (uint32_t j0 = 1; j0 < windows; j0 += MSM_PRE_BATCH) {
        uint32_t m = windows - j0 < MSM_PRE_BATCH ? windows - j0 : MSM_PRE_BATCH;
        Fq run = Fq::one();
        for (uint32_t k = 0; k < m; k++) {
            for (uint32_t d = 0; d < c; d++) t = xyzz_double(t);          // never the identity: the group order is odd
            Affine* o = out + (size_t)(j0 + k) * stride + i;
            fe_store(&o->x, t.x);
            fe_store(&o->y, t.y);
            zz[k] = t.zz;
            den[k] = fe_mul(t.zz, t.zzz);
            run = fe_mul(run, den[k]);
            pref[k] = run;
        }
        Fq inv = fe_inv(run);                                             // 1 / (den_0 * ... * den_{m-1})
        for (uint32_t k = m; k-- > 0;) {
            Fq ik = k ? fe_mul(inv, pref[k - 1]) : inv;                   // 1 / (ZZ_k * ZZZ_k)
            if (k) inv = fe_mul(inv, den[k]);
            Affine* o = out + (size_t)(j0 + k) * stride + i;
            // only ZZ was kept: 1/ZZZ = ik * ZZ, and 1/ZZ = (1/ZZZ)^2 * ZZ^2  (Z^-6 * Z^4)
            Fq izzz = fe_mul(ik, zz[k]);
            Fq izz = fe_mul(fe_sqr(izzz), fe_sqr(zz[k]));
            fe_store(&o->x, fe_mul(fe_load(&o->x), izz));
            fe_store(&o->y, fe_mul(fe_load(&o->y), izzz));
        }
    }
}

// ---- 4. order buckets by population (descending), and list the oversized ones --------------------
// size_hist[MSM_SIZE_BINS]: bin = min(size, BINS-1); stored reversed so that an ascending scan gives descending size.
// Uniform scalars put almost every bucket into the same two or three bins, so both kernels first combine their block's keys
// in shared memory and touch each global bin once per block (global atomics on one address serialise: 100 us -> a few us).
__global__ void __launch_bounds__(256)
msm_size_hist_kernel(const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends, uint32_t nkeys,
                     uint32_t* __restrict__ size_hist) {
    __shared__ uint32_t sh[MSM_SIZE_BINS];
    for (uint32_t i = threadIdx.x; i < MSM_SIZE_BINS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nkeys) {
        uint32_t s = ends[k] - offsets[k];
        uint32_t bin = s < MSM_SIZE_BINS - 1 ? s : MSM_SIZE_BINS - 1;
        atomicAdd(&sh[MSM_SIZE_BINS - 1 - bin], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < MSM_SIZE_BINS; i += blockDim.x)
        if (sh[i]) atomicAdd(size_hist + i, sh[i]);
}
__global__ void __launch_bounds__(256)
msm_size_scatter_kernel(const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends, uint32_t nkeys,
                        uint32_t* __restrict__ size_cursor, uint32_t* __restrict__ order) {
    __shared__ uint32_t sh[MSM_SIZE_BINS];          // per-block count of each bin, then the block's base position in that bin
    for (uint32_t i = threadIdx.x; i < MSM_SIZE_BINS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t rbin = 0, rank = 0;
    if (k < nkeys) {
        uint32_t s = ends[k] - offsets[k];
        uint32_t bin = s < MSM_SIZE_BINS - 1 ? s : MSM_SIZE_BINS - 1;
        rbin = MSM_SIZE_BINS - 1 - bin;
        rank = atomicAdd(&sh[rbin], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < MSM_SIZE_BINS; i += blockDim.x)
        if (sh[i]) sh[i] = atomicAdd(size_cursor + i, sh[i]);
    __syncthreads();
    if (k < nkeys) order[sh[rbin] + rank] = k;
}

// ---- 5. bucket accumulation: one thread per bucket, in population order --------------------------
// Buckets larger than `long_threshold` are left to msm_accumulate_long_kernel.
// MINB = resident 128-thread blocks per SM the register allocation must allow (4 -> 128 regs, 5 -> 96, 6 -> 80);
// PF = pull the next point into L2 with prefetch.global.L2 instead of double-buffering it in registers.
template <int MINB, bool PF>
__global__ void __launch_bounds__(128, MINB)
msm_accumulate_kernel(const Affine* __restrict__ bases, const uint32_t* __restrict__ entries,
                      const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends,
                      const uint32_t* __restrict__ order, uint32_t nkeys, uint32_t long_threshold,
                      Xyzz* __restrict__ buckets, uint32_t add_to) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= nkeys) return;
    uint32_t key = order[tid];
    uint32_t e = offsets[key], end = ends[key];
    if (end - e > long_threshold) return;
    if (add_to && e == end) return;                        // later chunk of a streamed MSM, nothing new for this bucket
    Xyzz acc = add_to ? xyzz_load(buckets + key) : Xyzz::identity();
    if (PF) {
    // No register double-buffer: the next point is pulled into L2 while the current addition runs, which frees
    // ~16 registers for one more resident warp per scheduler.
    uint32_t ent = e < end ? entries[e] : 0u;
    while (e < end) {
        uint32_t e1 = e + 1;
        uint32_t ent_n = e1 < end ? entries[e1] : 0u;
        if (e1 < end) asm volatile("prefetch.global.L2 [%0];" ::"l"(bases + (ent_n & 0x7fffffffu)));
        Affine p = affine_load(bases + (ent & 0x7fffffffu));
        if (!p.is_identity()) xyzz_madd(acc, p.x, p.y, (ent >> 31) != 0);
        ent = ent_n; e = e1;
    }
    } else if (e < end) {
        uint32_t ent = entries[e];
        Affine p = affine_load(bases + (ent & 0x7fffffffu));
        while (true) {
            uint32_t e1 = e + 1;
            uint32_t ent_n = 0;
            Affine pn;
            bool more = e1 < end;
            if (more) {                                   // prefetch the next point under the current addition
                ent_n = entries[e1];
                pn = affine_load(bases + (ent_n & 0x7fffffffu));
            }
            if (!p.is_identity()) xyzz_madd(acc, p.x, p.y, (ent >> 31) != 0);
            if (!more) break;
            p = pn; ent = ent_n; e = e1;
        }
    }
    xyzz_store(buckets + key, acc);
}

// Oversized buckets (skewed scalars: selector-like columns put most points into one bucket) are cut into
// chunks of MSM_LONG_CHUNK entries; one block per chunk strides over its entries and tree-reduces through
// shared memory; a second kernel sums the chunk partials of each oversized bucket.
__global__ void msm_long_list_kernel(const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends, uint32_t nkeys,
                                     uint32_t long_threshold, MsmLongInfo* __restrict__ long_info,
                                     MsmLongTask* __restrict__ tasks, uint32_t* __restrict__ counters /* [0]=#long, [1]=#tasks */) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nkeys) return;
    uint32_t s = ends[k] - offsets[k];
    if (s > long_threshold) {
        uint32_t nt = (s + MSM_LONG_CHUNK - 1) / MSM_LONG_CHUNK;
        uint32_t li = atomicAdd(counters + 0, 1u);
        uint32_t base = atomicAdd(counters + 1, nt);
        MsmLongInfo info; info.key = k; info.task_base = base; info.ntasks = nt; info.pad = 0;
        long_info[li] = info;
        for (uint32_t t = 0; t < nt; t++) { MsmLongTask tk; tk.key = k; tk.chunk = t; tasks[base + t] = tk; }
    }
}
__global__ void __launch_bounds__(MSM_LONG_THREADS)
msm_accumulate_long_kernel(const Affine* __restrict__ bases, const uint32_t* __restrict__ entries,
                           const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ends,
                           const MsmLongTask* __restrict__ tasks, const uint32_t* __restrict__ counters,
                           Xyzz* __restrict__ partials) {
    __shared__ Xyzz sh[MSM_LONG_THREADS];
    uint32_t total = counters[1];
    for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
        MsmLongTask tk = tasks[task];
        uint32_t lo = offsets[tk.key] + tk.chunk * MSM_LONG_CHUNK, end = ends[tk.key];
        uint32_t hi = lo + MSM_LONG_CHUNK < end ? lo + MSM_LONG_CHUNK : end;
        Xyzz acc = Xyzz::identity();
        for (uint32_t e = lo + threadIdx.x; e < hi; e += MSM_LONG_THREADS) {
            uint32_t ent = entries[e];
            Affine p = affine_load(bases + (ent & 0x7fffffffu));
            if (!p.is_identity()) xyzz_madd(acc, p.x, p.y, (ent >> 31) != 0);
        }
        sh[threadIdx.x] = acc;
        __syncthreads();
        for (int s = MSM_LONG_THREADS / 2; s > 0; s >>= 1) {
            if (threadIdx.x < (uint32_t)s) sh[threadIdx.x] = xyzz_add(sh[threadIdx.x], sh[threadIdx.x + s]);
            __syncthreads();
        }
        if (threadIdx.x == 0) xyzz_store(partials + task, sh[0]);
        __syncthreads();
    }
}
__global__ void __launch_bounds__(MSM_LONG_THREADS)
msm_long_combine_kernel(const MsmLongInfo* __restrict__ long_info, const uint32_t* __restrict__ counters,
                        const Xyzz* __restrict__ partials, Xyzz* __restrict__ buckets, uint32_t add_to) {
    __shared__ Xyzz sh[MSM_LONG_THREADS];
    uint32_t nlong = counters[0];
    for (uint32_t li = blockIdx.x; li < nlong; li += gridDim.x) {
        MsmLongInfo info = long_info[li];
        Xyzz acc = Xyzz::identity();
        for (uint32_t t = threadIdx.x; t < info.ntasks; t += MSM_LONG_THREADS)
            acc = xyzz_add(acc, xyzz_load(partials + info.task_base + t));
        sh[threadIdx.x] = acc;
        __syncthreads();
        for (int s = MSM_LONG_THREADS / 2; s > 0; s >>= 1) {
            if (threadIdx.x < (uint32_t)s) sh[threadIdx.x] = xyzz_add(sh[threadIdx.x], sh[threadIdx.x + s]);
            __syncthreads();
        }
        if (threadIdx.x == 0) xyzz_store(buckets + info.key, add_to ? xyzz_add(sh[0], xyzz_load(buckets + info.key)) : sh[0]);
        __syncthreads();
    }
}

// ---- 6. bucket reduction --------------------------------------------------------------------------
// sum_v v * B_v per bucket set, with no scalar multiplication other than by powers of two:
//   level 1  grid = (slices, sets), 128 threads, L = 2^log_l buckets per thread.  Thread t: run_t = sum B,
//            acc_t = sum (i+1) B over its L buckets.  Block: A = sum_t acc_t, T = sum_t t * run_t (suffix scan + tree),
//            R = sum_t run_t.  Emits P = A + L * T (weights relative to the block's first bucket) and R.
//   level 2  one block per set over the `slices` (<= 256) block results:
//            sum_b P_b + (128 * L) * sum_b b * R_b      (same suffix-scan trick).
// block-wide helpers on a shared array of n (power of two) points
template <int N>
H2B_D void block_suffix_scan(Xyzz* sh, uint32_t t) {       // sh[t] <- sum_{u >= t} sh[u]
    for (int o = 1; o < N; o <<= 1) {
        Xyzz v = Xyzz::identity();
        bool take = t + (uint32_t)o < (uint32_t)N;
        if (take) v = sh[t + o];
        __syncthreads();
        if (take) sh[t] = xyzz_add(sh[t], v);
        __syncthreads();
    }
}
template <int N>
H2B_D void block_tree_sum(Xyzz* sh, uint32_t t) {          // sh[0] <- sum_t sh[t]
    for (int st = N / 2; st > 0; st >>= 1) {
        if (t < (uint32_t)st) sh[t] = xyzz_add(sh[t], sh[t + st]);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(MSM_REDUCE_THREADS)
msm_reduce_kernel(MsmGeom g, const Xyzz* __restrict__ buckets, uint32_t slices, uint32_t log_l,
                  Xyzz* __restrict__ block_p, Xyzz* __restrict__ block_r) {
    __shared__ Xyzz sh[MSM_REDUCE_THREADS];
    const uint32_t j = blockIdx.y, s = blockIdx.x, t = threadIdx.x;
    const uint32_t L = 1u << log_l;
    const uint32_t first = (s * MSM_REDUCE_THREADS + t) << log_l;       // bucket index within the set
    Xyzz run;
    Xyzz acc = msm_chunk_sums(buckets + (size_t)j * g.nb + first, L, run);
    // A = sum acc_t
    sh[t] = acc;
    __syncthreads();
    block_tree_sum<MSM_REDUCE_THREADS>(sh, t);
    Xyzz A = sh[0];
    __syncthreads();
    // suffix sums of run_t: S_0 = R, sum_{t>=1} S_t = sum_t t * run_t
    sh[t] = run;
    __syncthreads();
    block_suffix_scan<MSM_REDUCE_THREADS>(sh, t);
    Xyzz R = sh[0];
    __syncthreads();
    if (t == 0) sh[0] = Xyzz::identity();
    __syncthreads();
    block_tree_sum<MSM_REDUCE_THREADS>(sh, t);
    if (t == 0) {
        Xyzz P = xyzz_add(A, xyzz_mul_pow2(sh[0], log_l));
        xyzz_store(block_p + (size_t)j * slices + s, P);
        xyzz_store(block_r + (size_t)j * slices + s, R);
    }
}
__global__ void __launch_bounds__(MSM_REDUCE_MAX_SLICES)
msm_reduce_final_kernel(uint32_t slices, uint32_t log_block_buckets, const Xyzz* __restrict__ block_p,
                        const Xyzz* __restrict__ block_r, Xyzz* __restrict__ set_sums) {
    __shared__ Xyzz sh[MSM_REDUCE_MAX_SLICES];
    const uint32_t j = blockIdx.x, t = threadIdx.x;
    sh[t] = t < slices ? xyzz_load(block_p + (size_t)j * slices + t) : Xyzz::identity();
    __syncthreads();
    block_tree_sum<MSM_REDUCE_MAX_SLICES>(sh, t);
    Xyzz sumP = sh[0];
    __syncthreads();
    sh[t] = t < slices ? xyzz_load(block_r + (size_t)j * slices + t) : Xyzz::identity();
    __syncthreads();
    block_suffix_scan<MSM_REDUCE_MAX_SLICES>(sh, t);
    __syncthreads();
    if (t == 0) sh[0] = Xyzz::identity();
    __syncthreads();
    block_tree_sum<MSM_REDUCE_MAX_SLICES>(sh, t);
    if (t == 0) xyzz_store(set_sums + j, xyzz_add(sumP, xyzz_mul_pow2(sh[0], log_block_buckets)));
}
// ---- 6b. low-latency reduction for large bucket sets (nb >= 2^14) ----------------------------------------------------
// sum_i (i + 1) B_i with i = hi * W + lo (W = 2^w columns, H = 2^h rows):
//     = 2^w * sum_hi hi * D_hi  +  sum_lo lo * C_lo  +  sum_i B_i,        D = row sums, C = column sums.
// Every step is a plain (unweighted) sum, so nothing is serial beyond log-depth trees:
//   lines   one warp per SEGMENT of a row / column (seg_len buckets, 2 additions per bucket in total): strided lane sums, then
//           the four warps of a block fold their 4 x 32 lane sums through shared memory together (6 warp-steps per block
//           instead of 5 per warp).  Row segments land at D[seg * H + hi], column segments at C[seg * W + lo].
//   masked  grid (h + w + 1, sets): S^D_b = sum of the row partials whose hi has bit b set, S^C_b likewise over lo, V = sum D.
//   host    sum_b 2^(b + w) S^D_b + sum_b 2^b S^C_b + V  (msm_horner_terms): ~20 dependent point operations, 0.02 ms on the
//           host against 0.24 ms for a lone GPU thread.
__global__ void __launch_bounds__(MSM_REDUCE_THREADS, 4)      // 128 registers: four blocks per SM (at 130 only three fit)
msm_reduce_lines_kernel(const Xyzz* __restrict__ buckets, uint32_t nb, uint32_t w, uint32_t rs, uint32_t cs, Xyzz* __restrict__ lines) {
    __shared__ Xyzz sh[MSM_REDUCE_THREADS];
    const uint32_t W = 1u << w, H = nb >> w, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, set = blockIdx.y;
    const uint32_t njobs = rs * H + cs * W;
    const uint32_t job = blockIdx.x * (MSM_REDUCE_THREADS / 32) + warp;
    const Xyzz* b = buckets + (size_t)set * nb;
    Xyzz acc = Xyzz::identity();
    if (job < njobs) {
        uint32_t start, stride, count;
        if (job < rs * H) {                                                // segment `sg` of row hi
            const uint32_t sg = job / H, hi = job - sg * H, len = W / rs;
            start = hi * W + sg * len; stride = 1; count = len;
        } else {                                                           // segment `sg` of column lo
            const uint32_t j2 = job - rs * H, sg = j2 >> w, lo = j2 & (W - 1u), rows = H / cs;
            start = sg * rows * W + lo; stride = W; count = rows;
        }
        for (uint32_t i = lane; i < count; i += 32) acc = xyzz_add(acc, xyzz_load(b + start + (size_t)i * stride));
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    // fold 4 lines x 32 lane sums: at step `st` thread t works on line t / st, lane t % st
#pragma unroll 1
    for (uint32_t st = 16; st > 0; st >>= 1) {
        const uint32_t t = threadIdx.x;
        Xyzz v;
        const bool act = t < 4 * st;
        const uint32_t pos = (t / st) * 32 + (t % st);
        if (act) v = xyzz_add(sh[pos], sh[pos + st]);
        __syncthreads();
        if (act) sh[pos] = v;
        __syncthreads();
    }
    if (threadIdx.x < 4) {
        const uint32_t j = blockIdx.x * (MSM_REDUCE_THREADS / 32) + threadIdx.x;
        if (j < njobs) xyzz_store(lines + (size_t)set * njobs + j, sh[threadIdx.x * 32]);
    }
}
__global__ void __launch_bounds__(MSM_MASKED_THREADS)
msm_reduce_masked_kernel(const Xyzz* __restrict__ lines, uint32_t h, uint32_t w, uint32_t rs, uint32_t cs, Xyzz* __restrict__ terms) {
    __shared__ Xyzz sh[MSM_MASKED_THREADS];
    const uint32_t H = 1u << h, W = 1u << w, t = threadIdx.x, job = blockIdx.x, set = blockIdx.y;
    const Xyzz* D = lines + (size_t)set * (rs * H + cs * W);
    const bool cols = job >= h && job < h + w;
    const Xyzz* src = cols ? D + (size_t)rs * H : D;
    const uint32_t count = cols ? cs * W : rs * H;                    // segments of one line share its weight: index mod H (mod W)
    const uint32_t mask = job < h ? (1u << job) : (cols ? (1u << (job - h)) : 0u);     // 0: plain sum (V)
    Xyzz acc = Xyzz::identity();
    for (uint32_t i = t; i < count; i += MSM_MASKED_THREADS)
        if (mask == 0 || (i & mask)) acc = xyzz_add(acc, xyzz_load(src + i));
    sh[t] = acc;
    __syncthreads();
    block_tree_sum<MSM_MASKED_THREADS>(sh, t);
    if (t == 0) xyzz_store(terms + (size_t)set * (h + w + 1) + job, sh[0]);
}

// ---- 7. peer exchange (one process per GPU, MSM sharded by point range) ---------------------------------------------
// ONE kernel publishes this rank's reduced terms into every peer's mailbox over NVLink (plain peer stores + a release flag),
// waits for the same epoch from every peer, and folds the `world` contributions per term; the sums go straight into page-locked
// host memory.  No NCCL call, no host staging: the exchange is the epilogue of the reduction on the compute stream.
constexpr unsigned long long MSM_PEER_GAVE_UP = 1ull << 63;      // flag bit: "published for this epoch, then failed the call"
H2B_D void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
H2B_D unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
H2B_D unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__global__ void __launch_bounds__(256)
msm_peer_exchange_kernel(const Xyzz* __restrict__ my_terms, uint32_t n_terms, MsmPeerTable peers, uint32_t rank, uint32_t world,
                         uint32_t parity, unsigned long long epoch, unsigned long long tag, unsigned long long timeout_ns,
                         Xyzz* __restrict__ out_terms, uint32_t* __restrict__ status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Xyzz* sh = reinterpret_cast<Xyzz*>(smem_raw);                      // [n_terms][world]
    __shared__ uint32_t failed;
    const uint32_t t = threadIdx.x, nt = blockDim.x;
    if (t == 0) failed = 0;
    // 1. publish: n_terms * 128 bytes into slot (parity, rank) of every mailbox, 16 bytes per store
    const uint4* src = reinterpret_cast<const uint4*>(my_terms);
    const uint32_t vec = n_terms * 8;
    for (uint32_t r = 0; r < world; r++) {
        uint4* dst = reinterpret_cast<uint4*>(static_cast<unsigned char*>(peers.mailbox[r]) +
                                              ((size_t)parity * MSM_PEER_MAX_RANKS + rank) * msm_peer_slot_bytes());
        for (uint32_t i = t; i < vec; i += nt) dst[i] = src[i];
        if (t == 0)      // geometry tag (window bits, sets, terms): every rank must have reduced to the same weighted terms
            (reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(peers.mailbox[r]) + msm_peer_tags_offset()))[parity * MSM_PEER_MAX_RANKS + rank] = tag;
    }
    __threadfence_system();
    __syncthreads();
    if (t < world) {
        unsigned long long* flag = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(peers.mailbox[t]) +
                                                                         msm_peer_flags_offset()) + parity * MSM_PEER_MAX_RANKS + rank;
        st_release_sys_u64(flag, epoch);
    }
    // 2. wait for epoch from every peer (my own mailbox)
    if (t < world) {
        const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(
            static_cast<unsigned char*>(peers.mailbox[rank]) + msm_peer_flags_offset()) + parity * MSM_PEER_MAX_RANKS + t;
        const unsigned long long t0 = global_timer_ns();
        // exactly this epoch: a smaller value is the peer's previous use of this slot (or the cleared mailbox), a larger one means
        // the peer has made more calls than this rank and the slot no longer holds the terms of this call
        // (bit 63 set: the peer published for that epoch but then gave up on the call -- see the failure path below)
        for (;;) {
            const unsigned long long f = ld_acquire_sys_u64(flag), fe = f & ~MSM_PEER_GAVE_UP;
            if (fe == epoch) { if (f & MSM_PEER_GAVE_UP) atomicExch(&failed, 0x300u + t); break; }
            if (fe > epoch) { atomicExch(&failed, 0x200u + t); break; }
            if (global_timer_ns() - t0 > timeout_ns) { atomicExch(&failed, 1u + t); break; }
            __nanosleep(200);
        }
        const unsigned long long* tags = reinterpret_cast<const unsigned long long*>(
            static_cast<unsigned char*>(peers.mailbox[rank]) + msm_peer_tags_offset()) + parity * MSM_PEER_MAX_RANKS;
        if (!failed && __ldcg(tags + t) != tag) atomicExch(&failed, 0x100u + t);
    }
    __syncthreads();
    if (failed) {
        // This rank will report an error for this call.  A peer that has not looked yet must not fold the terms published above as
        // if the call had gone through everywhere (it might be a call this rank's group has already abandoned): mark them.
        if (t < world) {
            unsigned long long* flag = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(peers.mailbox[t]) +
                                                                             msm_peer_flags_offset()) + parity * MSM_PEER_MAX_RANKS + rank;
            st_release_sys_u64(flag, epoch | MSM_PEER_GAVE_UP);
        }
        if (t == 0) *status = failed;
        return;
    }
    // 3. fold over the ranks, per term (tree through shared memory)
    const unsigned char* mine = static_cast<const unsigned char*>(peers.mailbox[rank]) + (size_t)parity * MSM_PEER_MAX_RANKS * msm_peer_slot_bytes();
    for (uint32_t i = t; i < n_terms * world; i += nt) {
        const uint32_t term = i / world, r = i - term * world;
        // (L2 loads: the slots were written by other GPUs; nothing of them may be served from this SM's L1)
        const uint4* q = reinterpret_cast<const uint4*>(reinterpret_cast<const Xyzz*>(mine + (size_t)r * msm_peer_slot_bytes()) + term);
        uint4* o = reinterpret_cast<uint4*>(sh + i);
#pragma unroll
        for (int k = 0; k < 8; k++) o[k] = __ldcg(q + k);
    }
    __syncthreads();
    uint32_t width = world;
    while (width > 1) {
        const uint32_t half = (width + 1) / 2;
        for (uint32_t i = t; i < n_terms * (width / 2); i += nt) {
            const uint32_t term = i / (width / 2), r = i - term * (width / 2);
            sh[term * world + r] = xyzz_add(sh[term * world + r], sh[term * world + r + half]);
        }
        __syncthreads();
        width = half;
    }
    for (uint32_t i = t; i < n_terms; i += nt) xyzz_store(out_terms + i, sh[i * world]);
    if (t == 0) *status = 0;
}

// small-window fallback: nb < MSM_REDUCE_THREADS -> one thread per window does the whole running sum
__global__ void msm_reduce_small_kernel(MsmGeom g, const Xyzz* __restrict__ buckets, Xyzz* __restrict__ window_partials) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.windows) return;
    xyzz_store(window_partials + j, msm_reduce_chunk(buckets + (size_t)j * g.nb, g.nb, 1));
}
#endif  // __CUDACC__ && !H2B_NO_KERNELS

}  // namespace h2b
