// msm_bin.cuh -- two-level binning of the signed window digits through shared memory: an ALTERNATIVE to stages 1-4 of msm.cuh's
// pipeline, selected with H2B_MSM_SORT=1 / h2b_set_option("msm_sort", "1").  NOT the default: measured on B200 it is slower.
//
// The default pipeline counts and scatters with one GLOBAL atomic per digit (27 M atomics + 27 M scattered 4-byte stores per
// 2^21-scalar chunk); both kernels sit at ~70 % of the L2 transaction rate and the scatter writes 2.3 x its output to DRAM as
// partially filled sectors (profiles/r2_msm_bin_before_2p24_details.csv).  This file is the attempt to replace that with
// shared-memory atomics and contiguous global stores only:
//   digits     one pass over the scalars: Montgomery -> canonical, signed digits, one packed word per (window, scalar)
//              [(key + 1) | sign << 31, 0 = zero digit] written coalesced; a block histograms the COARSE bin (key >> lb) of
//              its 2048-scalar tile in shared memory and adds the non-zero counters to the global bin totals;
//   (scan)     exclusive scan of the <= 512 bin totals;
//   partition  a block re-histograms its 1024-scalar tile, reserves one run per coarse bin with a single global atomicAdd each,
//              sorts its records by coarse bin in SHARED memory and copies run after run to global memory (a warp writes one
//              contiguous run: ~7 sectors per 26 records instead of 52) -- records of one coarse bin end up contiguous;
//   bin        a few blocks per coarse bin: shared-memory histogram of its <= 2^lb keys, block scan -> the bucket offsets / ends
//              (no global scan kernels), shared-memory cursors -> the slice's output assembled in shared memory and written
//              with coalesced stores, plus the population histogram the ordering stage needs (msm_size_hist_kernel folded in).
// Measured per 2^21-scalar chunk, each kernel alone under ncu (profiles/r2_msm_bin_tiled_2p24.txt): digits 0.116 ms, partition
// 0.143 ms, bin 0.294 ms, scan 0.02 ms = 0.57 ms against count 0.154 + scan 0.03 + scatter 0.353 + size_hist 0.011 = 0.55 ms for
// the global-atomic pipeline -- DRAM writes of the placement fall from 2.3 x to 0.7-1 x of the entries, but three passes of
// dependent shared-memory atomics cost more than the two passes of L2 atomics they replace, and inside the whole MSM the
// shared-memory-heavy blocks take more from the accumulation they overlap with (2^24: 38.1 ms against 36.5 ms).  What was tried on
// the way: `match.any` aggregation (~200 cycles per call), a ballot / shuffle / vote uniformity test per record (three dependent
// warp collectives), one load in flight per thread (DRAM-latency bound, 1.4 ms), scattered 4-byte stores from the bin kernel
// (0.42 ms; ~0.3 sectors / clock / SM is all an SM retires of them).  Kept as a tested option because its DRAM traffic is lower.
// Bucket contents are an unordered set (the group is commutative; the result is normalised to affine), so the order in which
// entries land inside a bucket is free, exactly as with the global-atomic version.
#pragma once
#include "msm.cuh"

namespace h2b {

constexpr int MSM_BIN_THREADS = 256;            // digits kernel
constexpr int MSM_PART_THREADS = 512;           // partition and bin kernels
constexpr uint32_t MSM_BIN_TILE = 2048;          // scalars per block of the digits kernel
constexpr uint32_t MSM_PART_TILE = 1024;         // scalars per block of the partition kernel (its records are staged in shared memory)
constexpr uint32_t MSM_PART_MAX_WINDOWS = 16;    // staging capacity: MSM_PART_TILE * 16 records (more windows: more, smaller tiles)
constexpr uint32_t MSM_BIN_MAX_COARSE = 512;     // coarse bins (shared-memory counters of the tile kernels)
constexpr uint32_t MSM_BIN_MIN_LB = 10, MSM_BIN_MAX_LB = 14;

// low key bits handled inside one coarse bin: at least 2^10 keys per bin, and never more than 512 bins
H2B_HD uint32_t msm_bin_low_bits(uint32_t nkeys) {
    uint32_t lb = MSM_BIN_MIN_LB;
    while (((nkeys + (1u << lb) - 1u) >> lb) > MSM_BIN_MAX_COARSE) lb++;
    return lb;
}
H2B_HD bool msm_bin_supported(uint32_t nkeys) { return msm_bin_low_bits(nkeys) <= MSM_BIN_MAX_LB; }

struct MsmBinKernelTable {
    void (*digits)(MsmGeom, const Fr*, uint32_t*, uint32_t*, uint32_t, uint32_t);
    void (*partition)(MsmGeom, const uint32_t*, uint32_t*, unsigned short*, uint32_t*, uint32_t, uint32_t, uint32_t);
    void (*bin)(const uint32_t*, const unsigned short*, const uint32_t*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t*, uint32_t*, uint32_t*,
                uint32_t*);
    int (*prepare)();                             // opt in to the dynamic shared memory of `bin`; returns a cudaError_t
};
const MsmBinKernelTable& msm_bin_kernel_table();
// shared memory of one bin block: counters + population histogram + the staged output image (two blocks per SM)
constexpr size_t MSM_BIN_SMEM_BUDGET = 110 * 1024;
H2B_HD size_t msm_bin_tables(uint32_t lb) { return ((size_t)1 << lb) * 4 + (size_t)MSM_SIZE_BINS * 4; }
H2B_HD uint32_t msm_bin_cap(uint32_t lb) { return (uint32_t)((MSM_BIN_SMEM_BUDGET - msm_bin_tables(lb)) / 4); }
// key slices per coarse bin so that a slice's expected output (records / bins / slices, plus a third) fits the staging image
H2B_HD uint32_t msm_bin_slices(uint32_t lb, uint64_t records, uint32_t ncoarse) {
    const uint64_t per_bin = records / (ncoarse ? ncoarse : 1) + 1;
    uint32_t s = 1;
    while (s < 16 && per_bin * 4 / 3 > (uint64_t)msm_bin_cap(lb) * s) s <<= 1;
    return s;
}
// scalars per partition tile for a given window count (the staging area holds MSM_PART_TILE * MSM_PART_MAX_WINDOWS records)
H2B_HD uint32_t msm_part_tile(uint32_t windows) {
    uint32_t t = MSM_PART_TILE;
    while (t > 32 && t * windows > MSM_PART_TILE * MSM_PART_MAX_WINDOWS) t >>= 1;
    return t;
}
H2B_HD size_t msm_part_smem() { return (size_t)MSM_PART_TILE * MSM_PART_MAX_WINDOWS * 6; }

#if defined(__CUDACC__) && defined(H2B_BIN_KERNELS)
// Shared-memory atomics in PTX.  `red.shared.add` of 1 becomes ATOMS.POPC.INC, which folds lanes holding the same address in
// hardware, so the histogram passes need no software aggregation even when a selector-like column sends every digit of a tile
// to one bin.  The slot-taking passes use the returning form; lanes with equal addresses serialise there (32-way at worst,
// only for skewed columns).  An explicit `match.any` aggregation costs ~200 cycles per call on sm_100a and a ballot / shuffle /
// vote uniformity test put three dependent warp collectives in front of every record: both measured slower (profiles/README.md).
__device__ __forceinline__ void smem_red_add(uint32_t* p, uint32_t v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t smem_atom_add(uint32_t* p, uint32_t v) {
    uint32_t r;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(r) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
    return r;
}
__device__ __forceinline__ void msm_bin_hist_add(uint32_t* sh, uint32_t bin, bool valid) {
    if (valid) smem_red_add(&sh[bin], 1u);
}
__device__ __forceinline__ uint32_t msm_bin_cursor_take(uint32_t* sh, uint32_t bin, bool valid) {
    return valid ? smem_atom_add(&sh[bin], 1u) : 0u;
}

// digits[j * n + i] = packed digit j of scalar i; coarse_totals[b] += digits of this tile falling into coarse bin b
__global__ void __launch_bounds__(MSM_BIN_THREADS)
msm_digits_kernel(MsmGeom g, const Fr* __restrict__ scalars, uint32_t* __restrict__ digits, uint32_t* __restrict__ coarse_totals,
                  uint32_t lb, uint32_t ncoarse) {
    __shared__ uint32_t sh[MSM_BIN_MAX_COARSE];
    for (uint32_t b = threadIdx.x; b < ncoarse; b += MSM_BIN_THREADS) sh[b] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * MSM_BIN_TILE;
    // four scalars are fetched before the first is decomposed: one 32-byte load in flight per thread would leave the kernel
    // waiting on DRAM latency (whole warps stay together: the trip counts are uniform)
    constexpr uint32_t G = 4;
    for (uint32_t t0 = threadIdx.x; t0 < MSM_BIN_TILE; t0 += G * MSM_BIN_THREADS) {
        Fr raw[G];
#pragma unroll
        for (uint32_t u = 0; u < G; u++) {
            const uint32_t i = base + t0 + u * MSM_BIN_THREADS;
            raw[u] = i < g.n ? fe_load(scalars + i) : Fr::zero();
        }
#pragma unroll
        for (uint32_t u = 0; u < G; u++) {
            const uint32_t i = base + t0 + u * MSM_BIN_THREADS;
            const bool in = i < g.n;
            Fr canon = fe_from_mont(raw[u]);
            msm_signed_digits(canon, g.c, g.windows, [&](uint32_t j, int32_t d) {
                uint32_t word = 0, bin = 0;
                if (d != 0) {
                    const uint32_t mag = d < 0 ? (uint32_t)(-d) : (uint32_t)d;
                    const uint32_t key = (g.pre_stride ? 0u : j * g.nb) + (mag - 1u);
                    word = (key + 1u) | (d < 0 ? 0x80000000u : 0u);
                    bin = key >> lb;
                }
                if (in) digits[(size_t)j * g.n + i] = word;
                msm_bin_hist_add(sh, bin, word != 0);
            });
        }
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < ncoarse; b += MSM_BIN_THREADS)
        if (sh[b]) atomicAdd(coarse_totals + b, sh[b]);
}

// coarse_cursor starts as the exclusive scan of the totals; on return rec_key / rec_val hold every non-zero digit grouped by
// coarse bin: rec_key = key & (2^lb - 1), rec_val = the entry word the accumulation reads (base index | sign << 31).
// `tile` scalars per block (msm_part_tile): their <= tile * windows records are sorted by coarse bin in shared memory first.
__global__ void __launch_bounds__(MSM_PART_THREADS)
msm_partition_kernel(MsmGeom g, const uint32_t* __restrict__ digits, uint32_t* __restrict__ coarse_cursor,
                     unsigned short* __restrict__ rec_key, uint32_t* __restrict__ rec_val, uint32_t lb, uint32_t ncoarse,
                     uint32_t tile) {
    extern __shared__ uint32_t stage[];
    uint32_t* sval = stage;                                                         // MSM_PART_TILE * MSM_PART_MAX_WINDOWS words
    unsigned short* skey = reinterpret_cast<unsigned short*>(stage + MSM_PART_TILE * MSM_PART_MAX_WINDOWS);
    __shared__ uint32_t cur[MSM_BIN_MAX_COARSE];          // counts -> local cursors
    __shared__ uint32_t lstart[MSM_BIN_MAX_COARSE + 1];   // local start of every bin's run
    __shared__ uint32_t gstart[MSM_BIN_MAX_COARSE];       // where the run goes in rec_key / rec_val
    __shared__ uint32_t warp_sums[MSM_PART_THREADS / 32];
    for (uint32_t b = threadIdx.x; b < MSM_BIN_MAX_COARSE; b += MSM_PART_THREADS) cur[b] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * tile;
    const uint32_t cnt = g.n - base < tile ? g.n - base : tile;
    // a tile is at most MSM_PART_TILE = R * MSM_PART_THREADS scalars: the R words a thread owns in one window are fetched
    // together (memory-level parallelism; a single load per round leaves the kernel latency bound)
    constexpr uint32_t R = MSM_PART_TILE / MSM_PART_THREADS;
    for (uint32_t j = 0; j < g.windows; j++) {
        uint32_t w[R];
#pragma unroll
        for (uint32_t r = 0; r < R; r++) {
            const uint32_t t = r * MSM_PART_THREADS + threadIdx.x;
            w[r] = t < cnt ? __ldg(digits + (size_t)j * g.n + base + t) : 0u;
        }
#pragma unroll
        for (uint32_t r = 0; r < R; r++) msm_bin_hist_add(cur, w[r] ? ((w[r] & 0x7fffffffu) - 1u) >> lb : 0u, w[r] != 0);
    }
    __syncthreads();
    {   // exclusive scan of the MSM_BIN_MAX_COARSE counters (two per thread), global reservation of every non-empty run
        constexpr uint32_t per = MSM_BIN_MAX_COARSE / MSM_PART_THREADS;
        const uint32_t b0 = threadIdx.x * per;
        uint32_t c[per], mine = 0;
        for (uint32_t k = 0; k < per; k++) { c[k] = cur[b0 + k]; mine += c[k]; }
        uint32_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= (unsigned)o) incl += v;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        uint32_t run = incl - mine;
        for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) run += warp_sums[w];
        for (uint32_t k = 0; k < per; k++) {
            lstart[b0 + k] = run;
            cur[b0 + k] = run;
            gstart[b0 + k] = (c[k] && b0 + k < ncoarse) ? atomicAdd(coarse_cursor + b0 + k, c[k]) : 0u;
            run += c[k];
        }
        if (threadIdx.x == MSM_PART_THREADS - 1) lstart[MSM_BIN_MAX_COARSE] = run;
    }
    __syncthreads();
    const uint32_t low_mask = (1u << lb) - 1u;
    for (uint32_t j = 0; j < g.windows; j++) {
        uint32_t w[R];
#pragma unroll
        for (uint32_t r = 0; r < R; r++) {
            const uint32_t t = r * MSM_PART_THREADS + threadIdx.x;
            w[r] = t < cnt ? __ldg(digits + (size_t)j * g.n + base + t) : 0u;
        }
#pragma unroll
        for (uint32_t r = 0; r < R; r++) {
            const uint32_t t = r * MSM_PART_THREADS + threadIdx.x;
            const uint32_t key = w[r] ? (w[r] & 0x7fffffffu) - 1u : 0u;
            const uint32_t pos = msm_bin_cursor_take(cur, key >> lb, w[r] != 0);
            if (w[r]) {
                const uint32_t idx = base + t + g.idx_offset;
                skey[pos] = (unsigned short)(key & low_mask);
                sval[pos] = (g.pre_stride ? j * g.pre_stride + idx : idx) | (w[r] & 0x80000000u);
            }
        }
    }
    __syncthreads();
    // copy-out: a warp takes one bin's run at a time -> contiguous global stores
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    for (uint32_t b = warp; b < ncoarse; b += MSM_PART_THREADS / 32) {
        const uint32_t ls = lstart[b], le = lstart[b + 1], gs = gstart[b];
        for (uint32_t p = ls + lane; p < le; p += 32) {
            rec_val[gs + (p - ls)] = sval[p];
            rec_key[gs + (p - ls)] = skey[p];
        }
    }
}

// `slices` blocks per coarse bin b (blockIdx.x = b * slices + slice): every block histograms ALL records
// [coarse_start[b], coarse_start[b + 1]) of its bin -- that gives the offsets of every key -- and then owns the keys of its slice:
// it writes their offsets / ends, counts their populations into size_hist (reversed bins, as msm_size_hist_kernel), and places
// their records.  Placement goes through a shared-memory image of the slice's output range (`cap` words), copied out with
// coalesced stores: 4-byte stores scattered straight to global memory retire at ~0.3 sectors / clock / SM and were the bound of
// both the global-atomic scatter and the first version of this kernel.  Records beyond `cap` (skewed columns) are stored directly.
__global__ void __launch_bounds__(MSM_PART_THREADS)
msm_bin_kernel(const uint32_t* __restrict__ coarse_start, const unsigned short* __restrict__ rec_key,
               const uint32_t* __restrict__ rec_val, uint32_t nkeys, uint32_t lb, uint32_t slices, uint32_t cap,
               uint32_t* __restrict__ offsets, uint32_t* __restrict__ ends, uint32_t* __restrict__ entries,
               uint32_t* __restrict__ size_hist) {
    extern __shared__ uint32_t dyn[];
    uint32_t* cnt = dyn;                              // 2^lb counters, then cursors
    uint32_t* sz = dyn + (1u << lb);                  // MSM_SIZE_BINS
    uint32_t* stage = sz + MSM_SIZE_BINS;             // cap words
    __shared__ uint32_t warp_sums[MSM_PART_THREADS / 32];
    __shared__ uint32_t slice_range[2];
    const uint32_t nlow = 1u << lb;
    const uint32_t bin = blockIdx.x / slices, slice = blockIdx.x % slices;
    const uint32_t klo = slice * (nlow / slices), khi = klo + nlow / slices;
    for (uint32_t k = threadIdx.x; k < nlow; k += MSM_PART_THREADS) cnt[k] = 0;
    for (uint32_t k = threadIdx.x; k < MSM_SIZE_BINS; k += MSM_PART_THREADS) sz[k] = 0;
    __syncthreads();
    const uint32_t s = coarse_start[bin], e = coarse_start[bin + 1];
    // U records per thread are fetched before the first is binned (one load in flight per thread = DRAM-latency bound)
    constexpr uint32_t U = 8;
    const uint32_t rounds = (e - s + U * MSM_PART_THREADS - 1) / (U * MSM_PART_THREADS);
    for (uint32_t r = 0; r < rounds; r++) {
        uint32_t k[U];
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
            const uint32_t i = s + (r * U + u) * MSM_PART_THREADS + threadIdx.x;
            k[u] = i < e ? (uint32_t)__ldg(rec_key + i) : 0xffffffffu;
        }
#pragma unroll
        for (uint32_t u = 0; u < U; u++) msm_bin_hist_add(cnt, k[u] & (nlow - 1u), k[u] < khi);     // keys above the slice are not needed
    }
    __syncthreads();
    // exclusive scan of the counters: each thread owns `per` consecutive ones
    const uint32_t per = nlow / MSM_PART_THREADS, k0 = threadIdx.x * per;
    uint32_t mine = 0;
    for (uint32_t k = 0; k < per; k++) mine += cnt[k0 + k];
    uint32_t incl = mine;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= (unsigned)o) incl += v;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t run = s + incl - mine;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) run += warp_sums[w];
    const uint32_t key0 = (bin << lb) + k0;
    for (uint32_t k = 0; k < per; k++) {
        const uint32_t c = cnt[k0 + k], low = k0 + k;
        if (low == klo) slice_range[0] = run;
        if (low >= klo && low < khi && key0 + k < nkeys) {
            offsets[key0 + k] = run;
            ends[key0 + k] = run + c;
            const uint32_t pop = c < MSM_SIZE_BINS - 1 ? c : MSM_SIZE_BINS - 1;
            smem_red_add(&sz[MSM_SIZE_BINS - 1 - pop], 1u);
        }
        cnt[k0 + k] = run;
        run += c;
        if (low + 1 == khi) slice_range[1] = run;
    }
    __syncthreads();
    const uint32_t out0 = slice_range[0], out1 = slice_range[1];
    for (uint32_t r = 0; r < rounds; r++) {
        uint32_t k[U], v[U];
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
            const uint32_t i = s + (r * U + u) * MSM_PART_THREADS + threadIdx.x;
            k[u] = i < e ? (uint32_t)__ldg(rec_key + i) : 0xffffffffu;
            v[u] = i < e ? __ldg(rec_val + i) : 0u;
        }
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
            const bool in = k[u] >= klo && k[u] < khi;
            const uint32_t pos = msm_bin_cursor_take(cnt, k[u] & (nlow - 1u), in);
            if (in) {
                if (pos - out0 < cap) stage[pos - out0] = v[u];
                else entries[pos] = v[u];
            }
        }
    }
    __syncthreads();
    const uint32_t staged = out1 - out0 < cap ? out1 - out0 : cap;
    for (uint32_t p = threadIdx.x; p < staged; p += MSM_PART_THREADS) entries[out0 + p] = stage[p];
    for (uint32_t k = threadIdx.x; k < MSM_SIZE_BINS; k += MSM_PART_THREADS)
        if (sz[k]) atomicAdd(size_hist + k, sz[k]);
}
#endif
}  // namespace h2b
