// lookup_kernels.cu -- see lookup_kernels.hpp.
#include <cub/cub.cuh>

#include "lookup_kernels.hpp"

namespace h2b {
namespace {

struct alignas(16) Key { uint64_t l[4]; };      // canonical value, little-endian 64-bit limbs (`Fr: Ord` compares these as one integer)

__device__ __forceinline__ bool key_ne(const Key& a, const Key& b) {
    return ((a.l[0] ^ b.l[0]) | (a.l[1] ^ b.l[1]) | (a.l[2] ^ b.l[2]) | (a.l[3] ^ b.l[3])) != 0;
}

// keys[0..u) = canonical input values, keys[u..2u) = canonical table values; perm = identity.
// bits[0..4) |= every key limb, bits[4..8) &= every key limb: a bit that differs between two keys is set in OR and clear in AND,
// which tells the host which limbs (and which bit range of them) the radix passes have to look at.
__global__ void __launch_bounds__(256)
lk_canon_kernel(const Fr* __restrict__ in, const Fr* __restrict__ tab, uint32_t u, Key* __restrict__ keys,
                uint32_t* __restrict__ perm, unsigned long long* __restrict__ bits) {
    __shared__ uint32_t sh_or[8], sh_and[8];
    if (threadIdx.x < 8) { sh_or[threadIdx.x] = 0u; sh_and[threadIdx.x] = 0xffffffffu; }
    __syncthreads();
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < 2 * u;
    Fr c = Fr::zero();
    if (live) {
        c = fe_from_mont(fe_load(i < u ? in + i : tab + (i - u)));
        Key k;
#pragma unroll
        for (int j = 0; j < 4; j++) k.l[j] = (uint64_t)c.l[2 * j] | ((uint64_t)c.l[2 * j + 1] << 32);
        keys[i] = k;
        perm[i] = i;
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint32_t o = __reduce_or_sync(0xffffffffu, live ? c.l[j] : 0u);
        uint32_t a = __reduce_and_sync(0xffffffffu, live ? c.l[j] : 0xffffffffu);
        if ((threadIdx.x & 31) == 0) { atomicOr(&sh_or[j], o); atomicAnd(&sh_and[j], a); }
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        atomicOr(&bits[threadIdx.x], (unsigned long long)sh_or[2 * threadIdx.x] | ((unsigned long long)sh_or[2 * threadIdx.x + 1] << 32));
        atomicAnd(&bits[4 + threadIdx.x], (unsigned long long)sh_and[2 * threadIdx.x] | ((unsigned long long)sh_and[2 * threadIdx.x + 1] << 32));
    }
}

// After sorting on limb `t` alone: two neighbours that agree on that limb but differ elsewhere mean the lower limbs decide an
// order somewhere, and the full least-significant-limb-first sequence has to run.
__global__ void lk_tie_check_kernel(const Key* __restrict__ keys, const uint32_t* __restrict__ perm, uint32_t m, int t,
                                    uint32_t* __restrict__ ties) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 || i >= m) return;
    Key a = keys[perm[i]], b = keys[perm[i - 1]];
    if (a.l[t] == b.l[t] && key_ne(a, b)) *ties = 1u;
}

__global__ void lk_gather_limb_kernel(const Key* __restrict__ keys, const uint32_t* __restrict__ perm, uint32_t m, int limb,
                                      uint64_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) out[i] = keys[perm[i]].l[limb];
}

// flag[i] = 1 where a new value starts in sorted order
__global__ void lk_flags_kernel(const Key* __restrict__ keys, const uint32_t* __restrict__ perm, uint32_t m, uint32_t* __restrict__ flag) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    flag[i] = (i == 0 || key_ne(keys[perm[i]], keys[perm[i - 1]])) ? 1u : 0u;
}

// dense rank g = gid1 - 1 of every element: the value of each rank, the rank of every input row, the multiplicity of every rank
// among the table rows
__global__ void lk_rank_kernel(const Key* __restrict__ keys, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ flag,
                               const uint32_t* __restrict__ gid1, uint32_t m, uint32_t u, uint32_t* __restrict__ rank_in,
                               uint32_t* __restrict__ counts, Key* __restrict__ values) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    uint32_t g = gid1[i] - 1, src = perm[i];
    if (flag[i]) values[g] = keys[src];
    if (src < u) rank_in[src] = g;
    else atomicAdd(counts + g, 1u);
}

// first occurrence of a value in the sorted input takes one copy of it out of the table multiset (lookup/prover.rs:432-446)
__global__ void lk_first_kernel(const uint32_t* __restrict__ sorted_in, uint32_t u, uint32_t* __restrict__ notfirst,
                                uint32_t* __restrict__ counts, uint32_t* __restrict__ status) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= u) return;
    bool first = i == 0 || sorted_in[i] != sorted_in[i - 1];
    notfirst[i] = first ? 0u : 1u;
    if (first) {
        uint32_t g = sorted_in[i];                      // one thread per distinct g: no race on counts[g]
        if (counts[g] == 0) atomicOr(status, 1u);
        else counts[g] -= 1;
    }
}

__device__ __forceinline__ Fr key_to_mont(const Key& k) {
    Fr c;
#pragma unroll
    for (int j = 0; j < 4; j++) { c.l[2 * j] = (uint32_t)k.l[j]; c.l[2 * j + 1] = (uint32_t)(k.l[j] >> 32); }
    return fe_to_mont(c);
}

// out_input = sorted input; out_table = the same value on first occurrences, otherwise the left-over table values in ascending order
// handed out from the LAST repeated row backwards (`repeated_input_rows.pop()`, lookup/prover.rs:455-459); rows >= u are zeroed.
__global__ void lk_emit_kernel(const Key* __restrict__ values, const uint32_t* __restrict__ sorted_in, const uint32_t* __restrict__ notfirst,
                               const uint32_t* __restrict__ rep_pos, const uint32_t* __restrict__ counts, const uint32_t* __restrict__ offs,
                               uint32_t m, uint32_t u, uint64_t n, Fr* __restrict__ out_in, Fr* __restrict__ out_tab,
                               uint32_t* __restrict__ status) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i >= u) { fe_store(out_in + i, Fr::zero()); fe_store(out_tab + i, Fr::zero()); return; }
    const uint32_t repeated = rep_pos[u - 1] + notfirst[u - 1], leftover = offs[m - 1] + counts[m - 1];
    if (i == 0 && repeated != leftover) atomicOr(status, 2u);
    uint32_t g_in = sorted_in[i], g_tab = g_in;
    if (notfirst[i]) {
        uint32_t p = repeated - 1 - rep_pos[i];
        if (p < leftover) {
            uint32_t lo = 0, hi = m;                      // first index with offs > p
            while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (offs[mid] <= p) lo = mid + 1; else hi = mid; }
            g_tab = lo - 1;
        }
    }
    fe_store(out_in + i, key_to_mont(values[g_in]));
    fe_store(out_tab + i, key_to_mont(values[g_tab]));
}

struct Layout {
    size_t keys, permA, permB, keyA, keyB, flag, gid1, rank_in, sorted_in, counts, offs, values, notfirst, rep_pos, status, bits, cub, total;
    size_t cub_bytes;
};

size_t cub_temp_bytes(uint32_t u) {
    uint32_t m = 2 * u;
    size_t best = 0, t = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)m);
    best = t > best ? t : best;
    cub::DeviceRadixSort::SortKeys(nullptr, t, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)u);
    best = t > best ? t : best;
    cub::DeviceScan::InclusiveSum(nullptr, t, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)m);
    best = t > best ? t : best;
    cub::DeviceScan::ExclusiveSum(nullptr, t, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)m);
    best = t > best ? t : best;
    return best;
}

Layout layout(uint32_t u) {
    Layout L;
    size_t m = 2 * (size_t)u, off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    L.keys = take(m * 32); L.values = take(m * 32);
    L.keyA = take(m * 8); L.keyB = take(m * 8);
    L.permA = take(m * 4); L.permB = take(m * 4);
    L.flag = take(m * 4); L.gid1 = take(m * 4); L.counts = take(m * 4); L.offs = take(m * 4);
    L.rank_in = take((size_t)u * 4); L.sorted_in = take((size_t)u * 4); L.notfirst = take((size_t)u * 4); L.rep_pos = take((size_t)u * 4);
    L.status = take(256);
    L.bits = take(256);                 // 4 x OR, 4 x AND of the key limbs, then the tie flag
    L.cub_bytes = cub_temp_bytes(u);
    L.cub = take(L.cub_bytes);
    L.total = off;
    return L;
}

}  // namespace

size_t lookup_permute_scratch_bytes(uint32_t usable) { return usable ? layout(usable).total : 256; }

cudaError_t lookup_permute_run(cudaStream_t stream, const Fr* d_input, const Fr* d_table, uint32_t u, uint64_t n, Fr* d_out_input,
                               Fr* d_out_table, void* d_scratch, size_t scratch_bytes, uint32_t** d_status_out, int* launches) {
    uint8_t* base = reinterpret_cast<uint8_t*>(d_scratch);
    int nl = 0;
    if (u == 0) {
        *d_status_out = reinterpret_cast<uint32_t*>(base);
        cudaError_t e = cudaMemsetAsync(base, 0, 4, stream);
        if (e != cudaSuccess) return e;
        if (n) { e = cudaMemsetAsync(d_out_input, 0, n * 32, stream); if (e != cudaSuccess) return e; e = cudaMemsetAsync(d_out_table, 0, n * 32, stream); }
        *launches = 0;
        return e;
    }
    Layout L = layout(u);
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    const uint32_t m = 2 * u;
    Key* keys = reinterpret_cast<Key*>(base + L.keys);
    Key* values = reinterpret_cast<Key*>(base + L.values);
    uint64_t* key_lo[2] = {reinterpret_cast<uint64_t*>(base + L.keyA), reinterpret_cast<uint64_t*>(base + L.keyB)};
    uint32_t* perm[2] = {reinterpret_cast<uint32_t*>(base + L.permA), reinterpret_cast<uint32_t*>(base + L.permB)};
    uint32_t* flag = reinterpret_cast<uint32_t*>(base + L.flag);
    uint32_t* gid1 = reinterpret_cast<uint32_t*>(base + L.gid1);
    uint32_t* counts = reinterpret_cast<uint32_t*>(base + L.counts);
    uint32_t* offs = reinterpret_cast<uint32_t*>(base + L.offs);
    uint32_t* rank_in = reinterpret_cast<uint32_t*>(base + L.rank_in);
    uint32_t* sorted_in = reinterpret_cast<uint32_t*>(base + L.sorted_in);
    uint32_t* notfirst = reinterpret_cast<uint32_t*>(base + L.notfirst);
    uint32_t* rep_pos = reinterpret_cast<uint32_t*>(base + L.rep_pos);
    uint32_t* status = reinterpret_cast<uint32_t*>(base + L.status);
    unsigned long long* bits = reinterpret_cast<unsigned long long*>(base + L.bits);
    void* tmp = base + L.cub;
    size_t tmp_bytes = L.cub_bytes;
    *d_status_out = status;
    cudaError_t e;
#define LK_CK(x) do { e = (x); if (e != cudaSuccess) return e; } while (0)
    const unsigned T = 256, gm = (m + T - 1) / T, gu = (u + T - 1) / T;
    LK_CK(cudaMemsetAsync(status, 0, 4, stream));
    LK_CK(cudaMemsetAsync(counts, 0, (size_t)m * 4, stream));
    LK_CK(cudaMemsetAsync(bits, 0, 32, stream));
    LK_CK(cudaMemsetAsync(bits + 4, 0xff, 32, stream));
    LK_CK(cudaMemsetAsync(bits + 8, 0, 8, stream));
    lk_canon_kernel<<<gm, T, 0, stream>>>(d_input, d_table, u, keys, perm[0], bits); nl++;
    // Which bits differ between any two keys?  Circuit columns are small integers (one low limb varies) or compressed, random
    // looking elements (the top limb separates them), so most of the 4 x 8 radix passes of a blind 254-bit sort look at constant
    // or irrelevant bits.  The plan needs the masks on the host: one 64-byte read-back and a stream synchronisation.
    unsigned long long h_bits[8];
    LK_CK(cudaMemcpyAsync(h_bits, bits, sizeof(h_bits), cudaMemcpyDeviceToHost, stream));
    LK_CK(cudaStreamSynchronize(stream));
    int vary[4], nvary = 0, lo_bit[4], hi_bit[4];
    for (int limb = 0; limb < 4; limb++) {
        unsigned long long v = h_bits[limb] & ~h_bits[4 + limb];
        if (!v) continue;
        lo_bit[limb] = __builtin_ctzll(v);
        hi_bit[limb] = 64 - __builtin_clzll(v);
        vary[nvary++] = limb;
    }
    int cur = 0;
    auto sort_on = [&](int limb) -> cudaError_t {       // stable pass on the varying bit range of one limb
        lk_gather_limb_kernel<<<gm, T, 0, stream>>>(keys, perm[cur], m, limb, key_lo[0]); nl++;
        size_t tb = tmp_bytes;
        cudaError_t r = cub::DeviceRadixSort::SortPairs(tmp, tb, (const uint64_t*)key_lo[0], key_lo[1], (const uint32_t*)perm[cur], perm[cur ^ 1],
                                                        (int)m, lo_bit[limb], hi_bit[limb], stream);
        nl += 2 + (hi_bit[limb] - lo_bit[limb] + 7) / 8;
        cur ^= 1;
        return r;
    };
    bool full = nvary > 1;
    if (nvary > 1) {
        // optimistic: the most significant varying limb alone; exact unless two keys tie on it and differ below
        const int t = vary[nvary - 1];
        LK_CK(sort_on(t));
        lk_tie_check_kernel<<<gm, T, 0, stream>>>(keys, perm[cur], m, t, reinterpret_cast<uint32_t*>(bits + 8)); nl++;
        uint32_t ties = 0;
        LK_CK(cudaMemcpyAsync(&ties, bits + 8, 4, cudaMemcpyDeviceToHost, stream));
        LK_CK(cudaStreamSynchronize(stream));
        full = ties != 0;
    }
    if (nvary == 1) LK_CK(sort_on(vary[0]));
    if (full)                                            // least significant limb first; every pass is stable
        for (int q = 0; q < nvary; q++) LK_CK(sort_on(vary[q]));
    lk_flags_kernel<<<gm, T, 0, stream>>>(keys, perm[cur], m, flag); nl++;
    { size_t tb = tmp_bytes; LK_CK(cub::DeviceScan::InclusiveSum(tmp, tb, (const uint32_t*)flag, gid1, (int)m, stream)); nl += 2; }
    int rank_bits = 1;
    while (rank_bits < 32 && (m >> rank_bits)) rank_bits++;       // ranks are < m
    lk_rank_kernel<<<gm, T, 0, stream>>>(keys, perm[cur], flag, gid1, m, u, rank_in, counts, values); nl++;
    { size_t tb = tmp_bytes; LK_CK(cub::DeviceRadixSort::SortKeys(tmp, tb, (const uint32_t*)rank_in, sorted_in, (int)u, 0, rank_bits, stream)); nl += 4; }
    lk_first_kernel<<<gu, T, 0, stream>>>(sorted_in, u, notfirst, counts, status); nl++;
    { size_t tb = tmp_bytes; LK_CK(cub::DeviceScan::ExclusiveSum(tmp, tb, (const uint32_t*)notfirst, rep_pos, (int)u, stream)); nl += 2; }
    { size_t tb = tmp_bytes; LK_CK(cub::DeviceScan::ExclusiveSum(tmp, tb, (const uint32_t*)counts, offs, (int)m, stream)); nl += 2; }
    lk_emit_kernel<<<(unsigned)((n + T - 1) / T), T, 0, stream>>>(values, sorted_in, notfirst, rep_pos, counts, offs, m, u, n, d_out_input,
                                                                  d_out_table, status); nl++;
    LK_CK(cudaGetLastError());
#undef LK_CK
    *launches = nl;
    return cudaSuccess;
}
}  // namespace h2b
