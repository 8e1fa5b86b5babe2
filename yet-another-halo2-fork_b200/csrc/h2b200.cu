// h2b200.cu -- libh2b200.so: context, resident data, kernel orchestration and the C ABI (include/h2b200.h).
//
// One `h2b_ctx` owns, per device: a compute stream, grow-only scratch buffers, resident MSM bases and cached
// NTT twiddle tables.  Every entry point enqueues its whole pipeline on that stream with no host round trip
// in the middle, then synchronises once and finishes on the host (window combine / normalisation).
#include <cuda_runtime.h>

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/h2b200.h"
#include "hostcopy.hpp"
#define H2B_NO_KERNELS 1      // kernels live in msm_kernels.cu / ntt_kernels_*.cu; this unit is host orchestration only
#include "msm.cuh"
#include "msm_kernels.hpp"
#include "msm_bin.cuh"
#include "msm_ba_kernels.hpp"
#include "poly_kernels.hpp"
#include "quotient_kernels.hpp"
#include "setup_kernels.hpp"
#include "lookup_kernels.hpp"
#include "ntt.cuh"
#include "ntt_plan.hpp"
#include "ntt_kernels.hpp"

using namespace h2b;

namespace {

// ------------------------------------------------------------------------------------------------
// small utilities
// ------------------------------------------------------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = (bytes + 255) & ~(size_t)255;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct ProfRec { const char* name; cudaEvent_t a, b; };

struct NttPlanDev {
    NttPlan plan;
    Fr* tables = nullptr;
    bool has_scale = false;
    Fr scale;
};

// Outputs of the binning stage of one MSM chunk (count / scan / scatter / order).  Two sets, so that chunk k+1 is binned
// on the sort stream (memory / atomic bound) while chunk k is accumulated on the compute stream (multiplier bound).
struct MsmSortSet {
    DevBuf counts, offsets, ends, entries, order, blocksums, sizehist, long_info, long_tasks, counters;
    DevBuf canon;                           // canonical scalars of the chunk (count -> scatter)
    DevBuf rec_key, rec_val, coarse;        // two-level binning (msm_bin.cuh): grouped records, coarse-bin totals / cursors / starts
    cudaEvent_t sorted = nullptr, consumed = nullptr;
    bool consumed_pending = false;
    void release() {
        DevBuf* bufs[] = {&counts, &offsets, &ends, &entries, &order, &blocksums, &sizehist, &long_info, &long_tasks, &counters,
                          &canon, &rec_key, &rec_val, &coarse};
        for (DevBuf* b : bufs) b->release();
        if (sorted) cudaEventDestroy(sorted);
        if (consumed) cudaEventDestroy(consumed);
        sorted = consumed = nullptr;
    }
};

struct Dev {
    int id = 0;
    int sms = 148;
    cudaStream_t stream = nullptr, copy_stream = nullptr, sort_stream = nullptr;
    cudaStream_t launch_stream = nullptr;         // where H2B_LAUNCH enqueues (stream, or sort_stream inside the binning stage)
    cudaEvent_t ev_first = nullptr, ev_last = nullptr;
    std::vector<cudaEvent_t> chunk_events;
    DevBuf ntt_a, ntt_b, ntt_tab;
    MsmSortSet ms[2];
    DevBuf m_scalars, m_buckets, m_wpart, m_long_partials, m_counters, m_tmp_bases, m_ba;
    StageRing ring;                               // pinned slots for pageable caller buffers (hostcopy.hpp)
    DevBuf q_stage, q_scratch;                    // quotient evaluation: uploaded program / pointer tables, intermediates
    DevBuf p_in, p_out, p_tiles;                  // polynomial helpers (poly.cuh): staged operands and per-tile summaries
    void* pinned = nullptr;
    size_t pinned_cap = 0;
    std::map<std::string, NttPlanDev> plans;
    bool prof = false;
    bool peer_attr = false;
    bool bin_attr = false;             // function attributes are per device: the tiled binning kernels' shared-memory opt-in was made here
    std::vector<ProfRec> recs;
    std::vector<cudaEvent_t> ev_pool;
    uint64_t launches = 0;

    cudaEvent_t get_event() {
        if (!ev_pool.empty()) { cudaEvent_t e = ev_pool.back(); ev_pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
    void prof_begin(const char* name) {
        if (!prof) return;
        ProfRec r; r.name = name; r.a = get_event(); r.b = get_event();
        cudaEventRecord(r.a, launch_stream);
        recs.push_back(r);
    }
    void prof_end() {
        if (!prof) return;
        cudaEventRecord(recs.back().b, launch_stream);
    }
    cudaError_t reserve_pinned(size_t bytes) {
        if (bytes <= pinned_cap) return cudaSuccess;
        if (pinned) cudaFreeHost(pinned);
        pinned = nullptr; pinned_cap = 0;
        cudaError_t e = cudaMallocHost(&pinned, bytes);
        if (e == cudaSuccess) pinned_cap = bytes;
        return e;
    }
};

struct BasesShard {
    void* d = nullptr;            // count affine points as registered
    size_t start = 0, count = 0;
    bool owned = true;
    void* pre = nullptr;          // optional: pre_w * count precomputed multiples 2^(pre_c*j) * P_i (window-major)
    uint32_t pre_c = 0, pre_w = 0;
    bool pre_skipped = false;     // the multiples did not fit: see h2b_last_notice
};
struct Bases { size_t n = 0; std::vector<BasesShard> shard; bool pre_skipped = false; };

#ifndef H2B_ACC_DEFAULT_VARIANT
#define H2B_ACC_DEFAULT_VARIANT 1
#endif
// Tuning knobs of one context.  The H2B_* environment variables are read ONCE, in h2b_init (getenv is not safe against a
// concurrent setenv, and a product path should not depend on the environment at call time); afterwards a knob changes only
// through h2b_set_option, under the context lock.
struct Tuning {
    int msm_c = 0;                 // H2B_MSM_C          window bits of the per-window layout (0 = auto)
    int msm_pre_c = 0;             // H2B_MSM_PRE_C      window bits of the precomputed-multiples layout (0 = auto)
    int msm_precompute = 1;        // H2B_MSM_PRECOMPUTE 0: registered bases keep only the 64 B per point
    int msm_chunks = 0;            // H2B_MSM_CHUNKS     number of scalar chunks of the streamed pipeline (0 = auto)
    int msm_no_ramp = 0;           // H2B_MSM_NO_RAMP    1: equal chunks also for host scalars
    int msm_no_batch = 0;          // H2B_MSM_NO_BATCH   1: h2b_msm_batch loops over single MSMs
    int msm_reduce_classic = 0;    // H2B_MSM_REDUCE_CLASSIC 1: running-sum reduction for every bucket-set size
    int msm_acc_variant = H2B_ACC_DEFAULT_VARIANT;   // H2B_ACC_VARIANT
    int msm_canon_once = 1;        // H2B_MSM_CANON_ONCE 1: the count kernel keeps the canonical scalars for the scatter kernel (one
                                   //                    Montgomery -> canonical multiplication per scalar instead of two)
    int msm_ba = 0;                // H2B_MSM_BA         rounds of batched affine additions in front of the mixed additions
                                   //                    (msm_ba.cuh): 0 = none (msm_accumulate_kernel), 1, 2
    int msm_ba_regs = 0;           // H2B_MSM_BA_REGS    0: 4 resident blocks per SM (128 registers), 1: 3 blocks (168)
    int msm_ba_min_mean = 8;       // H2B_MSM_BA_MIN_MEAN  mean bucket population of a chunk below which the plain kernel runs
    int msm_reduce_seg = 512;      // H2B_MSM_REDUCE_SEG buckets per row / column segment job of the lines reduction
    int msm_sort = 0;              // H2B_MSM_SORT       0: global-atomic count + scatter (faster, see msm_bin.cuh), 1: tiled two-level binning
    int msm_sort_blocks = 2;       // H2B_MSM_SORT_BLOCKS  blocks per SM of the count / scatter launches (grid-stride); 0 = one thread per scalar.
                                   //                    These kernels are bound by the L2, not by the SMs: two blocks per SM run them as fast
                                   //                    as a full grid and take fewer block slots from the accumulation beside them
                                   //                    (2^24: 36.29 -> 36.14 ms, from pageable scalars 36.6 -> 36.2 ms)
    int msm_sort_tb = 256;         // H2B_MSM_SORT_TB    threads per block of the count / scatter kernels (128 or 256)
    int sort_prio = 1;             // H2B_SORT_PRIO      priority of the binning stream against the compute stream: 1 higher,
                                   //                    0 equal, -1 lower (read when the context creates its streams)
    int ntt_full_table_max_log = 26;   // H2B_NTT_FULL_TABLE_MAX_LOG
    std::string ntt_split;         // H2B_NTT_SPLIT      e.g. "9,9,6"
    int l2_fetch = 0;              // H2B_L2_FETCH       cudaLimitMaxL2FetchGranularity hint (experiments)
    int peer_timeout_ms = 20000;   // H2B_PEER_TIMEOUT_MS how long the exchange kernel of a sharded MSM waits for a peer

    // name = the environment variable without its H2B_ prefix, lower case; value == nullptr or "" restores the default
    bool set(const std::string& name, const char* value) {
        const bool reset = !value || !*value;
        const int v = reset ? 0 : std::atoi(value);
        Tuning def;
        if (name == "msm_c") msm_c = (!reset && v >= 4 && v <= 24) ? v : def.msm_c;
        else if (name == "msm_pre_c") msm_pre_c = (!reset && v >= 4 && v <= 24) ? v : def.msm_pre_c;
        else if (name == "msm_precompute") msm_precompute = reset ? def.msm_precompute : (v != 0);
        else if (name == "msm_chunks") msm_chunks = (!reset && v >= 1 && v <= 64) ? v : def.msm_chunks;
        else if (name == "msm_no_ramp") msm_no_ramp = reset ? 0 : (v != 0);
        else if (name == "msm_no_batch") msm_no_batch = reset ? 0 : (v != 0);
        else if (name == "msm_reduce_classic") msm_reduce_classic = reset ? 0 : (v != 0);
        else if (name == "acc_variant") msm_acc_variant = (!reset && v >= 0 && v < 4) ? v : def.msm_acc_variant;
        else if (name == "msm_canon_once") msm_canon_once = reset ? def.msm_canon_once : (v != 0);
        else if (name == "msm_ba") msm_ba = (!reset && v >= 0 && v <= 2) ? v : def.msm_ba;
        else if (name == "msm_ba_regs") msm_ba_regs = (!reset && v >= 0 && v <= 1) ? v : def.msm_ba_regs;
        else if (name == "msm_ba_min_mean") msm_ba_min_mean = (!reset && v >= 1) ? v : def.msm_ba_min_mean;
        else if (name == "msm_reduce_seg") msm_reduce_seg = (!reset && v >= 32 && v <= 4096 && (v & (v - 1)) == 0) ? v : def.msm_reduce_seg;
        else if (name == "msm_sort") msm_sort = reset ? def.msm_sort : (v != 0);
        else if (name == "msm_sort_blocks") msm_sort_blocks = (!reset && v >= 0 && v <= 64) ? v : def.msm_sort_blocks;
        else if (name == "msm_sort_tb") msm_sort_tb = (!reset && (v == 64 || v == 128 || v == 256 || v == 512)) ? v : def.msm_sort_tb;
        else if (name == "sort_prio") sort_prio = (!reset && v >= -1 && v <= 1) ? v : def.sort_prio;
        else if (name == "ntt_full_table_max_log") ntt_full_table_max_log = reset ? def.ntt_full_table_max_log : v;
        else if (name == "ntt_split") ntt_split = reset ? std::string() : std::string(value);
        else if (name == "l2_fetch") l2_fetch = (v == 32 || v == 64 || v == 128) ? v : 0;
        else if (name == "peer_timeout_ms") peer_timeout_ms = (!reset && v >= 1) ? v : def.peer_timeout_ms;
        else return false;
        return true;
    }
    void from_environment() {
        static const char* names[] = {"msm_c", "msm_pre_c", "msm_precompute", "msm_chunks", "msm_no_ramp", "msm_no_batch",
                                      "msm_reduce_classic", "msm_reduce_seg", "msm_canon_once", "msm_ba", "msm_ba_regs", "msm_ba_min_mean", "acc_variant", "msm_sort", "msm_sort_blocks", "msm_sort_tb", "sort_prio", "ntt_full_table_max_log", "ntt_split", "l2_fetch", "peer_timeout_ms"};
        for (const char* nm : names) {
            std::string env = "H2B_";
            for (const char* c = nm; *c; c++) env.push_back((char)std::toupper((unsigned char)*c));
            if (const char* val = std::getenv(env.c_str())) set(nm, *val ? val : "1");
        }
    }
};

}  // namespace

// Mailboxes of a point-range sharded MSM across processes (msm.cuh 7): mine lives on device 0 of this context, the peers' are
// mapped through CUDA IPC (or handed over as plain pointers when the peer is a context of the same process).
struct PeerState {
    void* mailbox = nullptr;
    MsmPeerTable table = {};
    std::vector<void*> opened;
    int rank = -1, world = 0;
    uint64_t epoch = 0;
    uint32_t* status = nullptr;        // page-locked word the exchange kernel reports into
};

struct h2b_ctx {
    PeerState peer;
    std::vector<Dev> devs;
    std::mutex mu;
    std::string err;
    std::map<uint64_t, Bases> bases;
    uint64_t next_handle = 1;
    double last_ms = 0.0;
    Tuning tune;
    std::string status;                // last non-fatal notice (e.g. window precomputation skipped for lack of memory)
    mutable std::mutex err_mu;         // `fail` may be reached from the per-device host threads of a multi-GPU call
    int fail(int code, const std::string& msg) { std::lock_guard<std::mutex> lk(err_mu); err = msg; return code; }
};

namespace {

thread_local std::string g_init_error;
// Contexts that are alive: host languages with garbage collectors finalise objects in no particular order at exit, so a
// descriptor may be released (or a context destroyed) after its context is gone; those two calls check here first.
// Function-local statics: no destruction-order dependence on other translation units.
std::mutex& live_mu() { static std::mutex* m = new std::mutex(); return *m; }
std::set<h2b_ctx*>& live_set() { static std::set<h2b_ctx*>* s = new std::set<h2b_ctx*>(); return *s; }
bool ctx_is_live(h2b_ctx* c) { std::lock_guard<std::mutex> lk(live_mu()); return live_set().count(c) != 0; }

#define H2B_CK(expr)                                                                                         \
    do {                                                                                                     \
        cudaError_t e__ = (expr);                                                                            \
        if (e__ != cudaSuccess)                                                                              \
            return ctx->fail(e__ == cudaErrorMemoryAllocation ? H2B_E_NOMEM : H2B_E_CUDA,                     \
                             std::string(#expr) + ": " + cudaGetErrorString(e__));                           \
    } while (0)

#define H2B_LAUNCH(dev, name, kern, grid, block, smem, ...)                                                  \
    do {                                                                                                     \
        (dev).prof_begin(name);                                                                              \
        kern<<<(grid), (block), (smem), (dev).launch_stream>>>(__VA_ARGS__);                                        \
        (dev).prof_end();                                                                                    \
        (dev).launches++;                                                                                    \
    } while (0)

inline Fr fr_from_bytes(const void* p) { Fr f; std::memcpy(&f, p, 32); return f; }

// ------------------------------------------------------------------------------------------------
// NTT
// ------------------------------------------------------------------------------------------------
struct NttKernels {
    ntt_kernel_t tab[10][2];
    size_t smem[10];
    NttKernels() {
        std::memset(tab, 0, sizeof(tab));
        std::memset(smem, 0, sizeof(smem));
        h2b_ntt_register_a(tab, smem);
        h2b_ntt_register_b(tab, smem);
        h2b_ntt_register_c(tab, smem);
        h2b_ntt_register_d(tab, smem);
    }
};
const NttKernels& ntt_kernels() { static NttKernels k; return k; }

int ntt_prepare_device(h2b_ctx* ctx) {
    const NttKernels& K = ntt_kernels();
    for (int s = 1; s <= 9; s++)
        for (int r = 0; r < 2; r++)
            H2B_CK(cudaFuncSetAttribute((const void*)K.tab[s][r], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K.smem[s]));
    return H2B_OK;
}

int ntt_get_plan(h2b_ctx* ctx, Dev& d, uint32_t log_n, const Fr& omega, const Fr* scale, NttPlanDev** out) {
    std::string key((const char*)&log_n, 4);
    key.append((const char*)&omega, 32);
    if (scale) key.append((const char*)scale, 32);
    key.append(ctx->tune.ntt_split);
    key.push_back((char)ctx->tune.ntt_full_table_max_log);
    auto it = d.plans.find(key);
    if (it != d.plans.end()) { *out = &it->second; return H2B_OK; }
    NttPlanDev pd;
    pd.plan = ntt_make_plan(log_n, omega, scale, ctx->tune.ntt_split.c_str(), (uint32_t)ctx->tune.ntt_full_table_max_log);
    pd.has_scale = scale != nullptr;
    if (scale) pd.scale = *scale;
    H2B_CK(cudaMalloc((void**)&pd.tables, pd.plan.total_elems * sizeof(Fr)));
    for (const auto& seg : pd.plan.segments) {
        H2B_LAUNCH(d, "ntt_gen_powers", ntt_gen_powers_kernel, (seg.count + 127) / 128, 128, 0,
                   pd.tables + seg.offset, seg.root, seg.first, seg.count);
    }
    {
        uint32_t done = 0;
        for (int j = 0; j + 1 < pd.plan.n_pass; j++) {
            if (pd.plan.full_count[j]) {
                uint64_t cnt = pd.plan.full_count[j];
                uint32_t log_inner = log_n - done - pd.plan.log_s[j];
                const Fr* hi = pd.tables + (j == 0 ? pd.plan.off_hi_first : pd.plan.off_hi_rest);
                H2B_LAUNCH(d, "ntt_gen_powers", ntt_gen_boundary_kernel, (unsigned)((cnt + 127) / 128), 128, 0,
                           pd.tables + pd.plan.off_full[j], pd.tables + pd.plan.off_lo, hi, pd.plan.tw_h, log_inner, done, cnt);
            }
            done += pd.plan.log_s[j];
        }
    }
    H2B_CK(cudaGetLastError());
    auto ins = d.plans.emplace(key, std::move(pd));
    *out = &ins.first->second;
    return H2B_OK;
}

// Enqueue a (batched) transform: in -> out, `tmp` holds 2^log_n elements per column for the intermediate
// passes (unused for single-pass sizes).  in/tmp/out may be the same buffer only as documented at the call sites.
int ntt_enqueue(h2b_ctx* ctx, Dev& d, const Fr* in, uint64_t in_stride, Fr* tmp, uint64_t tmp_stride, Fr* out,
                uint64_t out_stride, uint32_t batch, uint32_t log_n, const Fr& omega, const Fr* scale, const NttFused& fz) {
    if (log_n == 0) {   // size-1 transform: identity (times scale / patterns are no-ops at index 0 except scale)
        return ctx->fail(H2B_E_INVALID, "ntt: log_n must be >= 1");
    }
    NttPlanDev* pd = nullptr;
    int rc = ntt_get_plan(ctx, d, log_n, omega, scale, &pd);
    if (rc) return rc;
    NttPass passes[4];
    ntt_build_passes(pd->plan, pd->tables, scale, fz, passes);
    const NttKernels& K = ntt_kernels();
    for (int j = 0; j < pd->plan.n_pass; j++) {
        bool last = j == pd->plan.n_pass - 1;
        const Fr* src = j == 0 ? in : tmp;
        uint64_t src_stride = j == 0 ? in_stride : tmp_stride;
        Fr* dst = last ? out : tmp;
        uint64_t dst_stride = last ? out_stride : tmp_stride;
        uint32_t s = passes[j].log_s;
        uint64_t tiles = ntt_pass_tiles(pd->plan, j);
        dim3 grid((unsigned)tiles, batch, 1);
        unsigned threads = (1u << s) < 32u ? 32u : (1u << s);
        ntt_kernel_t kern = K.tab[s][last ? 1 : 0];
        H2B_LAUNCH(d, last ? "ntt_row_pass" : "ntt_column_pass", kern, grid, threads, K.smem[s], passes[j], src, dst,
                   src_stride, dst_stride);
    }
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}

// ------------------------------------------------------------------------------------------------
// MSM
// ------------------------------------------------------------------------------------------------
uint32_t msm_choose_c(const h2b_ctx* ctx, size_t n_total) {
    if (ctx->tune.msm_c) return (uint32_t)ctx->tune.msm_c;   // c >= 4 keeps windows <= H2B_MSM_MAX_WINDOWS
    uint32_t lg = 0;
    while (((size_t)1 << (lg + 1)) <= n_total) lg++;
    int c = (int)lg - 4;
    if (c < 4) c = 4;
    if (c > 16) c = 16;
    return (uint32_t)c;
}

// What the reduction of one call leaves for the host: `windows` bucket sets, each reduced to `n_terms` points whose weighted
// sum sum_t 2^(exps[t]) * term_t is the set's bucket sum (msm_horner_terms); d_terms = where they sit on the device.
struct MsmLaunchInfo { uint32_t c, windows, n_terms; bool pre; uint32_t exps[MSM_MAX_TERMS]; Xyzz* d_terms; };
const MsmKernelTable& MK = msm_kernel_table();

// Enqueue the whole bucket pipeline for `n` resident (scalar, base) pairs on device d; the per-window partial
// sums (windows * slices XYZZ points) are copied asynchronously into d.pinned.
// The bucket pipeline in two stages so that a host-resident scalar vector can be streamed in chunks: stage A
// (count / scan / scatter / order / accumulate) runs per chunk and adds into the SAME bucket array; stage B (reduce +
// copy-back of the per-set sums) runs once.  pre_stride > 0: `bases` holds precomputed window multiples (see
// MsmGeom::pre_stride) and ONE bucket set is used.
struct MsmShape { MsmGeom g; uint32_t kw, nkeys; };

MsmShape msm_shape(uint32_t n, uint32_t c, uint32_t pre_stride, uint32_t idx_offset) {
    MsmShape sp;
    sp.g.c = c; sp.g.windows = msm_windows(c); sp.g.nb = 1u << (c - 1); sp.g.n = n; sp.g.pre_stride = pre_stride;
    sp.g.idx_offset = idx_offset;
    sp.kw = pre_stride ? 1u : sp.g.windows;
    sp.nkeys = sp.kw * sp.g.nb;
    return sp;
}

struct MsmChunkPlan { MsmShape sp; uint32_t long_threshold, max_long, max_tasks, nblk; };

int msm_chunk_plan(h2b_ctx* ctx, uint32_t n, uint32_t idx_offset, uint32_t c, uint32_t pre_stride, MsmChunkPlan* out) {
    MsmChunkPlan pl;
    pl.sp = msm_shape(n, c, pre_stride, idx_offset);
    const MsmGeom& g = pl.sp.g;
    const uint64_t max_entries = (uint64_t)n * g.windows;
    if (max_entries >= 0x7fffffffull || (uint64_t)(idx_offset + n) * (pre_stride ? g.windows : 1) >= 0x7fffffffull)
        return ctx->fail(H2B_E_INVALID, "msm: n * windows exceeds 2^31 (shard the MSM)");
    uint32_t mean = (uint32_t)(max_entries / pl.sp.nkeys);
    pl.long_threshold = std::max<uint32_t>(256u, 16u * std::max<uint32_t>(mean, 1u));
    pl.max_long = (uint32_t)(max_entries / pl.long_threshold) + 1;
    pl.max_tasks = (uint32_t)(max_entries / MSM_LONG_CHUNK) + pl.max_long + 1;
    pl.nblk = (pl.sp.nkeys + SCAN_BLOCK - 1) / SCAN_BLOCK;
    *out = pl;
    return H2B_OK;
}

// Batched-affine accumulation (msm_ba.cuh): how many rounds this chunk gets (0 = the plain kernel: sparse chunks have too few
// pairs per bucket to pay for the block-wide inversion), and the scratch it needs.
uint32_t msm_ba_rounds(const h2b_ctx* ctx, const MsmChunkPlan& pl) {
    if (!ctx->tune.msm_ba) return 0;
    const uint64_t mean = (uint64_t)pl.sp.g.n * pl.sp.g.windows / std::max<uint32_t>(pl.sp.nkeys, 1u);
    return mean >= (uint64_t)ctx->tune.msm_ba_min_mean ? (uint32_t)ctx->tune.msm_ba : 0u;
}
uint32_t msm_ba_stride(uint32_t nkeys) { return (nkeys + MSM_BA_THREADS - 1) / MSM_BA_THREADS * MSM_BA_THREADS; }
size_t msm_ba_bytes(uint32_t nkeys, uint32_t rounds) { return (size_t)(rounds == 2 ? 3 : 1) * MSM_BA_KCAP * msm_ba_stride(nkeys) * sizeof(Fq); }

// Binning stage of one chunk, enqueued on the SORT stream into sort set `S`: histogram, offsets, scatter, population order,
// oversized-bucket task list.
// Buffers of one sort set for a chunk of this plan (growing a buffer frees it, which waits for the device: the streamed MSM
// therefore sizes both of its sets for its largest chunk before the first one is enqueued).
int msm_sort_reserve(h2b_ctx* ctx, MsmSortSet& S, const MsmChunkPlan& pl) {
    const MsmGeom& g = pl.sp.g;
    const uint32_t nkeys = pl.sp.nkeys;
    const uint64_t max_entries = (uint64_t)g.n * g.windows;
    H2B_CK(S.counts.reserve((size_t)nkeys * 4));
    H2B_CK(S.offsets.reserve((size_t)nkeys * 4));
    H2B_CK(S.ends.reserve((size_t)nkeys * 4));
    H2B_CK(S.order.reserve((size_t)nkeys * 4));
    H2B_CK(S.entries.reserve(std::max<size_t>((size_t)max_entries * 4, 256)));
    H2B_CK(S.blocksums.reserve((size_t)pl.nblk * 4 + 256));
    H2B_CK(S.sizehist.reserve((size_t)MSM_SIZE_BINS * 4));
    H2B_CK(S.long_info.reserve((size_t)pl.max_long * sizeof(MsmLongInfo)));
    H2B_CK(S.long_tasks.reserve((size_t)pl.max_tasks * sizeof(MsmLongTask)));
    H2B_CK(S.counters.reserve(256));
    if (ctx->tune.msm_canon_once) H2B_CK(S.canon.reserve(std::max<size_t>((size_t)g.n * 32, 256)));
    if (ctx->tune.msm_sort && msm_bin_supported(nkeys)) {
        H2B_CK(S.rec_key.reserve(std::max<size_t>((size_t)max_entries * 2, 256)));
        H2B_CK(S.rec_val.reserve(std::max<size_t>((size_t)max_entries * 4, 256)));
        H2B_CK(S.coarse.reserve((size_t)(2 * MSM_BIN_MAX_COARSE + 8) * 4));
    }
    return H2B_OK;
}
int msm_stage_sort(h2b_ctx* ctx, Dev& d, MsmSortSet& S, const MsmChunkPlan& pl, const Fr* scalars) {
    const MsmGeom& g = pl.sp.g;
    const uint32_t nkeys = pl.sp.nkeys, n = g.n, nblk = pl.nblk;
    const uint64_t max_entries = (uint64_t)n * g.windows;
    { int rc0 = msm_sort_reserve(ctx, S, pl); if (rc0) return rc0; }
    uint32_t* counts = S.counts.as<uint32_t>();
    uint32_t* offsets = S.offsets.as<uint32_t>();
    uint32_t* ends = S.ends.as<uint32_t>();
    uint32_t* order = S.order.as<uint32_t>();
    uint32_t* entries = S.entries.as<uint32_t>();
    uint32_t* blocksums = S.blocksums.as<uint32_t>();
    uint32_t* sizehist = S.sizehist.as<uint32_t>();
    uint32_t* counters = S.counters.as<uint32_t>();

    d.launch_stream = d.sort_stream;
    cudaStream_t st = d.sort_stream;
    const bool tiled = ctx->tune.msm_sort && msm_bin_supported(nkeys);
    cudaError_t e1 = tiled ? cudaSuccess : cudaMemsetAsync(counts, 0, (size_t)nkeys * 4, st);
    cudaError_t e2 = cudaMemsetAsync(sizehist, 0, (size_t)MSM_SIZE_BINS * 4, st);
    cudaError_t e3 = cudaMemsetAsync(counters, 0, 256, st);
    if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) { d.launch_stream = d.stream; H2B_CK(e1); H2B_CK(e2); H2B_CK(e3); }
    const uint32_t tb = (uint32_t)ctx->tune.msm_sort_tb;
    if (tiled) {
        // two-level binning, shared-memory atomics only (msm_bin.cuh): digits + coarse totals, scan of the totals, partition into
        // coarse bins, one block per coarse bin for offsets / ends / placement / population histogram
        static const MsmBinKernelTable& BK = msm_bin_kernel_table();
        if (!d.bin_attr) {          // (this device is current: every caller of the sort stage has set it)
            if (BK.prepare() != 0) { cudaGetLastError(); d.launch_stream = d.stream; return ctx->fail(H2B_E_CUDA, "msm: cannot reserve the shared memory of the binning kernel"); }
            d.bin_attr = true;
        }
        const uint32_t lb = msm_bin_low_bits(nkeys), ncoarse = (nkeys + (1u << lb) - 1u) >> lb;
        uint32_t* cursor = S.coarse.as<uint32_t>();                       // [0, ncoarse]: totals -> exclusive scan -> run cursors
        uint32_t* start = cursor + MSM_BIN_MAX_COARSE + 4;                // [0, ncoarse]: the scan, kept; start[ncoarse] = total
        uint32_t* digits = entries;                                        // the digit words are dead once partitioned
        cudaError_t e4 = cudaMemsetAsync(cursor, 0, (size_t)(2 * MSM_BIN_MAX_COARSE + 8) * 4, st);
        if (e4 != cudaSuccess) { d.launch_stream = d.stream; H2B_CK(e4); }
        const uint32_t tiles = (n + MSM_BIN_TILE - 1) / MSM_BIN_TILE;
        if (n) H2B_LAUNCH(d, "msm_count", BK.digits, tiles, MSM_BIN_THREADS, 0, g, scalars, digits, cursor, lb, ncoarse);
        H2B_LAUNCH(d, "msm_scan", MK.scan_single_block, 1, SCAN_BLOCK, 0, cursor, ncoarse, start + ncoarse);
        cudaError_t e5 = cudaMemcpyAsync(start, cursor, (size_t)ncoarse * 4, cudaMemcpyDeviceToDevice, st);
        if (e5 != cudaSuccess) { d.launch_stream = d.stream; H2B_CK(e5); }
        const uint32_t ptile = msm_part_tile(g.windows);
        if (n) H2B_LAUNCH(d, "msm_scatter", BK.partition, (n + ptile - 1) / ptile, MSM_PART_THREADS, msm_part_smem(), g,
                          (const uint32_t*)digits, cursor, S.rec_key.as<unsigned short>(), S.rec_val.as<uint32_t>(), lb, ncoarse, ptile);
        const uint32_t slices = msm_bin_slices(lb, max_entries, ncoarse), cap = msm_bin_cap(lb);
        H2B_LAUNCH(d, "msm_scatter", BK.bin, ncoarse * slices, MSM_PART_THREADS, MSM_BIN_SMEM_BUDGET, (const uint32_t*)start,
                   (const unsigned short*)S.rec_key.as<unsigned short>(), (const uint32_t*)S.rec_val.as<uint32_t>(), nkeys, lb,
                   slices, cap, offsets, ends, entries, sizehist);
    } else {
        Fr* canon = ctx->tune.msm_canon_once ? S.canon.as<Fr>() : nullptr;
        uint32_t sgrid = (n + tb - 1) / tb;
        if (ctx->tune.msm_sort_blocks) sgrid = std::min<uint32_t>(sgrid, (uint32_t)(d.sms * ctx->tune.msm_sort_blocks));
        if (n) H2B_LAUNCH(d, "msm_count", MK.count, sgrid, tb, 0, g, scalars, counts, canon);
        H2B_LAUNCH(d, "msm_scan", MK.scan_block_sums, nblk, SCAN_BLOCK, 0, counts, nkeys, blocksums);
        H2B_LAUNCH(d, "msm_scan", MK.scan_single_block, 1, SCAN_BLOCK, 0, blocksums, nblk, (uint32_t*)nullptr);
        H2B_LAUNCH(d, "msm_scan", MK.scan_apply, nblk, SCAN_BLOCK, 0, counts, nkeys, blocksums, offsets, ends);
        if (n) H2B_LAUNCH(d, "msm_scatter", MK.scatter, sgrid, tb, 0, g, scalars, (const Fr*)canon, ends, entries);
        H2B_LAUNCH(d, "msm_order", MK.size_hist, (nkeys + tb - 1) / tb, tb, 0, offsets, ends, nkeys, sizehist);
    }
    H2B_LAUNCH(d, "msm_order", MK.scan_single_block, 1, SCAN_BLOCK, 0, sizehist, MSM_SIZE_BINS, (uint32_t*)nullptr);
    H2B_LAUNCH(d, "msm_order", MK.size_scatter, (nkeys + tb - 1) / tb, tb, 0, offsets, ends, nkeys, sizehist, order);
    H2B_LAUNCH(d, "msm_order", MK.long_list, (nkeys + tb - 1) / tb, tb, 0, offsets, ends, nkeys, pl.long_threshold,
               S.long_info.as<MsmLongInfo>(), S.long_tasks.as<MsmLongTask>(), counters);
    d.launch_stream = d.stream;
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}

// Accumulation stage of one chunk, enqueued on the COMPUTE stream: every bucket of sort set `S` is summed into the shared
// bucket array (add_to: on top of what earlier chunks left there).
int msm_stage_accumulate(h2b_ctx* ctx, Dev& d, MsmSortSet& S, const MsmChunkPlan& pl, const Affine* bases, bool add_to,
                         uint32_t set_index = 0, uint32_t n_sets = 1) {
    const uint32_t nkeys = pl.sp.nkeys;
    H2B_CK(d.m_buckets.reserve((size_t)nkeys * n_sets * sizeof(Xyzz)));
    H2B_CK(d.m_long_partials.reserve((size_t)pl.max_tasks * sizeof(Xyzz)));
    const uint32_t* offsets = S.offsets.as<uint32_t>();
    const uint32_t* ends = S.ends.as<uint32_t>();
    const uint32_t* order = S.order.as<uint32_t>();
    const uint32_t* entries = S.entries.as<uint32_t>();
    const uint32_t* counters = S.counters.as<uint32_t>();
    Xyzz* buckets = d.m_buckets.as<Xyzz>() + (size_t)set_index * nkeys;
    if (msm_ba_rounds(ctx, pl)) {
        // batched affine additions in front of the mixed additions (msm_ba.cuh): scratch rows [k][thread], thread = position in
        // the population order
        static const MsmBaKernelTable& BA = msm_ba_kernel_table();
        const uint32_t rounds = msm_ba_rounds(ctx, pl), stride = msm_ba_stride(nkeys);
        H2B_CK(d.m_ba.reserve(msm_ba_bytes(nkeys, rounds)));
        Fq* prefix = d.m_ba.as<Fq>();
        Fq* rx = rounds == 2 ? prefix + (size_t)MSM_BA_KCAP * stride : nullptr;
        Fq* ry = rounds == 2 ? rx + (size_t)MSM_BA_KCAP * stride : nullptr;
        H2B_LAUNCH(d, "msm_accumulate", BA.accumulate[rounds - 1][ctx->tune.msm_ba_regs], (nkeys + MSM_BA_THREADS - 1) / MSM_BA_THREADS,
                   MSM_BA_THREADS, 0, bases, entries, offsets, ends, order, nkeys, pl.long_threshold, buckets, add_to ? 1u : 0u,
                   prefix, rx, ry, stride);
    } else {
        auto kern = MK.accumulate[ctx->tune.msm_acc_variant];
        H2B_LAUNCH(d, "msm_accumulate", kern, (nkeys + 127) / 128, 128, 0, bases, entries, offsets, ends, order, nkeys,
                   pl.long_threshold, buckets, add_to ? 1u : 0u);
    }
    H2B_LAUNCH(d, "msm_accumulate_long", MK.accumulate_long, (unsigned)(d.sms * 8), MSM_LONG_THREADS, 0, bases, entries,
               offsets, ends, S.long_tasks.as<MsmLongTask>(), counters, d.m_long_partials.as<Xyzz>());
    H2B_LAUNCH(d, "msm_accumulate_long", MK.long_combine, (unsigned)d.sms, MSM_LONG_THREADS, 0,
               S.long_info.as<MsmLongInfo>(), counters, d.m_long_partials.as<Xyzz>(), buckets, add_to ? 1u : 0u);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}

// max_terms_total: upper bound on sets x terms the caller can take (the peer mailbox slot holds MSM_PEER_SLOT_POINTS); 0 = no bound.
// Beyond it the classic reduction (one term per bucket set) is used -- a pure function of the geometry, so every rank of a
// sharded MSM decides alike.
int msm_stage_reduce(h2b_ctx* ctx, Dev& d, uint32_t c, uint32_t pre_stride, MsmLaunchInfo* info, uint32_t batch_sets = 0,
                     bool copy_back = true, uint32_t max_terms_total = 0) {
    MsmShape sp = msm_shape(0, c, pre_stride, 0);
    const MsmGeom& g = sp.g;
    const uint32_t kw = batch_sets ? batch_sets : sp.kw;      // a batch of MSMs: one shared-window bucket set per column
    // reduce geometry: 128 threads x L buckets per block, at most MSM_REDUCE_MAX_SLICES blocks per bucket set
    uint32_t slices = 1, log_l = 0;
    if (g.nb >= (uint32_t)MSM_REDUCE_THREADS) {
        slices = g.nb / MSM_REDUCE_THREADS;
        while (slices > (uint32_t)MSM_REDUCE_MAX_SLICES) { slices >>= 1; log_l++; }
    }
    uint32_t lg_nb = 0;
    while ((1u << (lg_nb + 1)) <= g.nb) lg_nb++;
    Xyzz* buckets = d.m_buckets.as<Xyzz>();
    Xyzz* set_sums = nullptr;
    MsmGeom gr = g;
    gr.windows = kw;
    const bool lines_fit = !max_terms_total || (uint64_t)kw * (lg_nb + 1) <= max_terms_total;
    const bool lines_path = g.nb >= (1u << 14) && !ctx->tune.msm_reduce_classic && lines_fit;
    info->n_terms = 1;
    info->exps[0] = 0;
    if (lines_path) {
        // low-latency path: row / column segment sums, masked sums per index bit; the Horner over the bits is left to the host
        const uint32_t w = lg_nb / 2, h = lg_nb - w, H = 1u << h, W = 1u << w;
        const uint32_t seg = std::min<uint32_t>(W, (uint32_t)std::max(32, ctx->tune.msm_reduce_seg));   // buckets per line job
        const uint32_t rs = W / seg, cs = H / seg;
        const uint32_t nl = rs * H + cs * W, nt = h + w + 1;
        // (ONE reservation per call: growing a buffer frees the old one, and cudaFree waits for the whole device)
        H2B_CK(d.m_wpart.reserve(((size_t)kw * nl + (size_t)kw * nt) * sizeof(Xyzz)));
        Xyzz* wpart = d.m_wpart.as<Xyzz>();
        Xyzz* lines = wpart;
        Xyzz* terms = wpart + (size_t)kw * nl;
        set_sums = terms;
        H2B_LAUNCH(d, "msm_reduce", MK.reduce_lines, dim3((nl + 3) / 4, kw, 1), MSM_REDUCE_THREADS, 0, (const Xyzz*)buckets, g.nb, w, rs, cs, lines);
        H2B_LAUNCH(d, "msm_reduce", MK.reduce_masked, dim3(nt, kw, 1), MSM_MASKED_THREADS, 0, (const Xyzz*)lines, h, w, rs, cs, terms);
        info->n_terms = nt;
        for (uint32_t b = 0; b < h; b++) info->exps[b] = b + w;          // S^D_b carries 2^b * W
        for (uint32_t b = 0; b < w; b++) info->exps[h + b] = b;          // S^C_b carries 2^b
        info->exps[h + w] = 0;                                           // V
    } else if (g.nb >= (uint32_t)MSM_REDUCE_THREADS) {
        H2B_CK(d.m_wpart.reserve(((size_t)kw * slices * 2 + kw) * sizeof(Xyzz)));
        Xyzz* wpart = d.m_wpart.as<Xyzz>();
        Xyzz* block_p = wpart;
        Xyzz* block_r = wpart + (size_t)kw * slices;
        set_sums = wpart + (size_t)kw * slices * 2;
        dim3 grid(slices, kw, 1);
        H2B_LAUNCH(d, "msm_reduce", MK.reduce, grid, MSM_REDUCE_THREADS, 0, gr, buckets, slices, log_l, block_p, block_r);
        H2B_LAUNCH(d, "msm_reduce", MK.reduce_final, kw, MSM_REDUCE_MAX_SLICES, 0, slices, log_l + 7u, block_p, block_r, set_sums);
    } else {
        H2B_CK(d.m_wpart.reserve((size_t)kw * sizeof(Xyzz)));
        set_sums = d.m_wpart.as<Xyzz>();
        H2B_LAUNCH(d, "msm_reduce", MK.reduce_small, (kw + 31) / 32, 32, 0, gr, buckets, set_sums);
    }
    H2B_CK(cudaGetLastError());
    const size_t out_bytes = (size_t)kw * info->n_terms * sizeof(Xyzz);
    H2B_CK(d.reserve_pinned(std::max<size_t>(out_bytes, 4096)));
    if (copy_back) H2B_CK(d.ring.d2h(d.pinned, set_sums, out_bytes, d.stream));
    info->c = c; info->windows = kw; info->pre = pre_stride != 0; info->d_terms = set_sums;
    return H2B_OK;
}

// host: finish each bucket set (Horner over its terms) and add it to its window (after the stream has been synchronised)
void msm_fold_terms(const Dev& d, const MsmLaunchInfo& info, Xyzz* windows /* accumulated into */) {
    const Xyzz* part = reinterpret_cast<const Xyzz*>(d.pinned);
    // (with precomputed multiples there is one bucket set whose sum already carries every window's weight: it joins window 0)
    for (uint32_t j = 0; j < info.windows; j++)
        windows[j] = xyzz_add(windows[j], msm_horner_terms(part + (size_t)j * info.n_terms, info.exps, info.n_terms));
}

void begin_call(Dev& d) { cudaEventRecord(d.ev_first, d.stream); }
int end_call(h2b_ctx* ctx, Dev& d) {
    H2B_CK(cudaEventRecord(d.ev_last, d.stream));
    H2B_CK(cudaStreamSynchronize(d.stream));
    float ms = 0.f;
    H2B_CK(cudaEventElapsedTime(&ms, d.ev_first, d.ev_last));
    ctx->last_ms = ms;
    return H2B_OK;
}

// MSM over all devices of the context.  scalars: host pointer, or device-0 pointer when on_device.
int msm_peer_epilogue(h2b_ctx* ctx, Dev& d, const MsmLaunchInfo& info);

int msm_run(h2b_ctx* ctx, const Bases& B, const void* scalars, bool on_device, size_t n, size_t n_total,
            Xyzz* windows_out, uint32_t* c_out, uint32_t* windows_count, bool peer_exchange = false) {
    if (n > B.n) return ctx->fail(H2B_E_INVALID, "msm: more scalars than registered bases");
    uint32_t c = msm_choose_c(ctx, n_total);
    uint32_t nw = msm_windows(c);
    if (nw > H2B_MSM_MAX_WINDOWS) return ctx->fail(H2B_E_INVALID, "msm: window size too small");
    for (uint32_t j = 0; j < nw; j++) windows_out[j] = Xyzz::identity();
    *c_out = c; *windows_count = nw;
    if (n == 0) { ctx->last_ms = 0; return H2B_OK; }
    if (n >= ((size_t)1 << 31)) return ctx->fail(H2B_E_INVALID, "msm: n too large");
    std::vector<MsmLaunchInfo> infos(B.shard.size());
    std::vector<int> used;
    double max_ms = 0;
    // One host thread per device when several shards take part: the staged copies of pageable scalars block their caller, and
    // the devices must not queue up behind one another.
    auto run_shard = [&](size_t di) -> int {
        const BasesShard& sh = B.shard[di];
        size_t cnt = std::min(sh.count, n - sh.start);
        Dev& d = ctx->devs[di];
        H2B_CK(cudaSetDevice(d.id));
        begin_call(d);
        const Affine* dev_bases = reinterpret_cast<const Affine*>(sh.pre ? sh.pre : sh.d);
        const uint32_t use_c = sh.pre ? sh.pre_c : c;
        const uint32_t stride = sh.pre ? (uint32_t)sh.count : 0u;
        int rc = H2B_OK;
        // The scalars are processed in chunks: chunk k+1 crosses PCIe on the copy stream (host scalars) and is binned on the
        // sort stream while chunk k is accumulated on the compute stream; all chunks add into ONE bucket array.
        const bool staged = !(on_device && di == 0);
        size_t chunk = cnt;
        {   // at most 8 chunks of >= 2^20 points (device scalars) or >= 2^18 (host scalars, where the chunks also hide the
            // PCIe copy); measured on B200: profiles/README.md
            size_t k = std::min<size_t>(8, cnt >> (staged && !on_device ? 18 : 20));
            if (k > 1) chunk = (cnt + k - 1) / k;
        }
        // (host scalars only: there the first chunk also waits for its PCIe copy; with device-resident scalars the shorter
        // chunks cost the accumulation more than the earlier start saves -- profiles/README.md)
        bool ramp = chunk < cnt && staged && !on_device && !ctx->tune.msm_no_ramp;
        if (ctx->tune.msm_chunks) { chunk = (cnt + ctx->tune.msm_chunks - 1) / ctx->tune.msm_chunks; ramp = false; }
        // Chunk boundaries.  Nothing hides the binning of the FIRST chunk, so the pipeline starts with a short chunk and doubles
        // it until the steady chunk size is reached: each chunk's accumulation still covers the binning of the next one.
        std::vector<size_t> bounds{0};
        if (ramp)
            for (size_t sz = std::max<size_t>(chunk / 8, (size_t)1 << 17); sz < chunk && bounds.back() + sz + chunk <= cnt; sz *= 2)
                bounds.push_back(bounds.back() + sz);
        while (bounds.back() < cnt) bounds.push_back(std::min(cnt, bounds.back() + chunk));
        if (bounds.size() > 2 && cnt - bounds[bounds.size() - 2] < chunk / 4) bounds.erase(bounds.end() - 2);   // no runt at the end
        size_t nchunks = bounds.size() - 1;
        const Fr* dev_scalars = reinterpret_cast<const Fr*>(scalars) + sh.start;
        const uint8_t* host_src = reinterpret_cast<const uint8_t*>(scalars) + sh.start * 32;
        if (staged) {
            H2B_CK(d.m_scalars.reserve(cnt * 32));
            while (d.chunk_events.size() < nchunks) { cudaEvent_t e; H2B_CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); d.chunk_events.push_back(e); }
            H2B_CK(cudaStreamWaitEvent(d.copy_stream, d.ev_first, 0));      // previous call's use of m_scalars is over
            dev_scalars = d.m_scalars.as<Fr>();
        }
        H2B_CK(cudaStreamWaitEvent(d.sort_stream, d.ev_first, 0));
        {   // every buffer of the pipeline sized for the largest chunk up front: growing one later frees it, which waits for the
            // whole device in the middle of the sort / accumulate overlap
            size_t longest = 0;
            for (size_t k = 0; k < nchunks; k++) longest = std::max(longest, bounds[k + 1] - bounds[k]);
            MsmChunkPlan big;
            rc = msm_chunk_plan(ctx, (uint32_t)longest, 0u, use_c, stride, &big);
            if (rc) return rc;
            for (int q = 0; q < (nchunks > 1 ? 2 : 1); q++) { rc = msm_sort_reserve(ctx, d.ms[q], big); if (rc) return rc; }
            H2B_CK(d.m_buckets.reserve((size_t)big.sp.nkeys * sizeof(Xyzz)));
            H2B_CK(d.m_long_partials.reserve((size_t)big.max_tasks * sizeof(Xyzz)));
            if (msm_ba_rounds(ctx, big)) H2B_CK(d.m_ba.reserve(msm_ba_bytes(big.sp.nkeys, msm_ba_rounds(ctx, big))));
        }
        for (size_t k = 0; k < nchunks && rc == H2B_OK; k++) {
            size_t lo = bounds[k], len = bounds[k + 1] - lo;
            MsmSortSet& S = d.ms[k & 1];
            MsmChunkPlan pl;
            rc = msm_chunk_plan(ctx, (uint32_t)len, stride ? (uint32_t)lo : 0u, use_c, stride, &pl);
            if (rc) break;
            if (staged) {
                // chunk k crosses PCIe on the copy stream (pageable scalars: through the pinned ring, hostcopy.hpp) while the
                // kernels of chunk k-1, enqueued a moment ago, run
                if (on_device) H2B_CK(cudaMemcpyAsync(d.m_scalars.as<uint8_t>() + lo * 32, host_src + lo * 32, len * 32, cudaMemcpyDeviceToDevice, d.copy_stream));
                else H2B_CK(d.ring.h2d(d.m_scalars.as<uint8_t>() + lo * 32, host_src + lo * 32, len * 32, d.copy_stream));
                H2B_CK(cudaEventRecord(d.chunk_events[k], d.copy_stream));
            }
            if (S.consumed_pending) { H2B_CK(cudaStreamWaitEvent(d.sort_stream, S.consumed, 0)); S.consumed_pending = false; }
            if (staged) H2B_CK(cudaStreamWaitEvent(d.sort_stream, d.chunk_events[k], 0));
            rc = msm_stage_sort(ctx, d, S, pl, dev_scalars + lo);
            if (rc) break;
            H2B_CK(cudaEventRecord(S.sorted, d.sort_stream));
            H2B_CK(cudaStreamWaitEvent(d.stream, S.sorted, 0));
            const Affine* b = stride ? dev_bases : dev_bases + lo;           // classic mode indexes bases relative to the chunk
            rc = msm_stage_accumulate(ctx, d, S, pl, b, k > 0);
            if (rc) break;
            H2B_CK(cudaEventRecord(S.consumed, d.stream));
            S.consumed_pending = true;
        }
        if (rc) return rc;
        rc = msm_stage_reduce(ctx, d, use_c, stride, &infos[di], 0, !peer_exchange, peer_exchange ? (uint32_t)MSM_PEER_SLOT_POINTS : 0u);
        if (rc) return rc;
        if (peer_exchange) { rc = msm_peer_epilogue(ctx, d, infos[di]); if (rc) return rc; }
        H2B_CK(cudaEventRecord(d.ev_last, d.stream));
        return H2B_OK;
    };
    // a failed shard leaves copies / kernels in flight on three streams and `consumed` events pending: drain them so that the
    // next call on this context starts from a clean state
    auto drain = [&](int rc_in) -> int {
        for (size_t di = 0; di < B.shard.size() && di < ctx->devs.size(); di++) {
            Dev& d = ctx->devs[di];
            cudaSetDevice(d.id);
            cudaStreamSynchronize(d.copy_stream); cudaStreamSynchronize(d.sort_stream); cudaStreamSynchronize(d.stream);
            d.ms[0].consumed_pending = d.ms[1].consumed_pending = false;
            d.launch_stream = d.stream;
        }
        cudaGetLastError();
        cudaSetDevice(ctx->devs[0].id);
        return rc_in;
    };
    for (size_t di = 0; di < B.shard.size(); di++)
        if (B.shard[di].start < n && B.shard[di].count != 0) used.push_back((int)di);
    if (peer_exchange && used.size() != 1) return ctx->fail(H2B_E_INVALID, "msm_sharded: needs a single-device context with a non-empty range");
    if (used.size() == 1) {
        int rc = run_shard((size_t)used[0]);
        if (rc) return drain(rc);
    } else {
        std::vector<int> rcs(used.size(), H2B_OK);
        std::vector<std::thread> th;
        for (size_t u = 0; u < used.size(); u++) th.emplace_back([&, u] { rcs[u] = run_shard((size_t)used[u]); });
        for (auto& t : th) t.join();
        for (int rc : rcs) if (rc) return drain(rc);
    }
    for (int di : used) {
        Dev& d = ctx->devs[di];
        H2B_CK(cudaSetDevice(d.id));
        H2B_CK(cudaStreamSynchronize(d.stream));
        float ms = 0.f;
        H2B_CK(cudaEventElapsedTime(&ms, d.ev_first, d.ev_last));
        max_ms = std::max(max_ms, (double)ms);
        if (peer_exchange && *ctx->peer.status != 0) {
            const uint32_t st = *ctx->peer.status;
            return ctx->fail(H2B_E_CUDA, st >= 0x300 ? "msm_sharded: rank " + std::to_string(st - 0x300) + " gave up on this call (it timed out or failed in its exchange)"
                                         : st >= 0x200 ? "msm_sharded: rank " + std::to_string(st - 0x200) + " is ahead of this rank (the ranks have not made the same number of calls): every rank must h2b_peer_attach again"
                                         : st >= 0x100 ? "msm_sharded: rank " + std::to_string(st - 0x100) + " reduced with a different window geometry (register equal shard sizes)"
                                                     : "msm_sharded: timed out waiting for the terms of rank " + std::to_string(st - 1));
        }
        msm_fold_terms(d, infos[di], windows_out);
    }
    H2B_CK(cudaSetDevice(ctx->devs[0].id));      // the caller's (torch's) current device is device 0 of the context
    ctx->last_ms = max_ms;
    return H2B_OK;
}

// Epilogue of the reduction in a sharded MSM across processes: ONE kernel publishes this rank's terms to every peer mailbox,
// waits for theirs and folds them per term into the page-locked result buffer (msm.cuh 7).
int msm_peer_epilogue(h2b_ctx* ctx, Dev& d, const MsmLaunchInfo& info) {
    PeerState& P = ctx->peer;
    if (P.world <= 0 || !P.mailbox) return ctx->fail(H2B_E_INVALID, "msm_sharded: h2b_peer_attach has not been called");
    const uint32_t nt = info.windows * info.n_terms;
    if (nt > MSM_PEER_SLOT_POINTS) return ctx->fail(H2B_E_INVALID, "msm_sharded: too many terms for one mailbox slot");
    const size_t smem = (size_t)nt * (size_t)P.world * sizeof(Xyzz);
    if (smem > ((size_t)200 << 10)) return ctx->fail(H2B_E_INVALID, "msm_sharded: terms x ranks exceed the shared memory of the fold");
    if (!d.peer_attr) { H2B_CK(cudaFuncSetAttribute((const void*)MK.peer_exchange, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10)); d.peer_attr = true; }
    const uint64_t epoch = P.epoch;          // bumped by h2b_msm_sharded on entry, so that a call failing before this point still counts
    // geometry tag: every rank must have reduced to the same terms with the same weights
    uint64_t tag = ((uint64_t)info.c << 48) | ((uint64_t)info.windows << 32) | ((uint64_t)info.n_terms << 16) | (info.pre ? 1u : 0u);
    *P.status = 0xffffffffu;
    H2B_LAUNCH(d, "msm_peer_exchange", MK.peer_exchange, 1, 256, smem, (const Xyzz*)info.d_terms, nt, P.table, (uint32_t)P.rank,
               (uint32_t)P.world, (uint32_t)(epoch & 1), (unsigned long long)epoch, (unsigned long long)tag,
               (unsigned long long)ctx->tune.peer_timeout_ms * 1000000ull, reinterpret_cast<Xyzz*>(d.pinned), P.status);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}

// A batch of MSMs against the SAME registered bases (the per-column commits of one prover phase, plonk/prover.rs:436-440):
// column k is "chunk" k of the two-stream pipeline with its own bucket set, and ONE reduction launch handles every set, so the
// binning of column k+1 hides behind the accumulation of column k and the reduction latency is paid once.
// Needs the precomputed-window layout (one bucket set per MSM) on a single device; callers fall back to a loop otherwise.
int msm_run_batch_enqueue(h2b_ctx* ctx, const Bases& B, const void* const* cols, bool on_device, size_t m, size_t n, Xyzz* sums_out,
                          uint32_t* c_out);
int msm_run_batch(h2b_ctx* ctx, const Bases& B, const void* const* cols, bool on_device, size_t m, size_t n, Xyzz* sums_out,
                  uint32_t* c_out) {
    const BasesShard& sh = B.shard[0];
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    int rc = msm_run_batch_enqueue(ctx, B, cols, on_device, m, n, sums_out, c_out);
    if (rc) {          // copies / kernels may be in flight on three streams and `consumed` events pending: start the next call clean
        cudaStreamSynchronize(d.copy_stream); cudaStreamSynchronize(d.sort_stream); cudaStreamSynchronize(d.stream);
        d.ms[0].consumed_pending = d.ms[1].consumed_pending = false;
        d.launch_stream = d.stream;
        cudaGetLastError();
    }
    return rc;
}
int msm_run_batch_enqueue(h2b_ctx* ctx, const Bases& B, const void* const* cols, bool on_device, size_t m, size_t n, Xyzz* sums_out,
                          uint32_t* c_out) {
    const BasesShard& sh = B.shard[0];
    Dev& d = ctx->devs[0];
    begin_call(d);
    const Affine* dev_bases = reinterpret_cast<const Affine*>(sh.pre);
    const uint32_t use_c = sh.pre_c, stride = (uint32_t)sh.count;
    if (!on_device) {
        H2B_CK(d.m_scalars.reserve(m * n * 32));
        while (d.chunk_events.size() < m) { cudaEvent_t e; H2B_CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); d.chunk_events.push_back(e); }
        H2B_CK(cudaStreamWaitEvent(d.copy_stream, d.ev_first, 0));
    }
    H2B_CK(cudaStreamWaitEvent(d.sort_stream, d.ev_first, 0));
    for (size_t k = 0; k < m; k++) {
        MsmSortSet& S = d.ms[k & 1];
        MsmChunkPlan pl;
        int rc = msm_chunk_plan(ctx, (uint32_t)n, 0u, use_c, stride, &pl);
        if (rc) return rc;
        if (S.consumed_pending) { H2B_CK(cudaStreamWaitEvent(d.sort_stream, S.consumed, 0)); S.consumed_pending = false; }
        if (!on_device) {
            H2B_CK(d.ring.h2d(d.m_scalars.as<uint8_t>() + k * n * 32, cols[k], n * 32, d.copy_stream));
            H2B_CK(cudaEventRecord(d.chunk_events[k], d.copy_stream));
            H2B_CK(cudaStreamWaitEvent(d.sort_stream, d.chunk_events[k], 0));
        }
        const Fr* sc = on_device ? reinterpret_cast<const Fr*>(cols[k]) : d.m_scalars.as<Fr>() + k * n;
        rc = msm_stage_sort(ctx, d, S, pl, sc);
        if (rc) return rc;
        H2B_CK(cudaEventRecord(S.sorted, d.sort_stream));
        H2B_CK(cudaStreamWaitEvent(d.stream, S.sorted, 0));
        rc = msm_stage_accumulate(ctx, d, S, pl, dev_bases, false, (uint32_t)k, (uint32_t)m);
        if (rc) return rc;
        H2B_CK(cudaEventRecord(S.consumed, d.stream));
        S.consumed_pending = true;
    }
    MsmLaunchInfo info;
    int rc = msm_stage_reduce(ctx, d, use_c, stride, &info, (uint32_t)m);
    if (rc) return rc;
    rc = end_call(ctx, d);
    if (rc) return rc;
    const Xyzz* part = reinterpret_cast<const Xyzz*>(d.pinned);
    for (size_t k = 0; k < m; k++) sums_out[k] = msm_horner_terms(part + k * info.n_terms, info.exps, info.n_terms);
    *c_out = use_c;
    return H2B_OK;
}

// Window size used with precomputed multiples (one shared bucket set, so larger windows stay cheap to reduce).
uint32_t msm_choose_pre_c(const h2b_ctx* ctx, size_t n) {
    if (ctx->tune.msm_pre_c) return (uint32_t)ctx->tune.msm_pre_c;
    // Measured on B200 (profiles/README.md): c = 20 (13 windows, 2^19 buckets) wins from 2^19 points up, c = 17 from 2^16;
    // below that one bucket set has too few buckets to fill the GPU and the classic per-window layout is used (0).
    // Window sizes whose top window would hold only 1-2 scalar bits (18, 19, 21, 23) are avoided: every point would
    // land in a handful of buckets of that window.
    if (n >= ((size_t)1 << 19)) return 20;
    if (n >= ((size_t)1 << 16)) return 17;
    return 0;
}

// Registration = MsmAccel::get_base_descriptor's "expensive preprocessing" (zal.rs:71-77): besides making the bases
// resident, store 2^(c*j) * P_i for every window j so that an MSM needs ONE bucket set instead of one per window.
// Skipped (classic per-window buckets) for one-shot bases, tiny inputs, or when the multiples would not fit.
int precompute_shard(h2b_ctx* ctx, Dev& d, BasesShard& sh) {
    if (!ctx->tune.msm_precompute) return H2B_OK;
    if (sh.count < 256) return H2B_OK;
    uint32_t c = msm_choose_pre_c(ctx, sh.count);
    if (c == 0) return H2B_OK;
    uint32_t w = msm_windows(c);
    if ((uint64_t)sh.count * w >= 0x7fffffffull) return H2B_OK;
    size_t bytes = (size_t)sh.count * w * 64, free_b = 0, total_b = 0;
    H2B_CK(cudaMemGetInfo(&free_b, &total_b));
    if (bytes > free_b / 2) {
        // not an error, but a 2-3x slower MSM: say so (h2b_last_notice, h2b_bases_info) instead of dropping to the
        // per-window layout silently
        std::lock_guard<std::mutex> lk(ctx->err_mu);
        ctx->status = "bases_register: window multiples (" + std::to_string(bytes >> 20) + " MiB) exceed half of the free device memory (" +
                      std::to_string(free_b >> 20) + " MiB): registered WITHOUT precomputation, MSMs on this handle use one bucket set per window";
        sh.pre_skipped = true;
        return H2B_OK;
    }
    H2B_CK(cudaMalloc(&sh.pre, bytes));
    H2B_LAUNCH(d, "msm_precompute", MK.precompute, (unsigned)((sh.count + 127) / 128), 128, 0, (const Affine*)sh.d,
               (Affine*)sh.pre, (uint32_t)sh.count, (uint32_t)sh.count, c, w);
    H2B_CK(cudaGetLastError());
    H2B_CK(cudaStreamSynchronize(d.stream));
    sh.pre_c = c; sh.pre_w = w;
    return H2B_OK;
}

void release_bases(h2b_ctx* ctx, Bases& B) {
    for (size_t di = 0; di < B.shard.size(); di++)
        {
            cudaSetDevice(ctx->devs[di].id);
            if (B.shard[di].d && B.shard[di].owned) cudaFree(B.shard[di].d);
            if (B.shard[di].pre) cudaFree(B.shard[di].pre);
        }
    if (!ctx->devs.empty()) cudaSetDevice(ctx->devs[0].id);
}

int register_bases_impl(h2b_ctx* ctx, Bases& B, const void* bases, bool on_device, size_t n, bool allow_pre) {
    B.n = n;
    // Host bases are cut into one contiguous point range per device.  Device-resident bases (on device 0) are adopted in place
    // by a single-device context; a multi-device context keeps range 0 where it is and hands the other ranges to their devices
    // over NVLink (peer copy), so that a device-resident SRS (ParamsKZG::setup) shards exactly like a host one.
    size_t nd = ctx->devs.size();
    B.shard.resize(ctx->devs.size());
    size_t per = (n + nd - 1) / nd;
    for (size_t di = 0; di < nd; di++) {
        BasesShard& sh = B.shard[di];
        sh.start = std::min(n, di * per);
        sh.count = std::min(per, n - sh.start);
        if (sh.count == 0) continue;
        Dev& d = ctx->devs[di];
        H2B_CK(cudaSetDevice(d.id));
        const uint8_t* src = reinterpret_cast<const uint8_t*>(bases) + sh.start * 64;
        if (on_device && di == 0) { sh.d = const_cast<uint8_t*>(src); sh.owned = false; }
        else if (on_device) {
            H2B_CK(cudaMalloc(&sh.d, sh.count * 64));
            H2B_CK(cudaMemcpyPeerAsync(sh.d, d.id, src, ctx->devs[0].id, sh.count * 64, d.stream));
            H2B_CK(cudaStreamSynchronize(d.stream));
        } else {
            H2B_CK(cudaMalloc(&sh.d, sh.count * 64));
            H2B_CK(d.ring.h2d(sh.d, src, sh.count * 64, d.stream));
            H2B_CK(cudaStreamSynchronize(d.stream));
        }
        if (allow_pre) { int rc = precompute_shard(ctx, d, sh); if (rc) return rc; }
        if (sh.pre_skipped) B.pre_skipped = true;
    }
    return H2B_OK;
}

int register_bases(h2b_ctx* ctx, const void* bases, bool on_device, size_t n, bool allow_pre, uint64_t* handle_out) {
    Bases B;
    int rc = register_bases_impl(ctx, B, bases, on_device, n, allow_pre);
    if (rc) {                       // a failed copy / allocation / precomputation must not leak the shards already made
        for (Dev& d : ctx->devs) { cudaSetDevice(d.id); cudaStreamSynchronize(d.stream); }
        cudaGetLastError();
        release_bases(ctx, B);
        return rc;
    }
    H2B_CK(cudaSetDevice(ctx->devs[0].id));
    uint64_t h = ctx->next_handle++;
    ctx->bases.emplace(h, std::move(B));
    *handle_out = h;
    return H2B_OK;
}



// ------------------------------------------------------------------------------------------------
// polynomial helpers (poly.cuh): three-phase tiled scans
// ------------------------------------------------------------------------------------------------
const PolyKernelTable& PK = poly_kernel_table();

inline uint32_t poly_tiles(uint64_t n) { return (uint32_t)((n + POLY_TILE - 1) / POLY_TILE); }

// z[i] = init * prod_{j < i} v[j], i < out_len (out_len <= n + 1); z may alias v
int poly_grand_product_enqueue(h2b_ctx* ctx, Dev& d, const Fr* v, uint64_t n, const Fr& init, Fr* z, uint64_t out_len) {
    if (out_len == 0) return H2B_OK;
    const uint32_t tiles = poly_tiles(out_len);
    H2B_CK(d.p_tiles.reserve((size_t)tiles * 32 + 64));
    Fr* tt = d.p_tiles.as<Fr>();
    H2B_LAUNCH(d, "poly_grand_product", PK.prod_tile_reduce, tiles, POLY_T, 0, v, n, tt, 0u);
    H2B_LAUNCH(d, "poly_grand_product", PK.prod_scan_totals, 1, POLY_T, 0, tt, tiles, init);
    H2B_LAUNCH(d, "poly_grand_product", PK.prod_tile_scan, tiles, POLY_T, 0, v, n, (const Fr*)tt, z, out_len);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}
int poly_batch_invert_enqueue(h2b_ctx* ctx, Dev& d, Fr* a, uint64_t n) {
    if (n == 0) return H2B_OK;
    const uint32_t tiles = poly_tiles(n);
    H2B_CK(d.p_tiles.reserve((size_t)tiles * 64 + 64));
    Fr* tt = d.p_tiles.as<Fr>();
    H2B_LAUNCH(d, "poly_batch_invert", PK.prod_tile_reduce, tiles, POLY_T, 0, (const Fr*)a, n, tt, 1u);
    H2B_LAUNCH(d, "poly_batch_invert", PK.invert_totals, 1, POLY_T, 0, tt, tt + tiles, tiles);
    H2B_LAUNCH(d, "poly_batch_invert", PK.batch_invert, tiles, POLY_T, 0, a, n, (const Fr*)tt);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}
// Horner machinery: q (n - 1 entries) = a / (X - b) when q != nullptr; *d_final = sum a_i b^i when d_final != nullptr
int poly_horner_enqueue(h2b_ctx* ctx, Dev& d, const Fr* a, uint64_t n, const Fr& b, Fr* q, Fr* d_final) {
    const uint32_t tiles = poly_tiles(n);
    const uint64_t pad = (uint64_t)tiles * POLY_TILE - n;
    H2B_CK(d.p_tiles.reserve((size_t)tiles * 32 + 64));
    Fr* tv = d.p_tiles.as<Fr>();
    Fr bk = fe_pow_u64(b, (uint64_t)POLY_K), bt = fe_pow_u64(b, (uint64_t)POLY_TILE);
    H2B_LAUNCH(d, "poly_horner", PK.horner_tile_reduce, tiles, POLY_T, 0, a, n, pad, b, bk, tv);
    H2B_LAUNCH(d, "poly_horner", PK.horner_scan_totals, 1, POLY_T, 0, tv, tiles, bt, d_final);
    if (q) H2B_LAUNCH(d, "poly_horner", PK.horner_tile_scan, tiles, POLY_T, 0, a, n, pad, b, bk, (const Fr*)tv, q);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}


// ------------------------------------------------------------------------------------------------
// quotient evaluation (quotient.cuh)
// ------------------------------------------------------------------------------------------------
const QuotientKernelTable& QK = quotient_kernel_table();

// Bump allocator over one staging buffer: everything a call needs on the device (program words, constants, pointer tables)
// is packed on the host and uploaded with ONE copy.
struct Stage {
    std::vector<uint8_t> host;
    size_t put(const void* src, size_t bytes) {
        size_t off = (host.size() + 31) & ~(size_t)31;
        host.resize(off + bytes);
        if (bytes) std::memcpy(host.data() + off, src, bytes);
        return off;
    }
};

// Walk the program once on the host: every index must be in range (the kernel trusts it).
int quotient_validate(h2b_ctx* ctx, const h2b_graph* g, size_t n_fixed, size_t n_advice, size_t n_instance, size_t n_challenges,
                      uint32_t* last_target, bool* has_calcs) {
    auto bad = [&](const char* what) { return ctx->fail(H2B_E_INVALID, std::string("quotient_graph: ") + what); };
    if (g->n_rotations > Q_MAX_ROTATIONS) return bad("too many distinct rotations");
    auto src_ok = [&](uint32_t w0, uint32_t w1) {
        uint32_t kind = w0 & 0xffu, rot = w0 >> 8;
        switch (kind) {
            case QS_CONSTANT: return w1 < g->n_constants;
            case QS_INTERMEDIATE: return w1 < g->n_intermediates;
            case QS_FIXED: return w1 < n_fixed && rot < g->n_rotations;
            case QS_ADVICE: return w1 < n_advice && rot < g->n_rotations;
            case QS_INSTANCE: return w1 < n_instance && rot < g->n_rotations;
            case QS_CHALLENGE: return w1 < n_challenges;
            case QS_BETA: case QS_GAMMA: case QS_THETA: case QS_Y: case QS_PREVIOUS: return true;
            default: return false;
        }
    };
    size_t pc = 0;
    *has_calcs = false;
    while (pc < g->n_words) {
        if (pc + 6 > g->n_words) return bad("truncated calculation");
        const uint32_t* w = g->program + pc;
        if (w[0] > QC_STORE) return bad("unknown calculation");
        if (w[1] >= g->n_intermediates) return bad("target out of range");
        if (!src_ok(w[2], w[3])) return bad("value source out of range");
        if (w[0] == QC_HORNER) {
            if (pc + 7 > g->n_words) return bad("truncated Horner");
            size_t np = w[6];
            if (!src_ok(w[4], w[5]) || pc + 7 + 2 * np > g->n_words) return bad("truncated Horner");
            for (size_t i = 0; i < np; i++) if (!src_ok(w[7 + 2 * i], w[8 + 2 * i])) return bad("value source out of range");
            pc += 7 + 2 * np;
        } else {
            if (w[0] <= QC_MUL && !src_ok(w[4], w[5])) return bad("value source out of range");
            pc += 6;
        }
        *last_target = w[1];
        *has_calcs = true;
    }
    return H2B_OK;
}
}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

const char* h2b_version(void) { return "h2b200 0.1 (sm_100a)"; }

int h2b_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int h2b_init(h2b_ctx** out, const int* device_ids, int n_devices) {
    if (!out) return H2B_E_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_init_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
        return H2B_E_NODEVICE;
    }
    std::vector<int> ids;
    if (!device_ids || n_devices <= 0) ids.push_back(0);
    else ids.assign(device_ids, device_ids + n_devices);
    h2b_ctx* ctx = new h2b_ctx();
    ctx->tune.from_environment();          // the only place the H2B_* environment is consulted
    ctx->devs.resize(ids.size());
    for (size_t i = 0; i < ids.size(); i++) {
        Dev& d = ctx->devs[i];
        d.id = ids[i];
        if (d.id < 0 || d.id >= count) { g_init_error = "bad device id"; delete ctx; return H2B_E_INVALID; }
        cudaDeviceProp prop;
        if (cudaSetDevice(d.id) != cudaSuccess || cudaGetDeviceProperties(&prop, d.id) != cudaSuccess) {
            g_init_error = "cudaSetDevice failed"; delete ctx; return H2B_E_CUDA;
        }
        if (prop.major != 10) {
            g_init_error = "device is not sm_100 (this library is built for B200 only)"; delete ctx; return H2B_E_NODEVICE;
        }
        d.sms = prop.multiProcessorCount;
        if (ctx->tune.l2_fetch) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)ctx->tune.l2_fetch);   // experiment knob
        int prio_lo = 0, prio_hi = 0;
        cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        // numerically lower = higher priority; the binning stream runs above the compute stream by default (sort_prio)
        const int prio_compute = ctx->tune.sort_prio < 0 ? prio_hi : prio_lo;
        const int prio_sort = ctx->tune.sort_prio > 0 ? prio_hi : prio_lo;
        if (cudaStreamCreateWithPriority(&d.stream, cudaStreamNonBlocking, prio_compute) != cudaSuccess ||
            cudaStreamCreateWithFlags(&d.copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithPriority(&d.sort_stream, cudaStreamNonBlocking, prio_sort) != cudaSuccess ||
            cudaEventCreate(&d.ev_first) != cudaSuccess || cudaEventCreate(&d.ev_last) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.ms[0].sorted, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.ms[0].consumed, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.ms[1].sorted, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&d.ms[1].consumed, cudaEventDisableTiming) != cudaSuccess) {
            g_init_error = "stream/event creation failed"; delete ctx; return H2B_E_CUDA;
        }
        d.launch_stream = d.stream;
        if (d.reserve_pinned(64 << 10) != cudaSuccess) { g_init_error = "page-locked result buffer allocation failed"; delete ctx; return H2B_E_NOMEM; }
        int rc = ntt_prepare_device(ctx);
        if (rc) { g_init_error = ctx->err; delete ctx; return rc; }
    }
    // Device-resident scalars / columns live on device 0 and are handed to the other devices by cudaMemcpyAsync device-to-device:
    // with peer access these copies ride NVLink / NVSwitch; without it the driver stages them through host memory.
    for (size_t i = 0; i < ids.size(); i++)
        for (size_t j = 0; j < ids.size(); j++) {
            if (i == j || ids[i] == ids[j]) continue;      // (a device may be listed twice: two slots of streams / buffers on one GPU)
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, ids[i], ids[j]) == cudaSuccess && can) {
                cudaSetDevice(ids[i]);
                cudaError_t pe = cudaDeviceEnablePeerAccess(ids[j], 0);
                if (pe != cudaSuccess) cudaGetLastError();      // already enabled by the caller (torch, another context): fine
            }
        }
    cudaGetLastError();                                    // a refused peer query must not surface in a later launch check
    cudaSetDevice(ctx->devs[0].id);
    { std::lock_guard<std::mutex> lk(live_mu()); live_set().insert(ctx); }
    *out = ctx;
    return H2B_OK;
}

void h2b_destroy(h2b_ctx* ctx) {
    if (!ctx) return;
    { std::lock_guard<std::mutex> lk(live_mu()); if (!live_set().erase(ctx)) return; }      // unknown or already destroyed
    for (auto& kv : ctx->bases) release_bases(ctx, kv.second);
    if (ctx->peer.mailbox) {
        cudaSetDevice(ctx->devs[0].id);
        cudaStreamSynchronize(ctx->devs[0].stream);
        for (void* q : ctx->peer.opened) cudaIpcCloseMemHandle(q);
        cudaFree(ctx->peer.mailbox);
        if (ctx->peer.status) cudaFreeHost(ctx->peer.status);
    }
    for (Dev& d : ctx->devs) {
        cudaSetDevice(d.id);
        cudaStreamSynchronize(d.stream);
        for (auto& kv : d.plans) cudaFree(kv.second.tables);
        cudaStreamSynchronize(d.sort_stream);
        DevBuf* bufs[] = {&d.ntt_a, &d.ntt_b, &d.ntt_tab, &d.m_scalars, &d.m_buckets, &d.m_wpart, &d.m_long_partials, &d.m_ba, &d.m_counters,
                          &d.m_tmp_bases, &d.p_in, &d.p_out, &d.p_tiles, &d.q_stage, &d.q_scratch};
        for (DevBuf* b : bufs) b->release();
        d.ms[0].release(); d.ms[1].release();
        d.ring.release();
        if (d.pinned) cudaFreeHost(d.pinned);
        for (auto& r : d.recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        for (auto ev : d.ev_pool) cudaEventDestroy(ev);
        cudaEventDestroy(d.ev_first); cudaEventDestroy(d.ev_last);
        for (auto ev : d.chunk_events) cudaEventDestroy(ev);
        cudaStreamDestroy(d.copy_stream);
        cudaStreamDestroy(d.sort_stream);
        cudaStreamDestroy(d.stream);
    }
    delete ctx;
}

const char* h2b_last_error(const h2b_ctx* ctx) {
    if (!ctx) return g_init_error.c_str();
    // per-device host threads of a multi-GPU call may be writing ctx->err: hand out a per-thread copy taken under the lock
    thread_local std::string copy;
    { std::lock_guard<std::mutex> lk(ctx->err_mu); copy = ctx->err; }
    return copy.c_str();
}
const char* h2b_last_notice(const h2b_ctx* ctx) {
    if (!ctx) return "";
    thread_local std::string copy;
    { std::lock_guard<std::mutex> lk(ctx->err_mu); copy = ctx->status; }
    return copy.c_str();
}
int h2b_set_option(h2b_ctx* ctx, const char* name, const char* value) {
    if (!ctx || !name) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->tune.set(name, value)) return ctx->fail(H2B_E_INVALID, std::string("set_option: unknown option ") + name);
    return H2B_OK;
}

// ---- MSM -----------------------------------------------------------------------------------------
int h2b_bases_register(h2b_ctx* ctx, const void* bases, size_t n, uint64_t* handle_out) {
    if (!ctx || !handle_out || (!bases && n)) return ctx ? ctx->fail(H2B_E_INVALID, "bases_register: null argument") : H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    return register_bases(ctx, bases, false, n, true, handle_out);
}
int h2b_bases_register_device(h2b_ctx* ctx, const void* d_bases, size_t n, uint64_t* handle_out) {
    if (!ctx || !handle_out || (!d_bases && n)) return ctx ? ctx->fail(H2B_E_INVALID, "bases_register_device: null argument") : H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    return register_bases(ctx, d_bases, true, n, true, handle_out);
}
int h2b_bases_release(h2b_ctx* ctx, uint64_t handle) {
    if (!ctx || !ctx_is_live(ctx)) return H2B_E_INVALID;      // the context may have been destroyed first (it released every base)
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->bases.find(handle);
    if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "bases_release: unknown handle");
    release_bases(ctx, it->second);
    ctx->bases.erase(it);
    return H2B_OK;
}

static int msm_common(h2b_ctx* ctx, uint64_t handle, const void* scalars, bool on_device, size_t n, void* out96) {
    if (!ctx) return H2B_E_INVALID;
    if (!out96 || (!scalars && n)) return ctx->fail(H2B_E_INVALID, "msm: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->bases.find(handle);
    if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "msm: unknown bases handle");
    Xyzz windows[H2B_MSM_MAX_WINDOWS];
    uint32_t c = 0, nw = 0;
    int rc = msm_run(ctx, it->second, scalars, on_device, n, n, windows, &c, &nw);
    if (rc) return rc;
    msm_finish(windows, nw, c, out96);
    return H2B_OK;
}
int h2b_msm(h2b_ctx* ctx, uint64_t handle, const void* scalars, size_t n, void* out96) {
    return msm_common(ctx, handle, scalars, false, n, out96);
}
int h2b_msm_device(h2b_ctx* ctx, uint64_t handle, const void* d_scalars, size_t n, void* out96) {
    return msm_common(ctx, handle, d_scalars, true, n, out96);
}
static int msm_batch_common(h2b_ctx* ctx, uint64_t handle, const void* const* cols, bool on_device, size_t n_cols, size_t n, void* out96s) {
    if (!ctx) return H2B_E_INVALID;
    if (n_cols && (!cols || !out96s)) return ctx->fail(H2B_E_INVALID, "msm_batch: null argument");
    for (size_t k = 0; k < n_cols; k++) if (!cols[k] && n) return ctx->fail(H2B_E_INVALID, "msm_batch: null column");
    bool batched = false;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        auto it = ctx->bases.find(handle);
        if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "msm_batch: unknown bases handle");
        const Bases& B = it->second;
        if (n > B.n) return ctx->fail(H2B_E_INVALID, "msm_batch: more scalars than registered bases");
        // one bucket set per column needs the precomputed-window layout on a single device; a 4 GiB cap on the bucket sets
        batched = n_cols > 1 && n > 0 && ctx->devs.size() == 1 && B.shard[0].pre && B.shard[0].count >= n &&
                  (size_t)n_cols * ((size_t)1 << (B.shard[0].pre_c - 1)) * sizeof(Xyzz) <= ((size_t)4 << 30) &&
                  !ctx->tune.msm_no_batch;
        if (batched) {
            std::vector<Xyzz> sums(n_cols);
            uint32_t c = 0;
            int rc = msm_run_batch(ctx, B, cols, on_device, n_cols, n, sums.data(), &c);
            if (rc) return rc;
            for (size_t k = 0; k < n_cols; k++) msm_finish(&sums[k], 1, c, reinterpret_cast<uint8_t*>(out96s) + 96 * k);
            return H2B_OK;
        }
    }
    for (size_t k = 0; k < n_cols; k++) {          // fallback: one MSM per column (classic layout, several devices, ...)
        int rc = msm_common(ctx, handle, cols[k], on_device, n, reinterpret_cast<uint8_t*>(out96s) + 96 * k);
        if (rc) return rc;
    }
    return H2B_OK;
}
int h2b_msm_batch(h2b_ctx* ctx, uint64_t handle, const void* const* scalars_cols, size_t n_cols, size_t n, void* out96s) {
    return msm_batch_common(ctx, handle, scalars_cols, false, n_cols, n, out96s);
}
int h2b_msm_batch_device(h2b_ctx* ctx, uint64_t handle, const void* const* d_scalars_cols, size_t n_cols, size_t n, void* out96s) {
    return msm_batch_common(ctx, handle, d_scalars_cols, true, n_cols, n, out96s);
}
int h2b_msm_raw(h2b_ctx* ctx, const void* bases, const void* scalars, size_t n, void* out96) {
    if (!ctx) return H2B_E_INVALID;
    uint64_t h = 0;
    int rc;
    if ((!bases && n)) return ctx->fail(H2B_E_INVALID, "msm_raw: null argument");
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        rc = register_bases(ctx, bases, false, n, false, &h);     // one-shot bases: no window precomputation
    }
    if (rc) return rc;
    rc = h2b_msm(ctx, h, scalars, n, out96);
    int rc2 = h2b_bases_release(ctx, h);
    return rc ? rc : rc2;
}
int h2b_msm_windows(h2b_ctx* ctx, uint64_t handle, const void* scalars, int scalars_on_device, size_t n, size_t n_total,
                    void* out_windows, uint32_t* c_out, uint32_t* windows_out) {
    if (!ctx) return H2B_E_INVALID;
    if (!out_windows || !c_out || !windows_out || (!scalars && n)) return ctx->fail(H2B_E_INVALID, "msm_windows: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->bases.find(handle);
    if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "msm_windows: unknown bases handle");
    Xyzz windows[H2B_MSM_MAX_WINDOWS];
    int rc = msm_run(ctx, it->second, scalars, scalars_on_device != 0, n, n_total, windows, c_out, windows_out);
    if (rc) return rc;
    std::memcpy(out_windows, windows, (size_t)*windows_out * sizeof(Xyzz));
    return H2B_OK;
}
int h2b_bases_info(h2b_ctx* ctx, uint64_t handle, uint32_t* window_bits_out, uint64_t* device_bytes_out) {
    if (!ctx) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    auto it = ctx->bases.find(handle);
    if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "bases_info: unknown handle");
    uint32_t c = 0;
    uint64_t bytes = 0;
    for (const BasesShard& sh : it->second.shard) {
        if (sh.count == 0) continue;
        if (sh.owned) bytes += sh.count * 64;
        if (sh.pre) { bytes += (uint64_t)sh.count * sh.pre_w * 64; c = sh.pre_c; }
    }
    if (window_bits_out) *window_bits_out = c;
    if (device_bytes_out) *device_bytes_out = bytes;
    return H2B_OK;
}

// ---- sharded MSM across processes: mailboxes over CUDA IPC / NVLink (msm.cuh 7) --------------------------------------
int h2b_peer_export(h2b_ctx* ctx, void* handle64_out, void** local_ptr_out) {
    if (!ctx || !handle64_out) return ctx ? ctx->fail(H2B_E_INVALID, "peer_export: null argument") : H2B_E_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == H2B_PEER_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
    std::lock_guard<std::mutex> lk(ctx->mu);
    PeerState& P = ctx->peer;
    H2B_CK(cudaSetDevice(ctx->devs[0].id));
    if (!P.mailbox) {
        H2B_CK(cudaMalloc(&P.mailbox, msm_peer_mailbox_bytes()));
        H2B_CK(cudaMemset(P.mailbox, 0, msm_peer_mailbox_bytes()));
        H2B_CK(cudaMallocHost((void**)&P.status, 64));
        H2B_CK(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    H2B_CK(cudaIpcGetMemHandle(&h, P.mailbox));
    std::memcpy(handle64_out, &h, sizeof h);
    if (local_ptr_out) *local_ptr_out = P.mailbox;
    return H2B_OK;
}
int h2b_peer_attach(h2b_ctx* ctx, int rank, int world, const void* handles, void* const* local_ptrs) {
    if (!ctx) return H2B_E_INVALID;
    if (world < 1 || world > (int)MSM_PEER_MAX_RANKS || rank < 0 || rank >= world || (!handles && !local_ptrs && world > 1))
        return ctx->fail(H2B_E_INVALID, "peer_attach: bad rank / world (at most 16 ranks) or no handles");
    std::lock_guard<std::mutex> lk(ctx->mu);
    PeerState& P = ctx->peer;
    if (!P.mailbox) return ctx->fail(H2B_E_INVALID, "peer_attach: call h2b_peer_export first");
    if (ctx->devs.size() != 1) return ctx->fail(H2B_E_INVALID, "peer_attach: one device per process (a multi-device context shards by itself)");
    H2B_CK(cudaSetDevice(ctx->devs[0].id));
    for (void* q : P.opened) cudaIpcCloseMemHandle(q);
    P.opened.clear();
    P.table = MsmPeerTable{};
    for (int r = 0; r < world; r++) {
        if (r == rank) { P.table.mailbox[r] = P.mailbox; continue; }
        if (local_ptrs && local_ptrs[r]) {
            // a peer context of this process: its pointer is used as it is; make sure this device may store to that one
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, local_ptrs[r]) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != ctx->devs[0].id) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, ctx->devs[0].id, at.device) != cudaSuccess || !can)
                    return ctx->fail(H2B_E_CUDA, "peer_attach: no peer access from device " + std::to_string(ctx->devs[0].id) + " to device " + std::to_string(at.device));
                if (cudaDeviceEnablePeerAccess(at.device, 0) != cudaSuccess) cudaGetLastError();      // already enabled: fine
            }
            P.table.mailbox[r] = local_ptrs[r];
            continue;
        }
        if (!handles) return ctx->fail(H2B_E_INVALID, "peer_attach: no handle for rank " + std::to_string(r));
        cudaIpcMemHandle_t h;
        std::memcpy(&h, reinterpret_cast<const uint8_t*>(handles) + (size_t)r * H2B_PEER_HANDLE_BYTES, sizeof h);
        void* q = nullptr;
        H2B_CK(cudaIpcOpenMemHandle(&q, h, cudaIpcMemLazyEnablePeerAccess));
        P.opened.push_back(q);
        P.table.mailbox[r] = q;
    }
    // A fresh group starts at epoch 0 with an empty mailbox: flags, tags and slots left by an earlier group (or by calls that
    // failed half way) must not be taken for this group's.  Peers write into this mailbox as soon as THEY are attached and call
    // h2b_msm_sharded, hence the rule in the header: every rank finishes h2b_peer_attach before any rank calls h2b_msm_sharded.
    H2B_CK(cudaMemset(P.mailbox, 0, msm_peer_mailbox_bytes()));
    H2B_CK(cudaDeviceSynchronize());
    P.rank = rank; P.world = world; P.epoch = 0;
    // ranks attached through IPC handles are other processes on this host, each with a copy pool of its own for pageable host
    // buffers: share the host threads between them (at N = 8, eight full-size pools oversubscribe the cores and the staging
    // copies of a sharded MSM from pageable scalars take longer than the MSM)
    CopyPool::instance().set_share(1u + (unsigned)P.opened.size());
    return H2B_OK;
}
int h2b_msm_sharded(h2b_ctx* ctx, uint64_t handle, const void* scalars, int scalars_on_device, size_t n, size_t n_total, void* out96) {
    if (!ctx) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    // Every call counts, also one that is refused below or fails before its exchange: the epoch is what keeps the ranks' mailbox
    // slots in step, and a rank that skipped one would take a peer's previous terms for the current ones.
    if (ctx->peer.world > 0) ++ctx->peer.epoch;
    if (!out96 || (!scalars && n)) return ctx->fail(H2B_E_INVALID, "msm_sharded: null argument");
    auto it = ctx->bases.find(handle);
    if (it == ctx->bases.end()) return ctx->fail(H2B_E_INVALID, "msm_sharded: unknown bases handle");
    if (n == 0) return ctx->fail(H2B_E_INVALID, "msm_sharded: every rank needs a non-empty point range");
    Xyzz windows[H2B_MSM_MAX_WINDOWS];
    uint32_t c = 0, nw = 0;
    int rc = msm_run(ctx, it->second, scalars, scalars_on_device != 0, n, n_total, windows, &c, &nw, true);
    if (rc) return rc;
    msm_finish(windows, nw, c, out96);
    return H2B_OK;
}
int h2b_msm_combine(const void* parts, uint32_t n_parts, uint32_t c, uint32_t windows, void* out96) {
    if (!parts || !out96 || windows == 0 || windows > H2B_MSM_MAX_WINDOWS || c == 0 || c > 31) return H2B_E_INVALID;
    Xyzz acc[H2B_MSM_MAX_WINDOWS];
    const Xyzz* p = reinterpret_cast<const Xyzz*>(parts);
    for (uint32_t j = 0; j < windows; j++) {
        acc[j] = Xyzz::identity();
        for (uint32_t r = 0; r < n_parts; r++) {
            Xyzz t;
            std::memcpy(&t, p + (size_t)r * windows + j, sizeof(Xyzz));
            acc[j] = xyzz_add(acc[j], t);
        }
    }
    msm_finish(acc, windows, c, out96);
    return H2B_OK;
}

// ---- NTT -----------------------------------------------------------------------------------------
static int ntt_args_ok(h2b_ctx* ctx, uint32_t log_n) {
    if (log_n < 1 || log_n > 28) return ctx->fail(H2B_E_INVALID, "ntt: log_n out of range [1, 28] (Fr::S = 28, domain.rs:54)");
    return H2B_OK;
}

int h2b_ntt_device(h2b_ctx* ctx, void* d_a, uint32_t log_n, const void* omega32, const void* scale32) {
    if (!ctx) return H2B_E_INVALID;
    if (!d_a || !omega32) return ctx->fail(H2B_E_INVALID, "ntt: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, log_n);
    if (rc) return rc;
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    size_t N = (size_t)1 << log_n;
    H2B_CK(d.ntt_b.reserve(N * 32));
    Fr omega = fr_from_bytes(omega32), scale;
    if (scale32) scale = fr_from_bytes(scale32);
    begin_call(d);
    NttFused fz;
    // a -> (column passes) -> b -> (row pass) -> a ; single-pass sizes run a -> a inside one tile
    rc = ntt_enqueue(ctx, d, (Fr*)d_a, 0, d.ntt_b.as<Fr>(), 0, (Fr*)d_a, 0, 1, log_n, omega, scale32 ? &scale : nullptr, fz);
    if (rc) return rc;
    return end_call(ctx, d);
}

int h2b_ntt(h2b_ctx* ctx, void* a, uint32_t log_n, const void* omega32, const void* scale32) {
    if (!ctx) return H2B_E_INVALID;
    if (!a || !omega32) return ctx->fail(H2B_E_INVALID, "ntt: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, log_n);
    if (rc) return rc;
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    size_t N = (size_t)1 << log_n;
    H2B_CK(d.ntt_a.reserve(N * 32));
    H2B_CK(d.ntt_b.reserve(N * 32));
    Fr omega = fr_from_bytes(omega32), scale;
    if (scale32) scale = fr_from_bytes(scale32);
    begin_call(d);
    H2B_CK(d.ring.h2d(d.ntt_a.p, a, N * 32, d.stream));
    NttFused fz;
    rc = ntt_enqueue(ctx, d, d.ntt_a.as<Fr>(), 0, d.ntt_b.as<Fr>(), 0, d.ntt_a.as<Fr>(), 0, 1, log_n, omega,
                     scale32 ? &scale : nullptr, fz);
    if (rc) return rc;
    H2B_CK(d.ring.d2h(a, d.ntt_a.p, N * 32, d.stream));
    return end_call(ctx, d);
}

static int c2e_impl(h2b_ctx* ctx, const void* in_n, void* out_ne, bool on_device, uint32_t k, uint32_t ext_k,
                    const void* ext_omega32, const void* g_coset32, const void* g_coset_inv32) {
    if (!ctx) return H2B_E_INVALID;
    if (!in_n || !out_ne || !ext_omega32 || !g_coset32 || !g_coset_inv32) return ctx->fail(H2B_E_INVALID, "coeff_to_extended: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, ext_k);
    if (rc) return rc;
    if (k > ext_k) return ctx->fail(H2B_E_INVALID, "coeff_to_extended: k > extended_k");
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    size_t n = (size_t)1 << k, N = (size_t)1 << ext_k;
    Fr omega = fr_from_bytes(ext_omega32);
    Fr pat[2] = {fr_from_bytes(g_coset32), fr_from_bytes(g_coset_inv32)};
    NttFused fz;
    fz.in_len = n;
    fz.in_pat = pat;
    H2B_CK(d.ntt_b.reserve(N * 32));
    begin_call(d);
    const Fr* src;
    Fr* dst;
    if (on_device) { src = (const Fr*)in_n; dst = (Fr*)out_ne; }
    else {
        H2B_CK(d.ntt_a.reserve(N * 32));
        // stage the n inputs at the END of ntt_a's N slots?  No: the row pass writes `dst` while reading `tmp` only,
        // and the first pass has consumed `src` by then, so src may alias dst.
        H2B_CK(d.ring.h2d(d.ntt_a.p, in_n, n * 32, d.stream));
        src = d.ntt_a.as<Fr>(); dst = d.ntt_a.as<Fr>();
    }
    rc = ntt_enqueue(ctx, d, src, 0, d.ntt_b.as<Fr>(), 0, dst, 0, 1, ext_k, omega, nullptr, fz);
    if (rc) return rc;
    if (!on_device) H2B_CK(d.ring.d2h(out_ne, dst, N * 32, d.stream));
    return end_call(ctx, d);
}
int h2b_coeff_to_extended(h2b_ctx* ctx, const void* in_n, void* out_ne, uint32_t k, uint32_t ext_k, const void* ext_omega32,
                          const void* g_coset32, const void* g_coset_inv32) {
    return c2e_impl(ctx, in_n, out_ne, false, k, ext_k, ext_omega32, g_coset32, g_coset_inv32);
}
int h2b_coeff_to_extended_device(h2b_ctx* ctx, const void* d_in_n, void* d_out_ne, uint32_t k, uint32_t ext_k,
                                 const void* ext_omega32, const void* g_coset32, const void* g_coset_inv32) {
    return c2e_impl(ctx, d_in_n, d_out_ne, true, k, ext_k, ext_omega32, g_coset32, g_coset_inv32);
}

static int e2c_impl(h2b_ctx* ctx, const void* a_ne, void* out, bool on_device, uint32_t ext_k, const void* ext_omega_inv32,
                    const void* ext_ifft_divisor32, const void* g_coset32, const void* g_coset_inv32, size_t out_len,
                    const void* t_evals32, size_t n_t) {
    if (!ctx) return H2B_E_INVALID;
    if (!a_ne || !out || !ext_omega_inv32 || !ext_ifft_divisor32 || !g_coset32 || !g_coset_inv32)
        return ctx->fail(H2B_E_INVALID, "extended_to_coeff: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, ext_k);
    if (rc) return rc;
    size_t N = (size_t)1 << ext_k;
    if (out_len > N) return ctx->fail(H2B_E_INVALID, "extended_to_coeff: out_len exceeds the extended domain");
    if (t_evals32 && (n_t == 0 || (n_t & (n_t - 1)) != 0 || n_t > N)) return ctx->fail(H2B_E_INVALID, "extended_to_coeff: t_evals length must be a power of two");
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    Fr omega = fr_from_bytes(ext_omega_inv32), scale = fr_from_bytes(ext_ifft_divisor32);
    Fr pat[2] = {fr_from_bytes(g_coset_inv32), fr_from_bytes(g_coset32)};   // out of the coset: [1, zeta^2, zeta]
    NttFused fz;
    fz.out_len = out_len;
    fz.out_pat = pat;
    H2B_CK(d.ntt_b.reserve(N * 32));
    begin_call(d);
    if (t_evals32) {
        H2B_CK(d.ntt_tab.reserve(n_t * 32));
        H2B_CK(d.ring.h2d(d.ntt_tab.p, t_evals32, n_t * 32, d.stream));
        fz.in_tab = d.ntt_tab.as<Fr>();
        fz.in_tab_mask = (uint32_t)(n_t - 1);
    }
    const Fr* src;
    Fr* dst;
    if (on_device) { src = (const Fr*)a_ne; dst = (Fr*)out; }
    else {
        H2B_CK(d.ntt_a.reserve(N * 32));
        H2B_CK(d.ring.h2d(d.ntt_a.p, a_ne, N * 32, d.stream));
        src = d.ntt_a.as<Fr>(); dst = d.ntt_a.as<Fr>();
    }
    rc = ntt_enqueue(ctx, d, src, 0, d.ntt_b.as<Fr>(), 0, dst, 0, 1, ext_k, omega, &scale, fz);
    if (rc) return rc;
    if (!on_device) H2B_CK(d.ring.d2h(out, dst, out_len * 32, d.stream));
    return end_call(ctx, d);
}
int h2b_extended_to_coeff(h2b_ctx* ctx, const void* a_ne, void* out, uint32_t ext_k, const void* ext_omega_inv32,
                          const void* ext_ifft_divisor32, const void* g_coset32, const void* g_coset_inv32, size_t out_len,
                          const void* t_evals32, size_t n_t) {
    return e2c_impl(ctx, a_ne, out, false, ext_k, ext_omega_inv32, ext_ifft_divisor32, g_coset32, g_coset_inv32, out_len, t_evals32, n_t);
}
int h2b_extended_to_coeff_device(h2b_ctx* ctx, const void* d_a_ne, void* d_out, uint32_t ext_k, const void* ext_omega_inv32,
                                 const void* ext_ifft_divisor32, const void* g_coset32, const void* g_coset_inv32,
                                 size_t out_len, const void* t_evals32, size_t n_t) {
    return e2c_impl(ctx, d_a_ne, d_out, true, ext_k, ext_omega_inv32, ext_ifft_divisor32, g_coset32, g_coset_inv32, out_len, t_evals32, n_t);
}

// Batched transforms over host columns: columns are dealt round-robin over the devices; on each device the
// columns are staged contiguously and transformed by ONE launch per pass (grid.y = column).
static int batch_impl(h2b_ctx* ctx, const void* const* in_cols, void* const* out_cols, size_t n_cols, uint32_t log_in,
                      uint32_t log_n, const Fr& omega, const Fr* scale, const NttFused& fz) {
    size_t nin = (size_t)1 << log_in, N = (size_t)1 << log_n;
    size_t nd = ctx->devs.size();
    double max_ms = 0;
    // bound the staging footprint: process in rounds of at most `chunk` columns per device
    size_t chunk = std::max<size_t>(1, ((size_t)8 << 30) / (N * 32));
    for (size_t base = 0; base < n_cols; base += chunk * nd) {
        std::vector<std::vector<size_t>> mine(nd);
        for (size_t c = base; c < std::min(n_cols, base + chunk * nd); c++) mine[(c - base) % nd].push_back(c);
        // one host thread per device: staged copies of pageable columns block their caller (hostcopy.hpp)
        auto run_dev = [&](size_t di) -> int {
            Dev& d = ctx->devs[di];
            size_t m = mine[di].size();
            H2B_CK(cudaSetDevice(d.id));
            H2B_CK(d.ntt_a.reserve(m * N * 32));
            H2B_CK(d.ntt_b.reserve(m * N * 32));
            begin_call(d);
            for (size_t i = 0; i < m; i++)
                H2B_CK(d.ring.h2d(d.ntt_a.as<Fr>() + i * N, in_cols[mine[di][i]], nin * 32, d.stream));
            int rc = ntt_enqueue(ctx, d, d.ntt_a.as<Fr>(), N, d.ntt_b.as<Fr>(), N, d.ntt_a.as<Fr>(), N, (uint32_t)m, log_n, omega, scale, fz);
            if (rc) return rc;
            for (size_t i = 0; i < m; i++)
                H2B_CK(d.ring.d2h(out_cols[mine[di][i]], d.ntt_a.as<Fr>() + i * N, N * 32, d.stream));
            H2B_CK(cudaEventRecord(d.ev_last, d.stream));
            return H2B_OK;
        };
        std::vector<size_t> active;
        for (size_t di = 0; di < nd; di++) if (!mine[di].empty()) active.push_back(di);
        if (active.size() == 1) {
            int rc = run_dev(active[0]);
            if (rc) return rc;
        } else {
            std::vector<int> rcs(active.size(), H2B_OK);
            std::vector<std::thread> th;
            for (size_t u = 0; u < active.size(); u++) th.emplace_back([&, u] { rcs[u] = run_dev(active[u]); });
            for (auto& t : th) t.join();
            for (int rc : rcs) if (rc) return rc;
        }
        for (size_t di = 0; di < nd; di++) {
            if (mine[di].empty()) continue;
            Dev& d = ctx->devs[di];
            H2B_CK(cudaSetDevice(d.id));
            H2B_CK(cudaStreamSynchronize(d.stream));
            float ms = 0.f;
            H2B_CK(cudaEventElapsedTime(&ms, d.ev_first, d.ev_last));
            max_ms = std::max(max_ms, (double)ms);
        }
    }
    ctx->last_ms = max_ms;
    return H2B_OK;
}
int h2b_ntt_batch(h2b_ctx* ctx, void* const* cols, size_t n_cols, uint32_t log_n, const void* omega32, const void* scale32) {
    if (!ctx) return H2B_E_INVALID;
    if ((!cols && n_cols) || !omega32) return ctx->fail(H2B_E_INVALID, "ntt_batch: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, log_n);
    if (rc) return rc;
    Fr omega = fr_from_bytes(omega32), scale;
    if (scale32) scale = fr_from_bytes(scale32);
    NttFused fz;
    return batch_impl(ctx, (const void* const*)cols, cols, n_cols, log_n, log_n, omega, scale32 ? &scale : nullptr, fz);
}
int h2b_coeff_to_extended_batch(h2b_ctx* ctx, const void* const* in_cols, void* const* out_cols, size_t n_cols, uint32_t k,
                                uint32_t ext_k, const void* ext_omega32, const void* g_coset32, const void* g_coset_inv32) {
    if (!ctx) return H2B_E_INVALID;
    if ((n_cols && (!in_cols || !out_cols)) || !ext_omega32 || !g_coset32 || !g_coset_inv32)
        return ctx->fail(H2B_E_INVALID, "coeff_to_extended_batch: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, ext_k);
    if (rc) return rc;
    if (k > ext_k) return ctx->fail(H2B_E_INVALID, "coeff_to_extended_batch: k > extended_k");
    Fr omega = fr_from_bytes(ext_omega32);
    Fr pat[2] = {fr_from_bytes(g_coset32), fr_from_bytes(g_coset_inv32)};
    NttFused fz;
    fz.in_len = (uint64_t)1 << k;
    fz.in_pat = pat;
    return batch_impl(ctx, in_cols, out_cols, n_cols, k, ext_k, omega, nullptr, fz);
}

// Device-resident columns (all on device 0 of the context), extended on every device of the context: column c goes to device
// c mod G; a remote device pulls its n inputs over NVLink, transforms in its own buffers and pushes the n*e evaluations back into
// the caller's output buffer on device 0 (peer copies on the device's stream).  The independent coset NTTs of evaluate_h
// (plonk/evaluation.rs:344-361, 497-505, 571) are exactly this shape.  No collective: every column is independent.
int h2b_coeff_to_extended_batch_device(h2b_ctx* ctx, const void* const* d_in_cols, void* const* d_out_cols, size_t n_cols,
                                       uint32_t k, uint32_t ext_k, const void* ext_omega32, const void* g_coset32,
                                       const void* g_coset_inv32) {
    if (!ctx) return H2B_E_INVALID;
    if ((n_cols && (!d_in_cols || !d_out_cols)) || !ext_omega32 || !g_coset32 || !g_coset_inv32)
        return ctx->fail(H2B_E_INVALID, "coeff_to_extended_batch_device: null argument");
    for (size_t c = 0; c < n_cols; c++)
        if (!d_in_cols[c] || !d_out_cols[c]) return ctx->fail(H2B_E_INVALID, "coeff_to_extended_batch_device: null column");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = ntt_args_ok(ctx, ext_k);
    if (rc) return rc;
    if (k > ext_k) return ctx->fail(H2B_E_INVALID, "coeff_to_extended_batch_device: k > extended_k");
    if (!n_cols) { ctx->last_ms = 0; return H2B_OK; }
    const size_t n = (size_t)1 << k, N = (size_t)1 << ext_k, nd = ctx->devs.size();
    const Fr omega = fr_from_bytes(ext_omega32);
    Fr pat[2] = {fr_from_bytes(g_coset32), fr_from_bytes(g_coset_inv32)};
    const int dev0 = ctx->devs[0].id;
    auto run_dev = [&](size_t di) -> int {
        Dev& d = ctx->devs[di];
        H2B_CK(cudaSetDevice(d.id));
        H2B_CK(d.ntt_b.reserve(N * 32));
        if (di) H2B_CK(d.ntt_a.reserve(N * 32));
        NttFused fz;
        fz.in_len = n;
        fz.in_pat = pat;
        begin_call(d);
        for (size_t c = di; c < n_cols; c += nd) {
            if (di == 0) {
                int r = ntt_enqueue(ctx, d, (const Fr*)d_in_cols[c], 0, d.ntt_b.as<Fr>(), 0, (Fr*)d_out_cols[c], 0, 1, ext_k, omega, nullptr, fz);
                if (r) return r;
            } else {
                H2B_CK(cudaMemcpyPeerAsync(d.ntt_a.p, d.id, d_in_cols[c], dev0, n * 32, d.stream));
                int r = ntt_enqueue(ctx, d, d.ntt_a.as<Fr>(), 0, d.ntt_b.as<Fr>(), 0, d.ntt_a.as<Fr>(), 0, 1, ext_k, omega, nullptr, fz);
                if (r) return r;
                H2B_CK(cudaMemcpyPeerAsync(d_out_cols[c], dev0, d.ntt_a.p, d.id, N * 32, d.stream));
            }
        }
        H2B_CK(cudaEventRecord(d.ev_last, d.stream));
        return H2B_OK;
    };
    const size_t active = std::min(nd, n_cols);
    if (active == 1) {
        rc = run_dev(0);
        if (rc) return rc;
    } else {
        std::vector<int> rcs(active, H2B_OK);
        std::vector<std::thread> th;
        for (size_t u = 0; u < active; u++) th.emplace_back([&, u] { rcs[u] = run_dev(u); });
        for (auto& t : th) t.join();
        for (int r : rcs) if (r) return r;
    }
    double max_ms = 0;
    for (size_t di = 0; di < active; di++) {
        Dev& d = ctx->devs[di];
        H2B_CK(cudaSetDevice(d.id));
        H2B_CK(cudaStreamSynchronize(d.stream));
        float ms = 0.f;
        H2B_CK(cudaEventElapsedTime(&ms, d.ev_first, d.ev_last));
        max_ms = std::max(max_ms, (double)ms);
    }
    H2B_CK(cudaSetDevice(dev0));
    ctx->last_ms = max_ms;
    return H2B_OK;
}

// ---- polynomial helpers -----------------------------------------------------------
static int batch_invert_impl(h2b_ctx* ctx, void* a, size_t n, bool on_device) {
    if (!ctx) return H2B_E_INVALID;
    if (!a && n) return ctx->fail(H2B_E_INVALID, "batch_invert: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    Fr* buf = (Fr*)a;
    if (!on_device && n) {
        H2B_CK(d.p_in.reserve(n * 32));
        buf = d.p_in.as<Fr>();
        H2B_CK(d.ring.h2d(buf, a, n * 32, d.stream));
    }
    int rc = poly_batch_invert_enqueue(ctx, d, buf, n);
    if (rc) return rc;
    if (!on_device && n) H2B_CK(d.ring.d2h(a, buf, n * 32, d.stream));
    return end_call(ctx, d);
}
int h2b_fr_mul_device(h2b_ctx* ctx, const void* d_a, const void* d_b, const void* scale32_or_null, void* d_out, size_t n) {
    if (!ctx) return H2B_E_INVALID;
    if ((!d_a || !d_b || !d_out) && n) return ctx->fail(H2B_E_INVALID, "fr_mul: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    if (n) {
        const Fr scale = scale32_or_null ? fr_from_bytes(scale32_or_null) : Fr::zero();
        const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, (size_t)d.sms * 16);
        H2B_LAUNCH(d, "poly_mul", poly_kernel_table().mul, grid, 256, 0, (const Fr*)d_a, (const Fr*)d_b, scale,
                   scale32_or_null ? 1u : 0u, (Fr*)d_out, (uint64_t)n);
        H2B_CK(cudaGetLastError());
    }
    return end_call(ctx, d);
}
int h2b_batch_invert(h2b_ctx* ctx, void* a, size_t n) { return batch_invert_impl(ctx, a, n, false); }
int h2b_batch_invert_device(h2b_ctx* ctx, void* d_a, size_t n) { return batch_invert_impl(ctx, d_a, n, true); }

static int grand_product_impl(h2b_ctx* ctx, const void* v, size_t n, const void* init32, void* z, size_t out_len, bool on_device) {
    if (!ctx) return H2B_E_INVALID;
    if ((!v && n) || !init32 || (!z && out_len)) return ctx->fail(H2B_E_INVALID, "grand_product: null argument");
    if (out_len > n + 1) return ctx->fail(H2B_E_INVALID, "grand_product: out_len exceeds n + 1");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    Fr init = fr_from_bytes(init32);
    const Fr* src = (const Fr*)v;
    Fr* dst = (Fr*)z;
    if (!on_device) {
        H2B_CK(d.p_in.reserve(std::max<size_t>(n, 1) * 32));
        H2B_CK(d.p_out.reserve(std::max<size_t>(out_len, 1) * 32));
        if (n) H2B_CK(d.ring.h2d(d.p_in.p, v, n * 32, d.stream));
        src = d.p_in.as<Fr>(); dst = d.p_out.as<Fr>();
    }
    int rc = poly_grand_product_enqueue(ctx, d, src, n, init, dst, out_len);
    if (rc) return rc;
    if (!on_device && out_len) H2B_CK(d.ring.d2h(z, dst, out_len * 32, d.stream));
    return end_call(ctx, d);
}
int h2b_grand_product(h2b_ctx* ctx, const void* v, size_t n, const void* init32, void* z, size_t out_len) {
    return grand_product_impl(ctx, v, n, init32, z, out_len, false);
}
int h2b_grand_product_device(h2b_ctx* ctx, const void* d_v, size_t n, const void* init32, void* d_z, size_t out_len) {
    return grand_product_impl(ctx, d_v, n, init32, d_z, out_len, true);
}

static int horner_impl(h2b_ctx* ctx, const void* a, size_t n, const void* b32, void* q, void* out32, bool on_device, const char* what) {
    if (!ctx) return H2B_E_INVALID;
    if ((!a && n) || !b32 || (!q && !out32)) return ctx->fail(H2B_E_INVALID, std::string(what) + ": null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (q && n == 0) return ctx->fail(H2B_E_INVALID, "kate_division: empty polynomial (the reference panics on a.len() - 1)");
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    if (n == 0) { std::memset(out32, 0, 32); ctx->last_ms = 0; return H2B_OK; }          // empty sum
    begin_call(d);
    Fr b = fr_from_bytes(b32);
    const Fr* src = (const Fr*)a;
    Fr* dq = (Fr*)q;
    H2B_CK(d.m_counters.reserve(256));
    if (!on_device) {
        H2B_CK(d.p_in.reserve(n * 32));
        H2B_CK(d.ring.h2d(d.p_in.p, a, n * 32, d.stream));
        src = d.p_in.as<Fr>();
        if (q) { H2B_CK(d.p_out.reserve(std::max<size_t>(n - 1, 1) * 32)); dq = d.p_out.as<Fr>(); }
    }
    Fr* d_final = out32 ? d.m_counters.as<Fr>() : nullptr;
    int rc = poly_horner_enqueue(ctx, d, src, n, b, q ? dq : nullptr, d_final);
    if (rc) return rc;
    if (!on_device && q && n > 1) H2B_CK(d.ring.d2h(q, dq, (n - 1) * 32, d.stream));
    if (out32) H2B_CK(d.ring.d2h(out32, d_final, 32, d.stream));
    return end_call(ctx, d);
}
int h2b_eval_polynomial(h2b_ctx* ctx, const void* poly, size_t n, const void* point32, void* out32) {
    if (ctx && !out32) return ctx->fail(H2B_E_INVALID, "eval_polynomial: null argument");
    return horner_impl(ctx, poly, n, point32, nullptr, out32, false, "eval_polynomial");
}
int h2b_eval_polynomial_device(h2b_ctx* ctx, const void* d_poly, size_t n, const void* point32, void* out32) {
    if (ctx && !out32) return ctx->fail(H2B_E_INVALID, "eval_polynomial: null argument");
    return horner_impl(ctx, d_poly, n, point32, nullptr, out32, true, "eval_polynomial");
}
// count independent evaluations of length-n device polynomials in two launches and one result copy
int h2b_eval_polynomial_batch_device(h2b_ctx* ctx, const void* const* d_polys, size_t n, const void* points32, size_t count,
                                     void* out32s) {
    if (!ctx) return H2B_E_INVALID;
    if (count && (!d_polys || !points32 || !out32s)) return ctx->fail(H2B_E_INVALID, "eval_polynomial_batch: null argument");
    for (size_t i = 0; i < count; i++)
        if (!d_polys[i] && n) return ctx->fail(H2B_E_INVALID, "eval_polynomial_batch: null polynomial");
    if (count > 65535) return ctx->fail(H2B_E_INVALID, "eval_polynomial_batch: more than 65535 evaluations in one call");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    if (count == 0) { ctx->last_ms = 0; return H2B_OK; }
    if (n == 0) { std::memset(out32s, 0, count * 32); ctx->last_ms = 0; return H2B_OK; }          // empty sums
    const uint32_t tiles = poly_tiles(n);
    const uint64_t pad = (uint64_t)tiles * POLY_TILE - n;
    std::vector<HornerJob> jobs(count);
    for (size_t i = 0; i < count; i++) {
        HornerJob& J = jobs[i];
        J.a = (const Fr*)d_polys[i];
        J.reserved = 0;
        J.b = fr_from_bytes((const uint8_t*)points32 + 32 * i);
        J.bk = fe_pow_u64(J.b, (uint64_t)POLY_K);
        J.bt = fe_pow_u64(J.b, (uint64_t)POLY_TILE);
    }
    const size_t jobs_bytes = count * sizeof(HornerJob);
    H2B_CK(d.q_stage.reserve(jobs_bytes + count * 32 + 64));
    H2B_CK(d.p_tiles.reserve((size_t)tiles * count * 32 + 64));
    begin_call(d);
    H2B_CK(d.ring.h2d(d.q_stage.p, jobs.data(), jobs_bytes, d.stream));
    const HornerJob* dj = d.q_stage.as<HornerJob>();
    Fr* d_out = (Fr*)(d.q_stage.as<uint8_t>() + ((jobs_bytes + 31) & ~(size_t)31));
    Fr* tv = d.p_tiles.as<Fr>();
    H2B_LAUNCH(d, "poly_horner_batch", PK.horner_batch_tile_reduce, dim3(tiles, (unsigned)count), POLY_T, 0, dj, (uint64_t)n, pad, tiles, tv);
    H2B_LAUNCH(d, "poly_horner_batch", PK.horner_batch_totals, (unsigned)count, POLY_T, 0, dj, (const Fr*)tv, tiles, d_out);
    H2B_CK(cudaGetLastError());
    H2B_CK(d.ring.d2h(out32s, d_out, count * 32, d.stream));
    return end_call(ctx, d);       // synchronises: the job vector may go out of scope
}
int h2b_kate_division(h2b_ctx* ctx, const void* a, size_t n, const void* b32, void* q) {
    if (ctx && !q) return ctx->fail(H2B_E_INVALID, "kate_division: null argument");
    return horner_impl(ctx, a, n, b32, q, nullptr, false, "kate_division");
}
int h2b_kate_division_device(h2b_ctx* ctx, const void* d_a, size_t n, const void* b32, void* d_q) {
    if (ctx && !d_q) return ctx->fail(H2B_E_INVALID, "kate_division: null argument");
    return horner_impl(ctx, d_a, n, b32, d_q, nullptr, true, "kate_division");
}

// ---- quotient evaluation ---------------------------------------------------------
int h2b_quotient_graph_device(h2b_ctx* ctx, const h2b_graph* g, const void* const* d_fixed, size_t n_fixed,
                              const void* const* d_advice, size_t n_advice, const void* const* d_instance, size_t n_instance,
                              const void* challenges32, size_t n_challenges, const void* beta32, const void* gamma32,
                              const void* theta32, const void* y32, const void* d_prev, void* d_out, uint32_t k, uint32_t ext_k) {
    if (!ctx) return H2B_E_INVALID;
    if (!g || !d_out || !beta32 || !gamma32 || !theta32 || !y32 || (g->n_words && !g->program) || (g->n_constants && !g->constants32) ||
        (g->n_rotations && !g->rotations) || (n_fixed && !d_fixed) || (n_advice && !d_advice) || (n_instance && !d_instance) ||
        (n_challenges && !challenges32))
        return ctx->fail(H2B_E_INVALID, "quotient_graph: null argument");
    if (k > ext_k || ext_k > 28) return ctx->fail(H2B_E_INVALID, "quotient_graph: bad domain sizes");
    std::lock_guard<std::mutex> lk(ctx->mu);
    uint32_t last_target = 0;
    bool has_calcs = false;
    int rc = quotient_validate(ctx, g, n_fixed, n_advice, n_instance, n_challenges, &last_target, &has_calcs);
    if (rc) return rc;
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    const uint32_t size = 1u << ext_k;
    const unsigned threads = 128;
    unsigned blocks = std::min<unsigned>((size + threads - 1) / threads, (unsigned)d.sms * 8);
    const size_t nthreads = (size_t)blocks * threads;
    Stage st;
    size_t o_prog = st.put(g->program, g->n_words * 4), o_const = st.put(g->constants32, g->n_constants * 32),
           o_rot = st.put(g->rotations, g->n_rotations * 4), o_fixed = st.put(d_fixed, n_fixed * sizeof(void*)),
           o_adv = st.put(d_advice, n_advice * sizeof(void*)), o_inst = st.put(d_instance, n_instance * sizeof(void*)),
           o_chal = st.put(challenges32, n_challenges * 32);
    H2B_CK(d.q_stage.reserve(st.host.size() + 64));
    H2B_CK(d.q_scratch.reserve(std::max<size_t>((size_t)g->n_intermediates * nthreads * 32, 256)));
    begin_call(d);
    H2B_CK(d.ring.h2d(d.q_stage.p, st.host.data(), st.host.size(), d.stream));
    uint8_t* base = d.q_stage.as<uint8_t>();
    GraphArgs A;
    A.prog = (const uint32_t*)(base + o_prog); A.n_words = (uint32_t)g->n_words; A.n_rot = (uint32_t)g->n_rotations;
    A.n_inter = g->n_intermediates; A.last_target = last_target; A.has_calcs = has_calcs ? 1 : 0;
    A.constants = (const Fr*)(base + o_const); A.rotations = (const int32_t*)(base + o_rot);
    A.fixed = (const Fr* const*)(base + o_fixed); A.advice = (const Fr* const*)(base + o_adv); A.instance = (const Fr* const*)(base + o_inst);
    A.challenges = (const Fr*)(base + o_chal);
    A.beta = fr_from_bytes(beta32); A.gamma = fr_from_bytes(gamma32); A.theta = fr_from_bytes(theta32); A.y = fr_from_bytes(y32);
    A.prev = (const Fr*)d_prev; A.out = (Fr*)d_out; A.scratch = d.q_scratch.as<Fr>();
    A.size = size; A.rot_scale = 1u << (ext_k - k);
    H2B_LAUNCH(d, "quotient_graph", QK.graph, blocks, threads, 0, A);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);       // synchronises: the host staging vector may go out of scope
}

int h2b_quotient_permutation_device(h2b_ctx* ctx, void* d_values, const void* const* d_sets, size_t n_sets,
                                    const void* const* d_col_values, const void* const* d_col_perms, size_t n_cols, size_t chunk_len,
                                    const void* d_l0, const void* d_l_last, const void* d_l_active, const void* beta32,
                                    const void* gamma32, const void* y32, const void* delta_start32, const void* delta32,
                                    const void* ext_omega32, int32_t last_rotation, uint32_t k, uint32_t ext_k) {
    if (!ctx) return H2B_E_INVALID;
    if (!d_values || !d_l0 || !d_l_last || !d_l_active || !beta32 || !gamma32 || !y32 || !delta_start32 || !delta32 || !ext_omega32 ||
        (n_sets && !d_sets) || (n_cols && (!d_col_values || !d_col_perms)))
        return ctx->fail(H2B_E_INVALID, "quotient_permutation: null argument");
    if (k > ext_k || ext_k > 28) return ctx->fail(H2B_E_INVALID, "quotient_permutation: bad domain sizes");
    if (n_sets == 0) { ctx->last_ms = 0; return H2B_OK; }                       // `if !sets.is_empty()` (evaluation.rs:397)
    if (chunk_len == 0 || n_sets * chunk_len < n_cols || (n_sets - 1) * chunk_len >= std::max<size_t>(n_cols, 1))
        return ctx->fail(H2B_E_INVALID, "quotient_permutation: n_sets does not match columns.chunks(chunk_len)");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    Stage st;
    size_t o_sets = st.put(d_sets, n_sets * sizeof(void*)), o_cv = st.put(d_col_values, n_cols * sizeof(void*)),
           o_cp = st.put(d_col_perms, n_cols * sizeof(void*));
    H2B_CK(d.q_stage.reserve(st.host.size() + 64));
    begin_call(d);
    H2B_CK(d.ring.h2d(d.q_stage.p, st.host.data(), st.host.size(), d.stream));
    uint8_t* base = d.q_stage.as<uint8_t>();
    PermFoldArgs A;
    A.values = (Fr*)d_values; A.sets = (const Fr* const*)(base + o_sets); A.col_values = (const Fr* const*)(base + o_cv);
    A.col_perms = (const Fr* const*)(base + o_cp);
    A.n_sets = (uint32_t)n_sets; A.n_cols = (uint32_t)n_cols; A.chunk_len = (uint32_t)chunk_len;
    A.l0 = (const Fr*)d_l0; A.l_last = (const Fr*)d_l_last; A.l_active = (const Fr*)d_l_active;
    A.beta = fr_from_bytes(beta32); A.gamma = fr_from_bytes(gamma32); A.y = fr_from_bytes(y32);
    A.delta_start = fr_from_bytes(delta_start32); A.delta = fr_from_bytes(delta32); A.ext_omega = fr_from_bytes(ext_omega32);
    A.one = Fr::one(); A.last_rotation = last_rotation; A.size = 1u << ext_k; A.rot_scale = 1u << (ext_k - k);
    const unsigned threads = 128;
    unsigned blocks = std::min<unsigned>((A.size + threads - 1) / threads, (unsigned)d.sms * 8);
    H2B_LAUNCH(d, "quotient_permutation", QK.permutation, blocks, threads, 0, A);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

int h2b_quotient_lookup_device(h2b_ctx* ctx, void* d_values, const void* d_table_values, const void* d_product,
                               const void* d_permuted_input, const void* d_permuted_table, const void* d_l0, const void* d_l_last,
                               const void* d_l_active, const void* beta32, const void* gamma32, const void* y32, uint32_t k,
                               uint32_t ext_k) {
    if (!ctx) return H2B_E_INVALID;
    if (!d_values || !d_table_values || !d_product || !d_permuted_input || !d_permuted_table || !d_l0 || !d_l_last || !d_l_active ||
        !beta32 || !gamma32 || !y32)
        return ctx->fail(H2B_E_INVALID, "quotient_lookup: null argument");
    if (k > ext_k || ext_k > 28) return ctx->fail(H2B_E_INVALID, "quotient_lookup: bad domain sizes");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    LookupFoldArgs A;
    A.values = (Fr*)d_values; A.table_values = (const Fr*)d_table_values; A.product = (const Fr*)d_product;
    A.permuted_input = (const Fr*)d_permuted_input; A.permuted_table = (const Fr*)d_permuted_table;
    A.l0 = (const Fr*)d_l0; A.l_last = (const Fr*)d_l_last; A.l_active = (const Fr*)d_l_active;
    A.beta = fr_from_bytes(beta32); A.gamma = fr_from_bytes(gamma32); A.y = fr_from_bytes(y32); A.one = Fr::one();
    A.size = 1u << ext_k; A.rot_scale = 1u << (ext_k - k);
    const unsigned threads = 128;
    unsigned blocks = std::min<unsigned>((A.size + threads - 1) / threads, (unsigned)d.sms * 16);
    begin_call(d);
    H2B_LAUNCH(d, "quotient_lookup", QK.lookup, blocks, threads, 0, A);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

int h2b_quotient_shuffle_device(h2b_ctx* ctx, void* d_values, const void* d_input_values, const void* d_shuffle_values,
                                const void* d_product, const void* d_l0, const void* d_l_last, const void* d_l_active,
                                const void* y32, uint32_t k, uint32_t ext_k) {
    if (!ctx) return H2B_E_INVALID;
    if (!d_values || !d_input_values || !d_shuffle_values || !d_product || !d_l0 || !d_l_last || !d_l_active || !y32)
        return ctx->fail(H2B_E_INVALID, "quotient_shuffle: null argument");
    if (k > ext_k || ext_k > 28) return ctx->fail(H2B_E_INVALID, "quotient_shuffle: bad domain sizes");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    ShuffleFoldArgs A;
    A.values = (Fr*)d_values; A.input_values = (const Fr*)d_input_values; A.shuffle_values = (const Fr*)d_shuffle_values;
    A.product = (const Fr*)d_product; A.l0 = (const Fr*)d_l0; A.l_last = (const Fr*)d_l_last; A.l_active = (const Fr*)d_l_active;
    A.y = fr_from_bytes(y32); A.one = Fr::one(); A.size = 1u << ext_k; A.rot_scale = 1u << (ext_k - k);
    const unsigned threads = 128;
    unsigned blocks = std::min<unsigned>((A.size + threads - 1) / threads, (unsigned)d.sms * 16);
    begin_call(d);
    H2B_LAUNCH(d, "quotient_shuffle", QK.shuffle, blocks, threads, 0, A);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

// ---- measurement ---------------------------------------------------------------------------------
double h2b_last_call_ms(const h2b_ctx* ctx) { return ctx ? ctx->last_ms : 0.0; }

int h2b_profile_enable(h2b_ctx* ctx, int on) {
    if (!ctx) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    for (Dev& d : ctx->devs) d.prof = on != 0;
    return H2B_OK;
}

int h2b_profile_read(h2b_ctx* ctx, char* json_out, size_t cap) {
    if (!ctx || !json_out || cap < 3) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    std::map<std::string, std::pair<double, uint64_t>> agg;
    for (Dev& d : ctx->devs) {
        cudaSetDevice(d.id);
        cudaStreamSynchronize(d.stream);
        for (auto& r : d.recs) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
                auto& a = agg[r.name];
                a.first += ms; a.second += 1;
            }
            d.ev_pool.push_back(r.a); d.ev_pool.push_back(r.b);
        }
        d.recs.clear();
    }
    std::string s = "{";
    bool first = true;
    for (auto& kv : agg) {
        char buf[256];
        std::snprintf(buf, sizeof buf, "%s\"%s\": {\"ms\": %.6f, \"launches\": %llu}", first ? "" : ", ", kv.first.c_str(),
                      kv.second.first, (unsigned long long)kv.second.second);
        s += buf;
        first = false;
    }
    s += "}";
    if (s.size() + 1 > cap) return ctx->fail(H2B_E_INVALID, "profile_read: buffer too small");
    std::memcpy(json_out, s.c_str(), s.size() + 1);
    return H2B_OK;
}

uint64_t h2b_launch_count(const h2b_ctx* ctx) {
    if (!ctx) return 0;
    uint64_t t = 0;
    for (const Dev& d : ctx->devs) t += d.launches;
    return t;
}

int h2b_synth_bases(h2b_ctx* ctx, void* d_out, size_t n, uint64_t t0) {
    if (!ctx) return H2B_E_INVALID;
    if (!d_out && n) return ctx->fail(H2B_E_INVALID, "synth_bases: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    if (n) H2B_LAUNCH(d, "synth_bases", MK.synth_bases, (unsigned)((n + 127) / 128), 128, 0, (Affine*)d_out, (uint64_t)n, t0);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

// ---- set-up helpers (ParamsKZG::setup fixed-base multiplications, Fr::random wide reduction) ----------------
const SetupKernelTable& SK = setup_kernel_table();

int h2b_g1_fixed_base_mul_device(h2b_ctx* ctx, const void* d_scalars, size_t n, void* d_out) {
    if (!ctx) return H2B_E_INVALID;
    if (n && (!d_scalars || !d_out)) return ctx->fail(H2B_E_INVALID, "g1_fixed_base_mul: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    if (n) H2B_LAUNCH(d, "g1_fixed_base_mul", SK.fixed_base_mul, (unsigned)((n + 127) / 128), 128, 0, (const Fr*)d_scalars, (uint64_t)n, (Affine*)d_out);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

int h2b_fr_from_u512_device(h2b_ctx* ctx, const void* d_limbs, size_t n, void* d_out) {
    if (!ctx) return H2B_E_INVALID;
    if (n && (!d_limbs || !d_out)) return ctx->fail(H2B_E_INVALID, "fr_from_u512: null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    begin_call(d);
    if (n) H2B_LAUNCH(d, "fr_from_u512", SK.fr_from_u512, (unsigned)((n + 255) / 256), 256, 0, (const uint32_t*)d_limbs, (uint64_t)n, (Fr*)d_out);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

int h2b_fr_random_chacha20_device(h2b_ctx* ctx, const void* seeds32, const uint64_t* offsets, size_t n_streams, void* d_out) {
    if (!ctx) return H2B_E_INVALID;
    if (!n_streams) return H2B_OK;
    if (!seeds32 || !offsets || !d_out) return ctx->fail(H2B_E_INVALID, "fr_random_chacha20: null argument");
    for (size_t i = 0; i < n_streams; i++)
        if (offsets[i] > offsets[i + 1]) return ctx->fail(H2B_E_INVALID, "fr_random_chacha20: offsets must not decrease");
    if (offsets[0] != 0) return ctx->fail(H2B_E_INVALID, "fr_random_chacha20: offsets[0] must be 0");
    const uint64_t n = offsets[n_streams];
    if (!n) return H2B_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    Stage st;
    size_t o_seeds = st.put(seeds32, n_streams * 32), o_off = st.put(offsets, (n_streams + 1) * sizeof(uint64_t));
    H2B_CK(d.q_stage.reserve(st.host.size() + 64));
    begin_call(d);
    H2B_CK(d.ring.h2d(d.q_stage.p, st.host.data(), st.host.size(), d.stream));
    uint8_t* base = d.q_stage.as<uint8_t>();
    H2B_LAUNCH(d, "fr_random_chacha20", SK.fr_random_chacha20, (unsigned)((n + 255) / 256), 256, 0, (const uint32_t*)(base + o_seeds),
               (const uint64_t*)(base + o_off), (uint32_t)n_streams, n, (Fr*)d_out);
    H2B_CK(cudaGetLastError());
    return end_call(ctx, d);
}

// ---- lookup argument: permute_expression_pair (lookup/prover.rs:398-494) -----------------------------------
int h2b_lookup_permute_device(h2b_ctx* ctx, const void* d_input, const void* d_table, size_t usable, size_t n, void* d_out_input,
                              void* d_out_table) {
    if (!ctx) return H2B_E_INVALID;
    if (n && (!d_input || !d_table || !d_out_input || !d_out_table)) return ctx->fail(H2B_E_INVALID, "lookup_permute: null argument");
    if (usable > n || usable >= ((size_t)1 << 30)) return ctx->fail(H2B_E_INVALID, "lookup_permute: usable rows out of range");
    if (!n) return H2B_OK;
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    const size_t need = lookup_permute_scratch_bytes((uint32_t)usable);
    H2B_CK(d.q_scratch.reserve(need));
    H2B_CK(d.reserve_pinned(256));
    begin_call(d);
    uint32_t* d_status = nullptr;
    int launches = 0;
    d.prof_begin("lookup_permute");
    H2B_CK(lookup_permute_run(d.launch_stream, (const Fr*)d_input, (const Fr*)d_table, (uint32_t)usable, (uint64_t)n, (Fr*)d_out_input,
                              (Fr*)d_out_table, d.q_scratch.p, d.q_scratch.cap, &d_status, &launches));
    d.prof_end();
    d.launches += launches;
    H2B_CK(cudaMemcpyAsync(d.pinned, d_status, 4, cudaMemcpyDeviceToHost, d.launch_stream));
    int rc = end_call(ctx, d);
    if (rc) return rc;
    const uint32_t status = *reinterpret_cast<const uint32_t*>(d.pinned);
    if (status & 1u) return ctx->fail(H2B_E_CONSTRAINT, "lookup_permute: an input value is missing from the table (ConstraintSystemFailure)");
    if (status & 2u) return ctx->fail(H2B_E_CONSTRAINT, "lookup_permute: table multiset does not cover the repeated rows");
    return H2B_OK;
}

int h2b_imad_peak(h2b_ctx* ctx, double* wide_mads_per_s, double* field_muls_per_s) {
    if (!ctx || !wide_mads_per_s || !field_muls_per_s) return H2B_E_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    Dev& d = ctx->devs[0];
    H2B_CK(cudaSetDevice(d.id));
    H2B_CK(d.m_counters.reserve(256));
    const unsigned blocks = (unsigned)d.sms * 8, threads = 256;
    const uint32_t it1 = 4096, it2 = 512;
    cudaEvent_t a = d.get_event(), b = d.get_event();
    float ms = 0.f;
    for (int rep = 0; rep < 3; rep++) {   // warm-up + 2 timed; keep the last
        H2B_CK(cudaEventRecord(a, d.stream));
        MK.imad_wide_peak<<<blocks, threads, 0, d.stream>>>(12345u + rep, it1, (unsigned long long*)d.m_counters.p);
        H2B_CK(cudaEventRecord(b, d.stream));
        H2B_CK(cudaStreamSynchronize(d.stream));
        H2B_CK(cudaEventElapsedTime(&ms, a, b));
    }
    *wide_mads_per_s = (double)blocks * threads * it1 * 8.0 / (ms * 1e-3);   // 8 IMAD.WIDE per iteration
    for (int rep = 0; rep < 3; rep++) {
        H2B_CK(cudaEventRecord(a, d.stream));
        MK.field_mul_peak<<<blocks, threads, 0, d.stream>>>(it2, (Fq*)d.m_counters.p);
        H2B_CK(cudaEventRecord(b, d.stream));
        H2B_CK(cudaStreamSynchronize(d.stream));
        H2B_CK(cudaEventElapsedTime(&ms, a, b));
    }
    *field_muls_per_s = (double)blocks * threads * it2 * 2.0 / (ms * 1e-3);
    d.launches += 6;
    d.ev_pool.push_back(a); d.ev_pool.push_back(b);
    H2B_CK(cudaGetLastError());
    return H2B_OK;
}

}  // extern "C"
