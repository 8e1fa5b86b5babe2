// ntt_plan.hpp -- host-side planning for the multi-pass NTT (ntt.cuh): pass split, twiddle-table layout,
// and construction of the per-pass descriptors.  Shared by the library (tables generated on the device)
// and the host test shim (tables generated with the same host field code).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "ntt.cuh"

namespace h2b {

struct NttPowerSegment {      // table[offset + i] = first * root^i, i < count
    size_t offset;
    Fr root, first;
    uint32_t count;
};

struct NttPlan {
    uint32_t log_n = 0;
    int n_pass = 0;
    uint32_t log_s[4] = {0, 0, 0, 0};
    uint32_t tw_h = 0;
    bool has_scale = false;
    size_t off_tw_s[4] = {0, 0, 0, 0};
    size_t off_lo = 0, off_hi_first = 0, off_hi_rest = 0;
    size_t table_elems = 0;          // power segments
    size_t off_full[4] = {0, 0, 0, 0};   // full boundary tables (0 = none), appended after the power segments
    size_t full_count[4] = {0, 0, 0, 0};
    size_t total_elems = 0;
    std::vector<NttPowerSegment> segments;
};

// Pass split: every pass <= 2^9; column passes need >= 8 columns to the right (inner >= 8).
// `split_override` = "9,9,6" overrides (for tuning; the context's ntt_split knob) when it sums to log_n.
inline void ntt_choose_split(uint32_t log_n, int& n_pass, uint32_t log_s[4], const char* split_override = nullptr) {
    if (const char* env = (split_override && *split_override) ? split_override : nullptr) {
        uint32_t v[4]; int m = 0; uint32_t sum = 0; const char* p = env;
        while (*p && m < 4) { v[m] = (uint32_t)std::strtoul(p, (char**)&p, 10); sum += v[m]; m++; if (*p == ',') p++; }
        bool ok = m >= 1 && sum == log_n;
        for (int i = 0; i < m && ok; i++) ok = v[i] >= 1 && v[i] <= (uint32_t)NTT_MAX_LOG_S;
        if (ok && m > 1) { uint32_t inner = log_n; for (int i = 0; i + 1 < m; i++) { inner -= v[i]; if (inner < 3) ok = false; } if (v[0] < 3) ok = false; }
        if (ok) { n_pass = m; for (int i = 0; i < m; i++) log_s[i] = v[i]; return; }
    }
    if (log_n <= 9) { n_pass = 1; log_s[0] = log_n; return; }
    n_pass = log_n <= 18 ? 2 : (log_n <= 27 ? 3 : 4);
    uint32_t rem = log_n;
    for (int i = 0; i < n_pass; i++) {
        uint32_t s = (rem + (uint32_t)(n_pass - i) - 1) / (uint32_t)(n_pass - i);
        log_s[i] = s;
        rem -= s;
    }
}

// omega: primitive 2^log_n-th root (Montgomery); scale: optional multiplier folded into the transform.
inline NttPlan ntt_make_plan(uint32_t log_n, const Fr& omega, const Fr* scale, const char* split_override = nullptr,
                             uint32_t full_table_max_log = 26) {
    NttPlan pl;
    pl.log_n = log_n;
    ntt_choose_split(log_n, pl.n_pass, pl.log_s, split_override);
    pl.has_scale = scale != nullptr;
    size_t off = 0;
    for (int j = 0; j < pl.n_pass; j++) {
        // sub-NTT root of pass j: omega^(N / S_j)
        Fr root = omega;
        for (uint32_t i = 0; i < log_n - pl.log_s[j]; i++) root = fe_sqr(root);
        pl.off_tw_s[j] = off;
        pl.segments.push_back({off, root, Fr::one(), 1u << pl.log_s[j]});
        off += (size_t)1 << pl.log_s[j];
    }
    if (pl.n_pass > 1) {
        pl.tw_h = (log_n + 1) / 2;
        uint32_t hi_bits = log_n - pl.tw_h;
        Fr root_hi = omega;
        for (uint32_t i = 0; i < pl.tw_h; i++) root_hi = fe_sqr(root_hi);
        pl.off_lo = off;
        pl.segments.push_back({off, omega, Fr::one(), 1u << pl.tw_h});
        off += (size_t)1 << pl.tw_h;
        pl.off_hi_rest = off;
        pl.segments.push_back({off, root_hi, Fr::one(), 1u << hi_bits});
        off += (size_t)1 << hi_bits;
        if (scale) {
            pl.off_hi_first = off;
            pl.segments.push_back({off, root_hi, *scale, 1u << hi_bits});
            off += (size_t)1 << hi_bits;
        } else {
            pl.off_hi_first = pl.off_hi_rest;
        }
    }
    pl.table_elems = off;
    // Full boundary tables (one multiplication saved per element per column pass) while they stay affordable:
    // pass j needs N / outer_j entries.  full_table_max_log (default 26) caps the largest one.
    const uint32_t cap = full_table_max_log;
    uint32_t done = 0;
    for (int j = 0; j + 1 < pl.n_pass; j++) {
        uint32_t lg = log_n - done;
        if (lg <= cap) { pl.off_full[j] = off; pl.full_count[j] = (size_t)1 << lg; off += (size_t)1 << lg; }
        done += pl.log_s[j];
    }
    pl.total_elems = off;
    return pl;
}

struct NttFused {            // the EvaluationDomain wrappers' extra work, fused into first load / last store
    uint64_t in_len = ~0ull;          // inputs beyond in_len are zero (coeff_to_extended zero extension)
    uint64_t out_len = ~0ull;         // outputs beyond out_len are dropped (extended_to_coeff truncation)
    const Fr* in_pat = nullptr;       // [2]: multipliers for index%3 == 1, 2 on the way in
    const Fr* out_pat = nullptr;      // [2]: same on the way out
    const Fr* in_tab = nullptr;       // table in the memory space of the kernel: input[i] *= in_tab[i & in_tab_mask]
    uint32_t in_tab_mask = 0;
};

// Descriptors for all passes.  `tables` is the base of the generated table buffer (device or host).
inline void ntt_build_passes(const NttPlan& pl, const Fr* tables, const Fr* scale, const NttFused& fz, NttPass out[4]) {
    uint32_t done = 0;
    for (int j = 0; j < pl.n_pass; j++) {
        NttPass& P = out[j];
        std::memset(&P, 0, sizeof(P));
        P.log_s = pl.log_s[j];
        P.tw_s = tables + pl.off_tw_s[j];
        P.tw_h = pl.tw_h;
        P.in_len = ~0ull;
        P.out_len = ~0ull;
        bool last = j == pl.n_pass - 1;
        if (!last) {
            P.log_outer = done;
            P.log_inner = pl.log_n - done - pl.log_s[j];
            P.tw_lo = tables + pl.off_lo;
            P.tw_hi = tables + (j == 0 ? pl.off_hi_first : pl.off_hi_rest);
            P.tw_full = pl.full_count[j] ? tables + pl.off_full[j] : nullptr;
        } else {
            P.log_outer = done;
            P.n_prev = (uint32_t)j;
            for (int i = 0; i < j; i++) P.prev_log[i] = pl.log_s[i];
            P.out_len = fz.out_len;
            if (fz.out_pat) { P.flags |= NTT_OUT_PATTERN; P.out_pat[0] = fz.out_pat[0]; P.out_pat[1] = fz.out_pat[1]; }
            if (pl.n_pass == 1 && scale) { P.flags |= NTT_OUT_SCALE; P.out_scale = *scale; }
        }
        if (j == 0) {
            P.in_len = fz.in_len;
            if (fz.in_pat) { P.flags |= NTT_IN_PATTERN; P.in_pat[0] = fz.in_pat[0]; P.in_pat[1] = fz.in_pat[1]; }
            if (fz.in_tab) { P.flags |= NTT_IN_TABLE; P.in_tab = fz.in_tab; P.in_tab_mask = fz.in_tab_mask; }
        }
        done += pl.log_s[j];
    }
}

// number of tiles (CTAs along x) of pass j
inline uint64_t ntt_pass_tiles(const NttPlan& pl, int j) {
    uint64_t elems = 1ull << pl.log_n;
    uint64_t per_tile = (uint64_t)NTT_LANES << pl.log_s[j];
    return elems <= per_tile ? 1 : elems / per_tile;
}

}  // namespace h2b
