// host_check.cpp -- unit-test shim: compiles field.cuh / curve.cuh for the HOST (carry-flag PTX ops
// emulated in C) so the exact limb-level algorithms the kernels run are checked on the CPU against the
// oracle (tests/test_field_host.py).  Not part of the product path: libh2b200.so never links this.
#include <cstring>
#include "curve.cuh"
#include "field_wide.cuh"
#include "poly.cuh"
#include "msm_ba.cuh"
#include "streamcopy.hpp"
#include <vector>
using namespace h2b;

template <class F> static F ld(const void* p) { F f; std::memcpy(&f, p, 32); return f; }
template <class F> static void st(void* p, const F& f) { std::memcpy(p, &f, 32); }

extern "C" {
void hc_fr_mul(const void* a, const void* b, void* o) { st(o, fe_mul(ld<Fr>(a), ld<Fr>(b))); }
void hc_fq_mul(const void* a, const void* b, void* o) { st(o, fe_mul(ld<Fq>(a), ld<Fq>(b))); }
void hc_fr_add(const void* a, const void* b, void* o) { st(o, fe_add(ld<Fr>(a), ld<Fr>(b))); }
void hc_fr_sub(const void* a, const void* b, void* o) { st(o, fe_sub(ld<Fr>(a), ld<Fr>(b))); }
void hc_fq_add(const void* a, const void* b, void* o) { st(o, fe_add(ld<Fq>(a), ld<Fq>(b))); }
void hc_fq_sub(const void* a, const void* b, void* o) { st(o, fe_sub(ld<Fq>(a), ld<Fq>(b))); }
void hc_fr_neg(const void* a, void* o) { st(o, fe_neg(ld<Fr>(a))); }
void hc_fr_inv(const void* a, void* o) { st(o, fe_inv(ld<Fr>(a))); }
void hc_fr_inv_window(const void* a, void* o) { st(o, fe_inv_window(ld<Fr>(a))); }
void hc_fq_inv_window(const void* a, void* o) { st(o, fe_inv_window(ld<Fq>(a))); }
void hc_fr_from_mont(const void* a, void* o) { st(o, fe_from_mont(ld<Fr>(a))); }
void hc_fr_to_mont(const void* a, void* o) { st(o, fe_to_mont(ld<Fr>(a))); }

// acc = sum_i (+/-) pts[i] through xyzz_madd, returned affine (64 B)
void hc_madd_chain(const void* pts, const uint8_t* neg, size_t n, void* out64) {
    Xyzz acc = Xyzz::identity();
    for (size_t i = 0; i < n; i++) {
        Affine p; std::memcpy(&p, (const uint8_t*)pts + 64 * i, 64);
        if (p.is_identity()) continue;
        xyzz_madd(acc, p.x, p.y, neg[i] != 0);
    }
    Affine r = xyzz_to_affine(acc);
    std::memcpy(out64, &r, 64);
}
// tree-sum of affine points through xyzz_add (exercises add / double / cancellation), returned affine
void hc_add_tree(const void* pts, size_t n, void* out64) {
    Xyzz* v = new Xyzz[n ? n : 1];
    for (size_t i = 0; i < n; i++) { Affine p; std::memcpy(&p, (const uint8_t*)pts + 64 * i, 64); v[i] = Xyzz::from_affine(p); }
    // give the operands non-trivial ZZ by doubling+subtracting: P = 2P + (-P)
    for (size_t i = 0; i < n; i++) v[i] = xyzz_add(xyzz_double(v[i]), xyzz_neg(v[i]));
    for (size_t s = 1; s < n; s <<= 1)
        for (size_t i = 0; i + s < n; i += 2 * s) v[i] = xyzz_add(v[i], v[i + s]);
    Affine r = n ? xyzz_to_affine(v[0]) : xyzz_to_affine(Xyzz::identity());
    std::memcpy(out64, &r, 64);
    delete[] v;
}
// the staging ring's copy (streamcopy.hpp): any alignment of either side, any length
void hc_stream_copy(void* dst, const void* src, size_t len) { stream_copy(dst, src, len); }
// batched-affine bucket accumulation (msm_ba.cuh): one bucket through the same per-thread phases the kernel runs -- `rounds`
// (1 or 2) rounds of paired affine additions, the block-wide inversion replaced by the thread's own fe_inv_bgcd -- returned affine.
void hc_fq_inv_bgcd(const void* a, void* o) { st(o, fe_inv_bgcd(ld<Fq>(a))); }
void hc_fr_inv_bgcd(const void* a, void* o) { st(o, fe_inv_bgcd(ld<Fr>(a))); }
void hc_msm_ba_bucket(const void* pts, const uint8_t* neg, uint32_t n, int rounds, void* out64) {
    std::vector<uint32_t> entries(n ? n : 1);
    for (uint32_t i = 0; i < n; i++) entries[i] = i | (neg[i] ? 0x80000000u : 0u);
    std::vector<Fq> prefix(MSM_BA_KCAP), rx(MSM_BA_KCAP), ry(MSM_BA_KCAP);
    MsmBaSource s0;
    s0.bases = (const Affine*)pts; s0.entries = entries.data(); s0.rx = nullptr; s0.ry = nullptr; s0.stride = 1; s0.from_entries = true;
    uint32_t K0 = n >> 1;
    if (K0 > MSM_BA_KCAP) K0 = MSM_BA_KCAP;
    Fq inv = fe_inv_bgcd(msm_ba_phase_a(s0, K0, prefix.data(), 1));
    Xyzz acc = Xyzz::identity();
    if (rounds == 2) {
        uint32_t K1 = msm_ba_phase_b<true>(s0, K0, n, prefix.data(), 1, inv, acc, rx.data(), ry.data());
        MsmBaSource s1;
        s1.bases = nullptr; s1.entries = nullptr; s1.rx = rx.data(); s1.ry = ry.data(); s1.stride = 1; s1.from_entries = false;
        inv = fe_inv_bgcd(msm_ba_phase_a(s1, K1 >> 1, prefix.data(), 1));
        msm_ba_phase_b<false>(s1, K1 >> 1, K1, prefix.data(), 1, inv, acc, nullptr, nullptr);
    } else {
        msm_ba_phase_b<false>(s0, K0, n, prefix.data(), 1, inv, acc, nullptr, nullptr);
    }
    Affine r = xyzz_to_affine(acc);
    std::memcpy(out64, &r, 64);
}
}

// ------------------------------------------------------------------------------------------------
// NTT: run the very same per-thread level code (ntt.cuh) tile by tile, thread by thread.
// ------------------------------------------------------------------------------------------------
#include <vector>
#include "ntt_plan.hpp"

template <int LOG_S, bool ROW>
static void run_pass_host(const NttPass& P, uint64_t tiles, const Fr* in, Fr* out) {
    constexpr int S = 1 << LOG_S;
    constexpr int NLEV = detail::lev_count(LOG_S);
    std::vector<U4> slo((size_t)S * NTT_PITCH), shi((size_t)S * NTT_PITCH);
    const uint32_t threads = S < 32 ? 32 : S;
    for (uint64_t tile = 0; tile < tiles; tile++) {
        for (uint32_t t = 0; t < threads; t++) ntt_level<LOG_S, 0, ROW>(P, t, tile, in, out, slo.data(), shi.data(), P.tw_s, 0);
        if constexpr (NLEV > 1)
            for (uint32_t t = 0; t < threads; t++) ntt_level<LOG_S, 1, ROW>(P, t, tile, in, out, slo.data(), shi.data(), P.tw_s, 0);
        if constexpr (NLEV > 2)
            for (uint32_t t = 0; t < threads; t++) ntt_level<LOG_S, 2, ROW>(P, t, tile, in, out, slo.data(), shi.data(), P.tw_s, 0);
    }
}
template <bool ROW>
static void run_pass_dispatch(const NttPass& P, uint64_t tiles, const Fr* in, Fr* out) {
    switch (P.log_s) {
        case 1: run_pass_host<1, ROW>(P, tiles, in, out); break;
        case 2: run_pass_host<2, ROW>(P, tiles, in, out); break;
        case 3: run_pass_host<3, ROW>(P, tiles, in, out); break;
        case 4: run_pass_host<4, ROW>(P, tiles, in, out); break;
        case 5: run_pass_host<5, ROW>(P, tiles, in, out); break;
        case 6: run_pass_host<6, ROW>(P, tiles, in, out); break;
        case 7: run_pass_host<7, ROW>(P, tiles, in, out); break;
        case 8: run_pass_host<8, ROW>(P, tiles, in, out); break;
        case 9: run_pass_host<9, ROW>(P, tiles, in, out); break;
    }
}

extern "C" int hc_ntt(const void* in_, uint64_t in_len, void* out_, uint64_t out_len, uint32_t log_n, const void* omega_,
                      const void* scale_or_null, const void* in_pat_or_null, const void* out_pat_or_null) {
    if (log_n == 0) { std::memcpy(out_, in_, 32); return 0; }
    Fr omega = ld<Fr>(omega_);
    Fr scale; if (scale_or_null) scale = ld<Fr>(scale_or_null);
    // test shim only: the split override comes from the environment here (the library takes it from its context's knobs)
    NttPlan pl = ntt_make_plan(log_n, omega, scale_or_null ? &scale : nullptr, std::getenv("H2B_NTT_SPLIT"));
    std::vector<Fr> tables(pl.total_elems);
    for (auto& seg : pl.segments) {
        Fr acc = seg.first;
        for (uint32_t i = 0; i < seg.count; i++) { tables[seg.offset + i] = acc; acc = fe_mul(acc, seg.root); }
    }
    {   // full boundary tables, as ntt_gen_boundary_kernel fills them
        uint32_t done = 0;
        for (int j = 0; j + 1 < pl.n_pass; j++) {
            if (pl.full_count[j]) {
                uint32_t log_inner = log_n - done - pl.log_s[j];
                const Fr* hi = tables.data() + (j == 0 ? pl.off_hi_first : pl.off_hi_rest);
                for (uint64_t idx = 0; idx < pl.full_count[j]; idx++) {
                    uint64_t k = idx >> log_inner, i = idx & ((1ull << log_inner) - 1), e = (i * k) << done;
                    tables[pl.off_full[j] + idx] = fe_mul(tables[pl.off_lo + (e & ((1ull << pl.tw_h) - 1))], hi[e >> pl.tw_h]);
                }
            }
            done += pl.log_s[j];
        }
    }
    NttFused fz; fz.in_len = in_len; fz.out_len = out_len;
    Fr ip[2], op[2];
    if (in_pat_or_null) { std::memcpy(ip, in_pat_or_null, 64); fz.in_pat = ip; }
    if (out_pat_or_null) { std::memcpy(op, out_pat_or_null, 64); fz.out_pat = op; }
    NttPass passes[4];
    ntt_build_passes(pl, tables.data(), scale_or_null ? &scale : nullptr, fz, passes);
    const Fr* in = (const Fr*)in_;
    Fr* out = (Fr*)out_;
    std::vector<Fr> tmp(pl.n_pass > 1 ? (size_t)1 << log_n : 0);
    for (int j = 0; j < pl.n_pass; j++) {
        bool last = j == pl.n_pass - 1;
        const Fr* src = j == 0 ? in : tmp.data();
        Fr* dst = last ? out : tmp.data();
        uint64_t tiles = ntt_pass_tiles(pl, j);
        if (last) run_pass_dispatch<true>(passes[j], tiles, src, dst);
        else run_pass_dispatch<false>(passes[j], tiles, src, dst);
    }
    return pl.n_pass;
}

// ------------------------------------------------------------------------------------------------
// MSM: the same digit recoding, bucket reduction chunking and window combine as the kernels, run serially.
// ------------------------------------------------------------------------------------------------
#include "msm.cuh"
static void hc_msm_windows_impl(const void* scalars_, const void* bases_, uint32_t n, uint32_t c, uint32_t reduce_threads, std::vector<Xyzz>& windows);
extern "C" void hc_msm(const void* scalars_, const void* bases_, uint32_t n, uint32_t c, uint32_t reduce_threads, void* out96) {
    std::vector<Xyzz> windows;
    hc_msm_windows_impl(scalars_, bases_, n, c, reduce_threads, windows);
    msm_finish(windows.data(), msm_windows(c), c, out96);
}
// per-window sums of one point range (what h2b_msm_windows returns on the GPU): windows x 128 bytes
extern "C" void hc_msm_windows(const void* scalars_, const void* bases_, uint32_t n, uint32_t c, void* out_windows) {
    std::vector<Xyzz> windows;
    hc_msm_windows_impl(scalars_, bases_, n, c, 0, windows);
    std::memcpy(out_windows, windows.data(), windows.size() * sizeof(Xyzz));
}
static void hc_msm_windows_impl(const void* scalars_, const void* bases_, uint32_t n, uint32_t c, uint32_t reduce_threads, std::vector<Xyzz>& windows) {
    const Fr* scalars = (const Fr*)scalars_;
    const Affine* bases = (const Affine*)bases_;
    uint32_t W = msm_windows(c), nb = 1u << (c - 1);
    std::vector<std::vector<uint32_t>> seg((size_t)W * nb);
    for (uint32_t i = 0; i < n; i++) {
        Fr canon = fe_from_mont(fe_load(scalars + i));
        msm_signed_digits(canon, c, W, [&](uint32_t j, int32_t d) {
            if (d != 0) {
                uint32_t mag = d < 0 ? (uint32_t)(-d) : (uint32_t)d;
                seg[(size_t)j * nb + mag - 1].push_back(i | (d < 0 ? 0x80000000u : 0u));
            }
        });
    }
    std::vector<Xyzz> buckets((size_t)W * nb);
    for (size_t k = 0; k < buckets.size(); k++) {
        Xyzz acc = Xyzz::identity();
        for (uint32_t ent : seg[k]) {
            Affine p = affine_load(bases + (ent & 0x7fffffffu));
            if (!p.is_identity()) xyzz_madd(acc, p.x, p.y, (ent >> 31) != 0);
        }
        buckets[k] = acc;
    }
    windows.assign(W, Xyzz::identity());
    for (uint32_t j = 0; j < W; j++) {
        const Xyzz* bk = buckets.data() + (size_t)j * nb;
        if (!reduce_threads || nb < reduce_threads) {
            Xyzz run;
            windows[j] = msm_chunk_sums(bk, nb, run);
            continue;
        }
        // the kernels' hierarchy: blocks of `reduce_threads` threads x L buckets; P_b = A + L*T, R_b; then
        // window = sum_b P_b + (threads * L) * sum_b b * R_b
        uint32_t slices = nb / reduce_threads, log_l = 0;
        while (slices > 4) { slices >>= 1; log_l++; }          // small cap so that several levels are exercised
        uint32_t L = 1u << log_l;
        Xyzz sumP = Xyzz::identity(), sumbR = Xyzz::identity();
        for (uint32_t b = 0; b < slices; b++) {
            Xyzz A = Xyzz::identity(), T = Xyzz::identity(), Rn = Xyzz::identity();
            for (uint32_t t = 0; t < reduce_threads; t++) {
                Xyzz run;
                Xyzz acc = msm_chunk_sums(bk + (((size_t)b * reduce_threads + t) << log_l), L, run);
                A = xyzz_add(A, acc);
                for (uint32_t r = 0; r < t; r++) T = xyzz_add(T, run);      // t * run_t
                Rn = xyzz_add(Rn, run);
            }
            Xyzz P = xyzz_add(A, xyzz_mul_pow2(T, log_l));
            sumP = xyzz_add(sumP, P);
            for (uint32_t r = 0; r < b; r++) sumbR = xyzz_add(sumbR, Rn);
        }
        uint32_t log_bb = log_l; { uint32_t rt = reduce_threads; while (rt > 1) { rt >>= 1; log_bb++; } }
        windows[j] = xyzz_add(sumP, xyzz_mul_pow2(sumbR, log_bb));
    }
}

// split multiplier (field_wide.cuh)
extern "C" {
void hc_fq_mul_w(const void* a, const void* b, void* o) { st(o, fe_mul_w(ld<Fq>(a), ld<Fq>(b))); }
void hc_fr_mul_w(const void* a, const void* b, void* o) { st(o, fe_mul_w(ld<Fr>(a), ld<Fr>(b))); }
void hc_fq_sqr_w(const void* a, void* o) { st(o, fe_sqr_w(ld<Fq>(a))); }
void hc_fr_sqr_w(const void* a, void* o) { st(o, fe_sqr_w(ld<Fr>(a))); }
void hc_fq_mul_sub_mul(const void* a, const void* b, const void* c, const void* d, void* o) {
    st(o, fe_mul_sub_mul(ld<Fq>(a), ld<Fq>(b), ld<Fq>(c), ld<Fq>(d)));
}
void hc_mul_wide(const void* a, const void* b, void* o64) { Fq x = ld<Fq>(a), y = ld<Fq>(b); Wide w = mul_wide(x.l, y.l); std::memcpy(o64, &w, 64); }
void hc_sqr_wide(const void* a, void* o64) { Fq x = ld<Fq>(a); Wide w = sqr_wide(x.l); std::memcpy(o64, &w, 64); }
}
