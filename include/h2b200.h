/*
 * h2b200.h -- C ABI of libh2b200.so: B200-native (sm_100a) MSM and NTT for the halo2 KZG prover on BN254.
 *
 * This is the drop-in boundary for the two data-parallel kernels of the reference
 * (yetanotherco/yet-another-halo2-fork, paths below are under /root/reference):
 *
 *   MSM   halo2_middleware/src/zal.rs:57-103     trait MsmAccel<C> { msm, get_*_descriptor, msm_with_cached_* }
 *         halo2_middleware/src/zal.rs:136-138    H2cEngine::msm -> halo2curves::msm::best_multiexp
 *         halo2_backend/src/poly/kzg/commitment.rs:305-317,354-366   ParamsKZG::commit_lagrange / commit
 *   NTT   halo2_backend/src/poly/domain.rs:216   EvaluationDomain::lagrange_to_coeff
 *         halo2_backend/src/poly/domain.rs:230   EvaluationDomain::coeff_to_extended
 *         halo2_backend/src/poly/domain.rs:271   EvaluationDomain::extended_to_coeff
 *         halo2_backend/src/poly/domain.rs:297   EvaluationDomain::divide_by_vanishing_poly (fusable)
 *         halo2_backend/src/poly/domain.rs:343   ifft = best_fft + scale  (halo2curves::fft::best_fft)
 *
 * Data layouts are the reference's in-memory layouts (halo2curves 0.6 bn256), so Rust slices are passed
 * as-is with no conversion:
 *   Fr, Fq      32 B  = [u64; 4] little-endian limbs, Montgomery form (R = 2^256), fully reduced
 *   G1Affine    64 B  = (x, y), identity encoded as (0, 0)
 *   G1          96 B  = Jacobian (x, y, z), identity z = 0.  Results are returned NORMALISED (z = 1, or
 *                       (0, 1, 0) for the identity), i.e. the canonical representative Curve::to_affine /
 *                       batch_normalize would produce, so results are bit-identical whatever the bucket order.
 *
 * Conventions: every function returns 0 on success and a negative H2B_E_* code on failure (the reason is
 * kept per context: h2b_last_error).  Host pointers are borrowed for the duration of the call only.  All
 * calls are synchronous on return and serialised per context (one context may be shared by threads).
 * Nothing throws across this boundary.  There is no CPU fallback: without a usable sm_100 device
 * h2b_init fails.
 *
 * INTEGRATION.md shows the Rust binding (`impl MsmAccel<G1Affine> for B200Engine`, EvaluationDomain patch).
 */
#ifndef H2B200_H
#define H2B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct h2b_ctx h2b_ctx;

enum {
    H2B_OK = 0,
    H2B_E_INVALID = -1,   /* bad argument (null pointer, length mismatch, log_n out of range, unknown handle) */
    H2B_E_CUDA = -2,      /* CUDA runtime / launch failure */
    H2B_E_NOMEM = -3,     /* device or host allocation failed */
    H2B_E_NODEVICE = -4,  /* no usable GPU */
    H2B_E_CONSTRAINT = -5 /* the inputs violate the relation being proven (plonk::Error::ConstraintSystemFailure) */
};

/* ---- context ------------------------------------------------------------------------------------------
 * Replaces the engine object of zal.rs:182-243 (PlonkEngine { msm_backend }): one context owns the devices,
 * streams, resident bases and twiddle tables.  device_ids == NULL, n_devices == 0 selects device 0.
 * With n_devices > 1 the MSM is sharded by contiguous point range and batched NTT columns are dealt
 * round-robin (a single NTT stays on device_ids[0]).
 * h2b_destroy on a context that is unknown or already destroyed is ignored, and h2b_bases_release on such a context returns
 * H2B_E_INVALID: hosts with garbage collectors finalise descriptors and contexts in no particular order. */
int h2b_init(h2b_ctx** out, const int* device_ids, int n_devices);
void h2b_destroy(h2b_ctx* ctx);
const char* h2b_last_error(const h2b_ctx* ctx);
int h2b_device_count(void);
const char* h2b_version(void);
/* h2b_last_notice: the most recent non-fatal notice of the context ("" if none), e.g. that h2b_bases_register had to skip the
 *                  window precomputation for lack of device memory (MSMs on that handle are 2-3x slower).
 * h2b_set_option:  tuning knobs of one context.  Each knob is also read ONCE from the environment in h2b_init (variable =
 *                  "H2B_" + upper-case name); value NULL or "" restores the default.  Names: msm_c, msm_pre_c, msm_precompute,
 *                  msm_chunks, msm_no_ramp, msm_no_batch, msm_reduce_classic, msm_reduce_seg, msm_ba (rounds of batched affine additions in
 *                  the bucket accumulation: 0, 1, 2), msm_ba_regs, msm_ba_min_mean, msm_canon_once, msm_sort_blocks, acc_variant, msm_sort, msm_sort_tb,
 *                  sort_prio (read when the context creates its streams: environment only), ntt_split ("9,9,6"), ntt_full_table_max_log, peer_timeout_ms.  Unknown name -> H2B_E_INVALID. */
const char* h2b_last_notice(const h2b_ctx* ctx);
int h2b_set_option(h2b_ctx* ctx, const char* name, const char* value);

/* ---- MSM: MsmAccel<G1Affine> ---------------------------------------------------------------------------
 * h2b_bases_register     = MsmAccel::get_base_descriptor (zal.rs:82): copy n affine bases to the device(s)
 *                          once (ParamsKZG::g / g_lagrange stay resident across commits).  This is the descriptor's
 *                          "expensive preprocessing" (zal.rs:71-77): for n >= 2^16 the multiples 2^(c*j) * P_i of every
 *                          window j are computed and kept as well (13-15 x 64 B per point, ~0.04 s per 2^20 points),
 *                          so that an MSM needs one bucket set instead of one per window.  H2B_MSM_PRECOMPUTE=0
 *                          (environment) keeps only the 64 B per point.
 * h2b_bases_register_device: same, adopting a buffer that already lives on device 0 (not copied, not freed).
 * h2b_bases_release      = Drop of the descriptor (zal.rs:47-50).
 * h2b_msm                = MsmAccel::msm_with_cached_base (zal.rs:90): uses bases[0..n) of the handle; n may
 *                          be smaller than the registered count (commit on a shorter polynomial,
 *                          commitment.rs:313-316) but never larger -> H2B_E_INVALID (the reference asserts).
 * h2b_msm_raw            = MsmAccel::msm (zal.rs:58): uncached bases, copied for this call only.
 * h2b_msm_device         = same as h2b_msm with the scalars already resident on device 0 (descriptor path,
 *                          MsmAccel::msm_with_cached_inputs, zal.rs:96).
 * out96: G1 Jacobian, normalised. */
int h2b_bases_register(h2b_ctx* ctx, const void* bases, size_t n, uint64_t* handle_out);
int h2b_bases_register_device(h2b_ctx* ctx, const void* d_bases, size_t n, uint64_t* handle_out);
int h2b_bases_release(h2b_ctx* ctx, uint64_t handle);
int h2b_msm(h2b_ctx* ctx, uint64_t bases_handle, const void* scalars, size_t n, void* out96);
int h2b_msm_raw(h2b_ctx* ctx, const void* bases, const void* scalars, size_t n, void* out96);
int h2b_msm_device(h2b_ctx* ctx, uint64_t bases_handle, const void* d_scalars, size_t n, void* out96);
/* h2b_bases_info: what a registered handle holds: *window_bits_out = c of the precomputed window multiples (0 = none: the
 *                 multiples were disabled, too small to pay, or did not fit -- see h2b_last_notice), *device_bytes_out = device
 *                 memory the handle owns (13.9 GB for 2^24 points at c = 20).  Either pointer may be NULL. */
int h2b_bases_info(h2b_ctx* ctx, uint64_t bases_handle, uint32_t* window_bits_out, uint64_t* device_bytes_out);

/* Multi-process sharding (one process per GPU, point-range partition): each rank computes the per-window
 * bucket sums of its own range, the ranks exchange H2B_MSM_WINDOW_BYTES * windows bytes (a tiny all-gather),
 * and any rank folds them.  Window size c is a pure function of n_total so that all ranks agree.
 *   h2b_msm_windows : out_windows receives `*windows_out` points of H2B_MSM_WINDOW_BYTES each.
 *   h2b_msm_combine : host-only; parts = n_parts consecutive window arrays. */
#define H2B_MSM_WINDOW_BYTES 128
#define H2B_MSM_MAX_WINDOWS 64
int h2b_msm_windows(h2b_ctx* ctx, uint64_t bases_handle, const void* scalars, int scalars_on_device, size_t n,
                    size_t n_total, void* out_windows, uint32_t* c_out, uint32_t* windows_out);
int h2b_msm_combine(const void* parts, uint32_t n_parts, uint32_t c, uint32_t windows, void* out96);
/* The same sharding with the exchange INSIDE the library and on the device (SURVEY.md 8a `partials_gather` as the epilogue of
 * the window combine): every rank owns a small device mailbox; the last kernel of a rank's reduction stores its reduced terms
 * (<= 64 points) into every peer's mailbox over NVLink, waits for theirs and folds them, so each rank returns the WHOLE result
 * with no collective call and no host staging.
 *   h2b_peer_export : allocate this context's mailbox; handle64_out receives the 64-byte CUDA IPC handle the other processes
 *                     need, *local_ptr_out (optional) the device pointer for peers living in the SAME process.
 *   h2b_peer_attach : handles = world x 64 bytes in rank order (e.g. all-gathered by the launcher); local_ptrs (optional,
 *                     world entries) overrides the handle of a same-process peer with its pointer.  One device per context.
 *   h2b_msm_sharded : this rank's n scalars against its registered range; every rank must call it, with equal shard sizes
 *                     (the ranks must reduce with the same window geometry; a mismatch or a missing peer is reported as
 *                     H2B_E_CUDA after `peer_timeout_ms`).  out96 = the normalised sum over all ranks, on every rank.
 *   Ordering and recovery: h2b_peer_attach clears this rank's mailbox and restarts its call counter, so EVERY rank must have
 *   returned from h2b_peer_attach before ANY rank calls h2b_msm_sharded (a barrier in the launcher).  Every h2b_msm_sharded
 *   call counts, refused or failed ones included, so after an error that every rank saw (or a call every rank made) the group
 *   can simply call again.  A rank whose exchange fails marks the terms it had published, so a peer arriving later fails the
 *   same call ("gave up") rather than folding them.  If a rank skipped a call altogether the peers report it ("gave up" /
 *   "is ahead of this rank" / time-out); then all ranks attach again (same handles), barrier, and continue. */
#define H2B_PEER_HANDLE_BYTES 64
int h2b_peer_export(h2b_ctx* ctx, void* handle64_out, void** local_ptr_out);
int h2b_peer_attach(h2b_ctx* ctx, int rank, int world, const void* handles, void* const* local_ptrs_or_null);
int h2b_msm_sharded(h2b_ctx* ctx, uint64_t bases_handle, const void* scalars, int scalars_on_device, size_t n, size_t n_total,
                    void* out96);
/* h2b_msm_batch: n_cols MSMs of n scalars each against bases[0..n) of ONE handle -- the per-column commits of a prover phase
 * (plonk/prover.rs:436-440 `.iter().map(commit_lagrange)`, vanishing/prover.rs:136-140 the h pieces).  out96s receives
 * n_cols x 96 bytes.  Same results as n_cols calls of h2b_msm; the columns share one pipeline (binning of column k+1 overlaps
 * accumulation of column k, one reduction launch for all).  An API extension: the reference has no batched commit. */
int h2b_msm_batch(h2b_ctx* ctx, uint64_t bases_handle, const void* const* scalars_cols, size_t n_cols, size_t n, void* out96s);
int h2b_msm_batch_device(h2b_ctx* ctx, uint64_t bases_handle, const void* const* d_scalars_cols, size_t n_cols, size_t n,
                         void* out96s);

/* ---- NTT: EvaluationDomain ----------------------------------------------------------------------------
 * All roots / divisors / coset constants are the fields of the reference's EvaluationDomain (domain.rs:18-33)
 * passed in as 32-byte Montgomery Fr; nothing is re-derived on this side.
 *
 * h2b_ntt               best_fft(a, omega, log_n) in place; if scale32 != NULL every output is multiplied by
 *                       it (ifft, domain.rs:343-351: omega = omega_inv, scale = ifft_divisor)
 *                       -> lagrange_to_coeff(a) = h2b_ntt(a, k, omega_inv, ifft_divisor).
 * h2b_coeff_to_extended domain.rs:230-244: out[0..2^ext_k) = best_fft(zero_extend(a[i] * {1,g_coset,g_coset_inv}[i%3])).
 * h2b_extended_to_coeff domain.rs:271-293: ifft with ext_omega_inv / ext_ifft_divisor, then
 *                       a[i] *= {1,g_coset_inv,g_coset}[i%3], truncated to out_len (= n * quotient_poly_degree).
 *                       If t_evals32 != NULL (e = n_t entries, domain.rs:297-316) the input is first divided by
 *                       the vanishing polynomial: a[i] *= t_evals[i % n_t]  (divide_by_vanishing_poly fused).
 * *_device variants take device pointers (device 0) and leave the result on the device.
 * h2b_ntt_batch         n_cols independent transforms of the same shape (the per-column coset NTTs of
 *                       evaluate_h, plonk/evaluation.rs:344-361); columns are dealt across the context's devices. */
int h2b_ntt(h2b_ctx* ctx, void* a, uint32_t log_n, const void* omega32, const void* scale32_or_null);
int h2b_ntt_device(h2b_ctx* ctx, void* d_a, uint32_t log_n, const void* omega32, const void* scale32_or_null);
int h2b_coeff_to_extended(h2b_ctx* ctx, const void* in_n, void* out_ne, uint32_t k, uint32_t ext_k,
                          const void* ext_omega32, const void* g_coset32, const void* g_coset_inv32);
int h2b_coeff_to_extended_device(h2b_ctx* ctx, const void* d_in_n, void* d_out_ne, uint32_t k, uint32_t ext_k,
                                 const void* ext_omega32, const void* g_coset32, const void* g_coset_inv32);
int h2b_extended_to_coeff(h2b_ctx* ctx, const void* a_ne, void* out, uint32_t ext_k, const void* ext_omega_inv32,
                          const void* ext_ifft_divisor32, const void* g_coset32, const void* g_coset_inv32,
                          size_t out_len, const void* t_evals32_or_null, size_t n_t);
int h2b_extended_to_coeff_device(h2b_ctx* ctx, const void* d_a_ne, void* d_out, uint32_t ext_k,
                                 const void* ext_omega_inv32, const void* ext_ifft_divisor32, const void* g_coset32,
                                 const void* g_coset_inv32, size_t out_len, const void* t_evals32_or_null, size_t n_t);
int h2b_ntt_batch(h2b_ctx* ctx, void* const* cols, size_t n_cols, uint32_t log_n, const void* omega32,
                  const void* scale32_or_null);
int h2b_coeff_to_extended_batch(h2b_ctx* ctx, const void* const* in_cols, void* const* out_cols, size_t n_cols,
                                uint32_t k, uint32_t ext_k, const void* ext_omega32, const void* g_coset32,
                                const void* g_coset_inv32);
/* the same for device-resident columns (all on device 0 of the context): column c is transformed on device c mod G, moved there and
 * back by peer copies over NVLink (the independent per-column coset NTTs of evaluate_h, plonk/evaluation.rs:344-361). */
int h2b_coeff_to_extended_batch_device(h2b_ctx* ctx, const void* const* d_in_cols, void* const* d_out_cols, size_t n_cols,
                                       uint32_t k, uint32_t ext_k, const void* ext_omega32, const void* g_coset32,
                                       const void* g_coset_inv32);

/* ---- polynomial helpers between the accelerated steps ------------------------------------
 * The serial O(n) loops of create_proof that sit between MSM and NTT calls, as tiled prefix computations.
 * Elements are 32-byte Montgomery Fr.  *_device variants take device-0 pointers and leave results on the device.
 *
 * h2b_batch_invert   = ff::BatchInvert over a slice (permutation/prover.rs:124, lookup/prover.rs:225,
 *                      shuffle/prover.rs:150): a[i] <- 1 / a[i] in place; zeros are left untouched.
 * h2b_grand_product  = the running products of the Z polynomials (permutation/prover.rs:160-166,
 *                      lookup/prover.rs:254-259, shuffle/prover.rs:161-166): z[0] = init, z[i] = z[i-1] * v[i-1],
 *                      for i < out_len (out_len <= n + 1).  z may alias v.
 * h2b_eval_polynomial = arithmetic.rs:56-82 eval_polynomial: out = sum_i poly[i] * point^i  (n = 0 -> 0).
 * h2b_kate_division  = arithmetic.rs:99-120 kate_division: q (n - 1 entries) = a(X) / (X - b), remainder dropped;
 *                      n = 0 is an error (the reference panics on `a.len() - 1`). */
int h2b_batch_invert(h2b_ctx* ctx, void* a, size_t n);
int h2b_batch_invert_device(h2b_ctx* ctx, void* d_a, size_t n);
/* h2b_batch_invert_device_noalloc: the same inversion (same kernels, bit-identical results) for callers that must not allocate.
 *                      It has no counterpart in the reference, and none among the calls above either: it takes no context.
 *                      Who needs it: a prover that has already taken all of HBM into its own arena (a cudaMalloc would fail), a
 *                      pipeline that cannot afford the device-wide stall of a cudaMalloc / cudaFree while other streams run, or a
 *                      component that owns a buffer and a stream but no h2b_ctx (h2b_init allocates; so does the first call of a
 *                      new size on a context).  This call performs no cudaMalloc, cudaFree or other allocating call, ever.
 *                      Its tile scratch is a fixed H2B_NOALLOC_SCRATCH_BYTES array inside the library image: every GPU that
 *                      loads libh2b200.so carries it until the process exits, used or not.
 *   d_a          n Montgomery Fr on the CURRENT device (cudaSetDevice is the caller's), inverted in place; zeros stay zero.
 *   n            at most H2B_NOALLOC_MAX_ELEMS; more is refused with H2B_E_INVALID and a message (use the context call).
 *   cuda_stream  a cudaStream_t (NULL = the default stream).  Unlike the rest of this header the call is NOT synchronous: it
 *                returns once the three kernels are enqueued; the result is ready when the stream gets there.
 *   error_out    optional; receives a static string ("" on success) -- there is no context to keep a message in.
 *   The first call on each device looks the array's device address up (cudaGetSymbolAddress); make that first call before
 *   any stream capture begins.  All calls share the one array per GPU and it keeps the previous call's contents (nothing in
 *   it needs to start at zero).  Calls are safe back to back on one stream; two calls that can run at the same time on one GPU
 *   -- different streams with no dependency between them -- race on the scratch and return garbage: order them. */
#define H2B_NOALLOC_MAX_ELEMS (1u << 26)
#define H2B_NOALLOC_SCRATCH_BYTES (2u << 20)
int h2b_batch_invert_device_noalloc(void* d_a, size_t n, void* cuda_stream, const char** error_out);
/* h2b_fr_mul_device: d_out[i] = d_a[i] * d_b[i] (* scale32 when it is not NULL), n device-resident field elements; d_out may
 * alias d_a.  The coefficient scaling a_t * s^t that puts a polynomial on the coset s * H before a size-n transform --
 * distribute_powers_zeta (poly/domain.rs:325-341) for an arbitrary shift, given as its power table -- and the inverse scaling
 * after the inverse transform; what evaluating the quotient coset by coset needs (plonk/cosets.py).  An API extension. */
int h2b_fr_mul_device(h2b_ctx* ctx, const void* d_a, const void* d_b, const void* scale32_or_null, void* d_out, size_t n);
int h2b_grand_product(h2b_ctx* ctx, const void* v, size_t n, const void* init32, void* z, size_t out_len);
int h2b_grand_product_device(h2b_ctx* ctx, const void* d_v, size_t n, const void* init32, void* d_z, size_t out_len);
int h2b_eval_polynomial(h2b_ctx* ctx, const void* poly, size_t n, const void* point32, void* out32);
int h2b_eval_polynomial_device(h2b_ctx* ctx, const void* d_poly, size_t n, const void* point32, void* out32);
/* h2b_eval_polynomial_batch_device: out32s[i] = eval_polynomial(d_polys[i], points32[i]) for `count` device-resident polynomials
 * of n coefficients each -- the evaluations create_proof writes after the challenge x (plonk/prover.rs:742-811: advice, fixed,
 * random, permutation, lookup and shuffle polynomials at x and its rotations) and the ones the multi-open argument takes
 * (poly/query.rs:62-64 via shplonk/prover.rs:117-131) in two launches and one result copy instead of three launches and a
 * synchronisation each.  points32 = count x 32 bytes (Montgomery), out32s = count x 32 bytes on the host.  An API extension. */
int h2b_eval_polynomial_batch_device(h2b_ctx* ctx, const void* const* d_polys, size_t n, const void* points32, size_t count,
                                     void* out32s);
int h2b_kate_division(h2b_ctx* ctx, const void* a, size_t n, const void* b32, void* q);
int h2b_kate_division_device(h2b_ctx* ctx, const void* d_a, size_t n, const void* b32, void* d_q);

/* ---- quotient evaluation: the pointwise part of Evaluator::evaluate_h ----------------------
 * Everything is device resident (device 0): the inputs are the extended-domain columns coeff_to_extended produced, the
 * output feeds extended_to_coeff.  size = 2^ext_k rows; rotations are index offsets (idx + rot * 2^(ext_k-k)) mod size
 * (plonk/evaluation.rs:21-24).
 *
 * h2b_graph             = the reference's GraphEvaluator (evaluation.rs:187-196) flattened: `constants`, `rotations`, and
 *                         the `calculations` as 32-bit words.  A ValueSource (evaluation.rs:26-47) is two words,
 *                         w0 = kind | rotation_index << 8 and w1 = index, with kind 0 Constant, 1 Intermediate, 2 Fixed,
 *                         3 Advice, 4 Instance, 5 Challenge, 6 Beta, 7 Gamma, 8 Theta, 9 Y, 10 PreviousValue.  A Calculation
 *                         (evaluation.rs:101-121) is [op, target, a.w0, a.w1, b.w0, b.w1] with op 0 Add, 1 Sub, 2 Mul,
 *                         3 Square, 4 Double, 5 Negate, 7 Store, and Horner is [6, target, start.w0, start.w1, factor.w0,
 *                         factor.w1, n_parts, part.w0, part.w1, ...].
 * h2b_quotient_graph_device = GraphEvaluator::evaluate (evaluation.rs:789-834) for every row: out[idx] = value of the last
 *                         calculation (0 if there is none); `d_prev` supplies PreviousValue per row (NULL = zero; may alias d_out).
 *                         Indices are validated on the host; a malformed program is H2B_E_INVALID.
 * h2b_quotient_permutation_device = evaluation.rs:397-493 in place on d_values.  d_sets: permutation_product_coset of each set;
 *                         d_col_values / d_col_perms: the cosets of the permuted columns and pk.permutation.cosets, in
 *                         p.columns order, cut into sets of chunk_len; delta_start = beta * ZETA, delta = Fr::DELTA,
 *                         ext_omega = domain.get_extended_omega(); last_rotation = -(blinding_factors + 1).
 * h2b_quotient_lookup_device = evaluation.rs:517-565 in place (d_table_values = the lookup GraphEvaluator's output).
 * h2b_quotient_shuffle_device = evaluation.rs:573-620 in place (input / shuffle GraphEvaluator outputs). */
typedef struct h2b_graph {
    const uint32_t* program;  size_t n_words;
    const void* constants32;  size_t n_constants;
    const int32_t* rotations; size_t n_rotations;
    uint32_t n_intermediates;
} h2b_graph;
int h2b_quotient_graph_device(h2b_ctx* ctx, const h2b_graph* g, const void* const* d_fixed, size_t n_fixed,
                              const void* const* d_advice, size_t n_advice, const void* const* d_instance, size_t n_instance,
                              const void* challenges32, size_t n_challenges, const void* beta32, const void* gamma32,
                              const void* theta32, const void* y32, const void* d_prev_or_null, void* d_out, uint32_t k,
                              uint32_t ext_k);
int h2b_quotient_permutation_device(h2b_ctx* ctx, void* d_values, const void* const* d_sets, size_t n_sets,
                                    const void* const* d_col_values, const void* const* d_col_perms, size_t n_cols,
                                    size_t chunk_len, const void* d_l0, const void* d_l_last, const void* d_l_active,
                                    const void* beta32, const void* gamma32, const void* y32, const void* delta_start32,
                                    const void* delta32, const void* ext_omega32, int32_t last_rotation, uint32_t k,
                                    uint32_t ext_k);
int h2b_quotient_lookup_device(h2b_ctx* ctx, void* d_values, const void* d_table_values, const void* d_product,
                               const void* d_permuted_input, const void* d_permuted_table, const void* d_l0,
                               const void* d_l_last, const void* d_l_active, const void* beta32, const void* gamma32,
                               const void* y32, uint32_t k, uint32_t ext_k);
int h2b_quotient_shuffle_device(h2b_ctx* ctx, void* d_values, const void* d_input_values, const void* d_shuffle_values,
                                const void* d_product, const void* d_l0, const void* d_l_last, const void* d_l_active,
                                const void* y32, uint32_t k, uint32_t ext_k);

/* ---- measurement hooks (bench.py / profiles) -----------------------------------------------------------
 * h2b_last_call_ms: device time of the most recent call on this context, CUDA events on the launching
 *                   stream from the first to the last enqueued operation (copies included if the call copied).
 * h2b_profile: when enabled every kernel launch is bracketed by events; h2b_profile_read writes a JSON object
 *              {"kernel": {"ms": total, "launches": count}, ...} (and resets the counters). */
double h2b_last_call_ms(const h2b_ctx* ctx);
int h2b_profile_enable(h2b_ctx* ctx, int on);
int h2b_profile_read(h2b_ctx* ctx, char* json_out, size_t cap);
uint64_t h2b_launch_count(const h2b_ctx* ctx);   /* kernels launched by this context since h2b_init */

/* synthetic benchmark input: d_out[i] = [t0 + i]G as 64-byte affine points, on device 0. */
int h2b_synth_bases(h2b_ctx* ctx, void* d_out, size_t n, uint64_t t0);

/* ---- lookup argument ---------------------------------------------------------------------------------------------------
 * h2b_lookup_permute_device = permute_expression_pair (plonk/lookup/prover.rs:398-494) on the first `usable` of n rows of the two
 *                       compressed columns (device 0, Montgomery Fr): d_out_input = the input values sorted by canonical value
 *                       (`Fr: Ord`), d_out_table = the table values arranged so that every first occurrence of an input value sits
 *                       beside the same table value and the left-over table values fill the repeated rows from the last one backwards
 *                       in ascending order (the reference's BTreeMap walk + `repeated_input_rows.pop()`).  Rows usable..n of both
 *                       outputs are zeroed (the caller blinds them, :469-471).  H2B_E_CONSTRAINT when an input value does not
 *                       occur in the table (Error::ConstraintSystemFailure, :440).  Outputs must not alias the inputs. */
int h2b_lookup_permute_device(h2b_ctx* ctx, const void* d_input, const void* d_table, size_t usable, size_t n, void* d_out_input,
                              void* d_out_table);

/* ---- set-up helpers (not on the proving hot path) -----------------------------------------------------------------
 * h2b_g1_fixed_base_mul_device: d_out[i] = [k_i] G (64-byte affine, G = (1, 2)) for n Montgomery scalars on device 0: the fixed-base
 *                       multiplications of ParamsKZG::setup (poly/kzg/commitment.rs:74-112: g[i] = [s^i] G, g_lagrange[i] = [L_i(s)] G).
 * h2b_fr_from_u512_device: `Fr::from_u512` / `from_uniform_bytes` / `Fr::random` (halo2curves 0.6): n x 64 bytes, each a little-endian
 *                       512-bit integer, reduced mod r into Montgomery form (vanishing/prover.rs:76-81 draws n of them). */
int h2b_g1_fixed_base_mul_device(h2b_ctx* ctx, const void* d_scalars, size_t n, void* d_out_affine);
int h2b_fr_from_u512_device(h2b_ctx* ctx, const void* d_limbs64, size_t n, void* d_out);
/* h2b_fr_random_chacha20_device: the random polynomial of the vanishing argument (plonk/vanishing/prover.rs:57-81).  Stream i is
 *                       rand_chacha::ChaCha20Rng::from_seed(seeds32[i]) (host, n_streams x 32 bytes) and fills
 *                       d_out[offsets[i] .. offsets[i+1]) (host offsets, n_streams + 1 entries, offsets[0] = 0) with consecutive
 *                       Fr::random draws: element j of a stream = from_u512(ChaCha20 block j).  One thread per element. */
int h2b_fr_random_chacha20_device(h2b_ctx* ctx, const void* seeds32, const uint64_t* offsets, size_t n_streams, void* d_out);

/* integer-multiply pipe calibration: runs a dependent-free IMAD.WIDE loop on every SM and returns the
 * sustained rate in units of 32x32->64 multiply-adds per second (the MSM / NTT roofline denominator). */
int h2b_imad_peak(h2b_ctx* ctx, double* wide_mads_per_s, double* field_muls_per_s);

#ifdef __cplusplus
}
#endif
#endif /* H2B200_H */
