//! Raw FFI declarations for libh2b200.so (include/h2b200.h).  Source only: no Rust toolchain exists in the build
//! image, so this crate has not been compiled there; it is the binding a maintainer of the reference adds.
#![allow(non_camel_case_types)]
use core::ffi::{c_char, c_int, c_void};

#[repr(C)]
pub struct h2b_ctx {
    _private: [u8; 0],
}

pub const H2B_OK: c_int = 0;
pub const H2B_MSM_WINDOW_BYTES: usize = 128;
pub const H2B_MSM_MAX_WINDOWS: usize = 64;
pub const H2B_PEER_HANDLE_BYTES: usize = 64;

/// `h2b_graph` of include/h2b200.h: the reference's `GraphEvaluator` (plonk/evaluation.rs:187-196) flattened.
#[repr(C)]
pub struct h2b_graph {
    pub program: *const u32,
    pub n_words: usize,
    pub constants32: *const c_void,
    pub n_constants: usize,
    pub rotations: *const i32,
    pub n_rotations: usize,
    pub n_intermediates: u32,
}

#[link(name = "h2b200")]
extern "C" {
    pub fn h2b_init(out: *mut *mut h2b_ctx, device_ids: *const c_int, n_devices: c_int) -> c_int;
    pub fn h2b_destroy(ctx: *mut h2b_ctx);
    pub fn h2b_last_error(ctx: *const h2b_ctx) -> *const c_char;
    pub fn h2b_device_count() -> c_int;
    pub fn h2b_last_notice(ctx: *const h2b_ctx) -> *const c_char;
    pub fn h2b_set_option(ctx: *mut h2b_ctx, name: *const c_char, value: *const c_char) -> c_int;
    pub fn h2b_bases_info(ctx: *mut h2b_ctx, bases_handle: u64, window_bits_out: *mut u32, device_bytes_out: *mut u64) -> c_int;
    pub fn h2b_bases_register(ctx: *mut h2b_ctx, bases: *const c_void, n: usize, handle_out: *mut u64) -> c_int;
    pub fn h2b_bases_release(ctx: *mut h2b_ctx, handle: u64) -> c_int;
    pub fn h2b_msm(ctx: *mut h2b_ctx, bases_handle: u64, scalars: *const c_void, n: usize, out96: *mut c_void) -> c_int;
    pub fn h2b_msm_batch(ctx: *mut h2b_ctx, bases_handle: u64, scalars_cols: *const *const c_void, n_cols: usize, n: usize, out96s: *mut c_void) -> c_int;
    pub fn h2b_msm_batch_device(ctx: *mut h2b_ctx, bases_handle: u64, d_scalars_cols: *const *const c_void, n_cols: usize, n: usize, out96s: *mut c_void) -> c_int;
    pub fn h2b_msm_raw(ctx: *mut h2b_ctx, bases: *const c_void, scalars: *const c_void, n: usize, out96: *mut c_void) -> c_int;
    pub fn h2b_ntt(ctx: *mut h2b_ctx, a: *mut c_void, log_n: u32, omega32: *const c_void, scale32_or_null: *const c_void) -> c_int;
    pub fn h2b_coeff_to_extended(ctx: *mut h2b_ctx, in_n: *const c_void, out_ne: *mut c_void, k: u32, ext_k: u32,
                                 ext_omega32: *const c_void, g_coset32: *const c_void, g_coset_inv32: *const c_void) -> c_int;
    pub fn h2b_extended_to_coeff(ctx: *mut h2b_ctx, a_ne: *const c_void, out: *mut c_void, ext_k: u32, ext_omega_inv32: *const c_void,
                                 ext_ifft_divisor32: *const c_void, g_coset32: *const c_void, g_coset_inv32: *const c_void,
                                 out_len: usize, t_evals32_or_null: *const c_void, n_t: usize) -> c_int;
    pub fn h2b_ntt_batch(ctx: *mut h2b_ctx, cols: *const *mut c_void, n_cols: usize, log_n: u32, omega32: *const c_void,
                         scale32_or_null: *const c_void) -> c_int;
    pub fn h2b_coeff_to_extended_batch(ctx: *mut h2b_ctx, in_cols: *const *const c_void, out_cols: *const *mut c_void, n_cols: usize,
                                       k: u32, ext_k: u32, ext_omega32: *const c_void, g_coset32: *const c_void,
                                       g_coset_inv32: *const c_void) -> c_int;
    // device-resident variants (device-0 pointers; results stay on the device)
    pub fn h2b_bases_register_device(ctx: *mut h2b_ctx, d_bases: *const c_void, n: usize, handle_out: *mut u64) -> c_int;
    pub fn h2b_msm_device(ctx: *mut h2b_ctx, bases_handle: u64, d_scalars: *const c_void, n: usize, out96: *mut c_void) -> c_int;
    pub fn h2b_ntt_device(ctx: *mut h2b_ctx, d_a: *mut c_void, log_n: u32, omega32: *const c_void, scale32_or_null: *const c_void) -> c_int;
    pub fn h2b_coeff_to_extended_device(ctx: *mut h2b_ctx, d_in_n: *const c_void, d_out_ne: *mut c_void, k: u32, ext_k: u32,
                                        ext_omega32: *const c_void, g_coset32: *const c_void, g_coset_inv32: *const c_void) -> c_int;
    pub fn h2b_extended_to_coeff_device(ctx: *mut h2b_ctx, d_a_ne: *const c_void, d_out: *mut c_void, ext_k: u32,
                                        ext_omega_inv32: *const c_void, ext_ifft_divisor32: *const c_void, g_coset32: *const c_void,
                                        g_coset_inv32: *const c_void, out_len: usize, t_evals32_or_null: *const c_void, n_t: usize) -> c_int;
    // polynomial helpers (arithmetic.rs:56-120, ff::BatchInvert, the Z running products)
    pub fn h2b_batch_invert(ctx: *mut h2b_ctx, a: *mut c_void, n: usize) -> c_int;
    pub fn h2b_batch_invert_device(ctx: *mut h2b_ctx, d_a: *mut c_void, n: usize) -> c_int;
    // context-free, allocation-free variant (scratch inside the library image; n <= 2^26; asynchronous on `cuda_stream`)
    pub fn h2b_batch_invert_device_noalloc(d_a: *mut c_void, n: usize, cuda_stream: *mut c_void, error_out: *mut *const c_char) -> c_int;
    pub fn h2b_fr_mul_device(ctx: *mut h2b_ctx, d_a: *const c_void, d_b: *const c_void, scale32_or_null: *const c_void, d_out: *mut c_void, n: usize) -> c_int;
    pub fn h2b_grand_product(ctx: *mut h2b_ctx, v: *const c_void, n: usize, init32: *const c_void, z: *mut c_void, out_len: usize) -> c_int;
    pub fn h2b_grand_product_device(ctx: *mut h2b_ctx, d_v: *const c_void, n: usize, init32: *const c_void, d_z: *mut c_void, out_len: usize) -> c_int;
    pub fn h2b_eval_polynomial(ctx: *mut h2b_ctx, poly: *const c_void, n: usize, point32: *const c_void, out32: *mut c_void) -> c_int;
    pub fn h2b_eval_polynomial_device(ctx: *mut h2b_ctx, d_poly: *const c_void, n: usize, point32: *const c_void, out32: *mut c_void) -> c_int;
    pub fn h2b_eval_polynomial_batch_device(ctx: *mut h2b_ctx, d_polys: *const *const c_void, n: usize, points32: *const c_void,
                                            count: usize, out32s: *mut c_void) -> c_int;
    pub fn h2b_kate_division(ctx: *mut h2b_ctx, a: *const c_void, n: usize, b32: *const c_void, q: *mut c_void) -> c_int;
    pub fn h2b_kate_division_device(ctx: *mut h2b_ctx, d_a: *const c_void, n: usize, b32: *const c_void, d_q: *mut c_void) -> c_int;
    // quotient evaluation (plonk/evaluation.rs:317-623), device resident
    pub fn h2b_quotient_graph_device(ctx: *mut h2b_ctx, g: *const h2b_graph, d_fixed: *const *const c_void, n_fixed: usize,
                                     d_advice: *const *const c_void, n_advice: usize, d_instance: *const *const c_void, n_instance: usize,
                                     challenges32: *const c_void, n_challenges: usize, beta32: *const c_void, gamma32: *const c_void,
                                     theta32: *const c_void, y32: *const c_void, d_prev_or_null: *const c_void, d_out: *mut c_void,
                                     k: u32, ext_k: u32) -> c_int;
    pub fn h2b_quotient_permutation_device(ctx: *mut h2b_ctx, d_values: *mut c_void, d_sets: *const *const c_void, n_sets: usize,
                                           d_col_values: *const *const c_void, d_col_perms: *const *const c_void, n_cols: usize,
                                           chunk_len: usize, d_l0: *const c_void, d_l_last: *const c_void, d_l_active: *const c_void,
                                           beta32: *const c_void, gamma32: *const c_void, y32: *const c_void,
                                           delta_start32: *const c_void, delta32: *const c_void, ext_omega32: *const c_void,
                                           last_rotation: i32, k: u32, ext_k: u32) -> c_int;
    pub fn h2b_quotient_lookup_device(ctx: *mut h2b_ctx, d_values: *mut c_void, d_table_values: *const c_void, d_product: *const c_void,
                                      d_permuted_input: *const c_void, d_permuted_table: *const c_void, d_l0: *const c_void,
                                      d_l_last: *const c_void, d_l_active: *const c_void, beta32: *const c_void, gamma32: *const c_void,
                                      y32: *const c_void, k: u32, ext_k: u32) -> c_int;
    pub fn h2b_quotient_shuffle_device(ctx: *mut h2b_ctx, d_values: *mut c_void, d_input_values: *const c_void,
                                       d_shuffle_values: *const c_void, d_product: *const c_void, d_l0: *const c_void,
                                       d_l_last: *const c_void, d_l_active: *const c_void, y32: *const c_void, k: u32, ext_k: u32) -> c_int;
    // device-resident columns spread over the context's GPUs (the independent coset NTTs of evaluate_h, evaluation.rs:344-361)
    pub fn h2b_coeff_to_extended_batch_device(ctx: *mut h2b_ctx, d_in_cols: *const *const c_void, d_out_cols: *const *mut c_void,
                                              n_cols: usize, k: u32, ext_k: u32, ext_omega32: *const c_void,
                                              g_coset32: *const c_void, g_coset_inv32: *const c_void) -> c_int;
    // set-up and randomness helpers: ParamsKZG::setup (kzg/commitment.rs:64-131), Fr::from_uniform_bytes, and the vanishing
    // argument's per-thread ChaCha20 streams (vanishing/prover.rs:57-81)
    pub fn h2b_g1_fixed_base_mul_device(ctx: *mut h2b_ctx, d_scalars: *const c_void, n: usize, d_out_affine: *mut c_void) -> c_int;
    pub fn h2b_fr_from_u512_device(ctx: *mut h2b_ctx, d_limbs64: *const c_void, n: usize, d_out: *mut c_void) -> c_int;
    pub fn h2b_fr_random_chacha20_device(ctx: *mut h2b_ctx, seeds32: *const c_void, offsets: *const u64, n_streams: usize,
                                         d_out: *mut c_void) -> c_int;
    // permute_expression_pair of the lookup argument (lookup/prover.rs:398-494); -5 = ConstraintSystemFailure
    pub fn h2b_lookup_permute_device(ctx: *mut h2b_ctx, d_input: *const c_void, d_table: *const c_void, usable: usize, n: usize,
                                     d_out_input: *mut c_void, d_out_table: *mut c_void) -> c_int;
    // multi-process sharding (one rank per GPU): each rank returns the window sums of its point range, any rank folds them
    pub fn h2b_msm_windows(ctx: *mut h2b_ctx, bases_handle: u64, scalars: *const c_void, scalars_on_device: c_int, n: usize,
                           n_total: usize, out_windows: *mut c_void, c_out: *mut u32, windows_out: *mut u32) -> c_int;
    pub fn h2b_msm_combine(parts: *const c_void, n_parts: u32, c: u32, windows: u32, out96: *mut c_void) -> c_int;
    // the same sharding with the exchange inside the library: peer mailboxes over CUDA IPC / NVLink, folded on the device
    pub fn h2b_peer_export(ctx: *mut h2b_ctx, handle64_out: *mut c_void, local_ptr_out: *mut *mut c_void) -> c_int;
    pub fn h2b_peer_attach(ctx: *mut h2b_ctx, rank: c_int, world: c_int, handles: *const c_void, local_ptrs_or_null: *const *mut c_void) -> c_int;
    pub fn h2b_msm_sharded(ctx: *mut h2b_ctx, bases_handle: u64, scalars: *const c_void, scalars_on_device: c_int, n: usize,
                           n_total: usize, out96: *mut c_void) -> c_int;
    // measurement hooks and synthetic benchmark input (bench only)
    pub fn h2b_version() -> *const c_char;
    pub fn h2b_last_call_ms(ctx: *const h2b_ctx) -> f64;
    pub fn h2b_profile_enable(ctx: *mut h2b_ctx, on: c_int) -> c_int;
    pub fn h2b_profile_read(ctx: *mut h2b_ctx, json_out: *mut c_char, cap: usize) -> c_int;
    pub fn h2b_launch_count(ctx: *const h2b_ctx) -> u64;
    pub fn h2b_synth_bases(ctx: *mut h2b_ctx, d_out: *mut c_void, n: usize, t0: u64) -> c_int;
    pub fn h2b_imad_peak(ctx: *mut h2b_ctx, wide_mads_per_s: *mut f64, field_muls_per_s: *mut f64) -> c_int;
}
