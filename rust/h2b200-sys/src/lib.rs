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

#[link(name = "h2b200")]
extern "C" {
    pub fn h2b_init(out: *mut *mut h2b_ctx, device_ids: *const c_int, n_devices: c_int) -> c_int;
    pub fn h2b_destroy(ctx: *mut h2b_ctx);
    pub fn h2b_last_error(ctx: *const h2b_ctx) -> *const c_char;
    pub fn h2b_device_count() -> c_int;
    pub fn h2b_bases_register(ctx: *mut h2b_ctx, bases: *const c_void, n: usize, handle_out: *mut u64) -> c_int;
    pub fn h2b_bases_release(ctx: *mut h2b_ctx, handle: u64) -> c_int;
    pub fn h2b_msm(ctx: *mut h2b_ctx, bases_handle: u64, scalars: *const c_void, n: usize, out96: *mut c_void) -> c_int;
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
}
