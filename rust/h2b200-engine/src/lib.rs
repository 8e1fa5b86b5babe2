//! `B200Engine`: a GPU implementation of the reference's ZK-acceleration-layer trait
//! (`halo2_middleware::zal::traits::MsmAccel`, zal.rs:57-103) over libh2b200.so.  The `EvaluationDomain` side of the
//! boundary is a patch of halo2_backend itself (rust/patches/0001-halo2_backend-b200-ntt.patch): those methods have no trait.
//!
//! Source only (no Rust toolchain in the build image).  Soundness of the pointer casts rests on halo2curves 0.6
//! `bn256::Fr` / `Fq` being wrappers of `[u64; 4]` in Montgomery form and `G1Affine` being `{x: Fq, y: Fq}`
//!; the static assertion below fails the build otherwise.
//!
//! Residency rules (zal.rs:47-50, 71-77):
//! * `msm(coeffs, base)` has no lifetime to hang a cache on (zal.rs:58), and the reference also calls it with temporary base
//!   vectors (kzg/msm.rs:82 `MSMKZG::eval`): it never caches by address.  It uses resident bases only when the slice lies
//!   inside a range that is CURRENTLY registered through `register_srs`, whose guard borrows the slice for as long as the
//!   registration lives, so a freed-and-reused allocation cannot match; everything else goes through `h2b_msm_raw`.
//! * `get_base_descriptor(base)` registers, and the descriptor (lifetime `'b` = the borrow of `base`) releases on `Drop`.
use core::ffi::c_void;
use std::sync::{Arc, Mutex};

use h2b200_sys as sys;
use halo2_middleware::zal::traits::MsmAccel;
use halo2curves::bn256::{Fr, G1Affine, G1};

const _: () = assert!(core::mem::size_of::<Fr>() == 32 && core::mem::size_of::<G1Affine>() == 64 && core::mem::size_of::<G1>() == 96);

/// The library context, shared by the engine and every descriptor / registration it handed out (whichever is dropped last
/// destroys it, so a descriptor can never outlive the context it releases into).
struct Ctx(*mut sys::h2b_ctx);
unsafe impl Send for Ctx {}
unsafe impl Sync for Ctx {} // every C entry point locks the context
impl Drop for Ctx {
    fn drop(&mut self) { unsafe { sys::h2b_destroy(self.0) } }
}
impl Ctx {
    fn check(&self, rc: i32) {
        // MsmAccel::msm returns C::Curve, not Result (zal.rs:58): the only error channel is a panic.
        assert_eq!(rc, sys::H2B_OK, "h2b200: {}", last_error(self.0));
    }
    fn register(&self, base: &[G1Affine]) -> u64 {
        let mut h = 0u64;
        self.check(unsafe { sys::h2b_bases_register(self.0, base.as_ptr() as *const c_void, base.len(), &mut h) });
        h
    }
    fn msm_handle(&self, handle: u64, coeffs: &[Fr]) -> G1 {
        let mut out = core::mem::MaybeUninit::<G1>::uninit();
        self.check(unsafe { sys::h2b_msm(self.0, handle, coeffs.as_ptr() as *const c_void, coeffs.len(), out.as_mut_ptr() as *mut c_void) });
        unsafe { out.assume_init() }
    }
}
fn last_error(ctx: *const sys::h2b_ctx) -> String {
    unsafe { std::ffi::CStr::from_ptr(sys::h2b_last_error(ctx)).to_string_lossy().into_owned() }
}

/// address range of a registered SRS and its handle
struct Resident { start: usize, len: usize, handle: u64 }

pub struct B200Engine {
    ctx: Arc<Ctx>,
    resident: Arc<Mutex<Vec<Resident>>>,
}

/// Guard of `B200Engine::register_srs`: while it lives, `msm(coeffs, &srs[0..m])` runs on the resident copy.  It BORROWS the
/// slice, so the allocation cannot be freed, moved or mutated before the registration is gone.
pub struct ResidentSrs<'a> {
    ctx: Arc<Ctx>,
    resident: Arc<Mutex<Vec<Resident>>>,
    handle: u64,
    _srs: &'a [G1Affine],
}
impl Drop for ResidentSrs<'_> {
    fn drop(&mut self) {
        self.resident.lock().unwrap().retain(|r| r.handle != self.handle);
        unsafe { sys::h2b_bases_release(self.ctx.0, self.handle) };
    }
}

impl B200Engine {
    /// `devices`: CUDA ordinals (zal.rs:27-30 "the device(s) to run on"); empty = device 0.
    pub fn new(devices: &[i32]) -> Self {
        let mut raw = core::ptr::null_mut();
        let rc = unsafe { sys::h2b_init(&mut raw, if devices.is_empty() { core::ptr::null() } else { devices.as_ptr() }, devices.len() as i32) };
        assert_eq!(rc, sys::H2B_OK, "h2b_init failed: {}", last_error(core::ptr::null()));
        Self { ctx: Arc::new(Ctx(raw)), resident: Arc::new(Mutex::new(Vec::new())) }
    }
    /// Make an SRS resident (`ParamsKZG::g`, `g_lagrange`): `commit` / `commit_lagrange` pass `&g[0..size]` to plain `msm`
    /// on every call (kzg/commitment.rs:313-316, 362-365), which then finds the prefix of this registration.
    pub fn register_srs<'a>(&self, srs: &'a [G1Affine]) -> ResidentSrs<'a> {
        let handle = self.ctx.register(srs);
        self.resident.lock().unwrap().push(Resident { start: srs.as_ptr() as usize, len: srs.len(), handle });
        ResidentSrs { ctx: self.ctx.clone(), resident: self.resident.clone(), handle, _srs: srs }
    }
    fn resident_prefix(&self, base: &[G1Affine]) -> Option<u64> {
        let start = base.as_ptr() as usize;
        self.resident.lock().unwrap().iter().find(|r| r.start == start && base.len() <= r.len).map(|r| r.handle)
    }
    pub fn raw(&self) -> *mut sys::h2b_ctx { self.ctx.0 }
}

pub struct B200BaseDesc<'b> { ctx: Arc<Ctx>, handle: u64, len: usize, _raw: &'b [G1Affine] }
impl Drop for B200BaseDesc<'_> {
    fn drop(&mut self) { unsafe { sys::h2b_bases_release(self.ctx.0, self.handle) }; }
}
pub struct B200CoeffsDesc<'c> { raw: &'c [Fr] }

impl MsmAccel<G1Affine> for B200Engine {
    fn msm(&self, coeffs: &[Fr], bases: &[G1Affine]) -> G1 {
        assert_eq!(coeffs.len(), bases.len()); // best_multiexp asserts the same
        if let Some(handle) = self.resident_prefix(bases) {
            return self.ctx.msm_handle(handle, coeffs);
        }
        let mut out = core::mem::MaybeUninit::<G1>::uninit();
        self.ctx.check(unsafe {
            sys::h2b_msm_raw(self.ctx.0, bases.as_ptr() as *const c_void, coeffs.as_ptr() as *const c_void, coeffs.len(), out.as_mut_ptr() as *mut c_void)
        });
        unsafe { out.assume_init() }
    }
    type CoeffsDescriptor<'c> = B200CoeffsDesc<'c>;
    type BaseDescriptor<'b> = B200BaseDesc<'b>;
    fn get_coeffs_descriptor<'c>(&self, coeffs: &'c [Fr]) -> B200CoeffsDesc<'c> { B200CoeffsDesc { raw: coeffs } }
    fn get_base_descriptor<'b>(&self, base: &'b [G1Affine]) -> B200BaseDesc<'b> {
        B200BaseDesc { ctx: self.ctx.clone(), handle: self.ctx.register(base), len: base.len(), _raw: base }
    }
    fn msm_with_cached_scalars(&self, coeffs: &B200CoeffsDesc<'_>, base: &[G1Affine]) -> G1 { self.msm(coeffs.raw, base) }
    fn msm_with_cached_base(&self, coeffs: &[Fr], base: &B200BaseDesc<'_>) -> G1 {
        assert!(coeffs.len() <= base.len);
        self.ctx.msm_handle(base.handle, coeffs)
    }
    fn msm_with_cached_inputs(&self, coeffs: &B200CoeffsDesc<'_>, base: &B200BaseDesc<'_>) -> G1 {
        self.msm_with_cached_base(coeffs.raw, base)
    }
}
