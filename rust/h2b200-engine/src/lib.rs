//! `B200Engine`: a GPU implementation of the reference's ZK-acceleration-layer trait
//! (`halo2_middleware::zal::traits::MsmAccel`, zal.rs:57-103) over libh2b200.so, plus the three
//! `EvaluationDomain` transforms (halo2_backend/src/poly/domain.rs:216,230,271) as free functions on raw slices.
//!
//! Source only (no Rust toolchain in the build image).  Soundness of the pointer casts rests on halo2curves 0.6
//! `bn256::Fr` / `Fq` being `#[repr(transparent)]`-like wrappers of `[u64; 4]` in Montgomery form and `G1Affine`
//! being `{x: Fq, y: Fq}`; the static assertions below fail the build otherwise.
use core::ffi::c_void;
use std::collections::HashMap;
use std::sync::Mutex;

use h2b200_sys as sys;
use halo2_middleware::zal::traits::MsmAccel;
use halo2curves::bn256::{Fq, Fr, G1Affine, G1};

const _: () = assert!(core::mem::size_of::<Fr>() == 32 && core::mem::size_of::<G1Affine>() == 64 && core::mem::size_of::<G1>() == 96);

pub struct B200Engine {
    ctx: *mut sys::h2b_ctx,
    /// resident bases keyed by (pointer, len) of the slice they were registered from
    cache: Mutex<HashMap<(usize, usize), u64>>,
}
unsafe impl Send for B200Engine {}
unsafe impl Sync for B200Engine {} // every C entry point locks the context

impl B200Engine {
    /// `devices`: CUDA ordinals (zal.rs:27-30 "the device(s) to run on"); empty = device 0.
    pub fn new(devices: &[i32]) -> Self {
        let mut ctx = core::ptr::null_mut();
        let rc = unsafe { sys::h2b_init(&mut ctx, if devices.is_empty() { core::ptr::null() } else { devices.as_ptr() }, devices.len() as i32) };
        assert_eq!(rc, sys::H2B_OK, "h2b_init failed: {}", last_error(core::ptr::null()));
        Self { ctx, cache: Mutex::new(HashMap::new()) }
    }
    fn check(&self, rc: i32) {
        // MsmAccel::msm returns C::Curve, not Result (zal.rs:58): the only error channel is a panic.
        assert_eq!(rc, sys::H2B_OK, "h2b200: {}", last_error(self.ctx));
    }
    fn handle_for(&self, base: &[G1Affine]) -> u64 {
        let key = (base.as_ptr() as usize, base.len());
        let mut cache = self.cache.lock().unwrap();
        *cache.entry(key).or_insert_with(|| {
            let mut h = 0u64;
            self.check(unsafe { sys::h2b_bases_register(self.ctx, base.as_ptr() as *const c_void, base.len(), &mut h) });
            h
        })
    }
    pub fn raw(&self) -> *mut sys::h2b_ctx { self.ctx }
}
impl Drop for B200Engine {
    fn drop(&mut self) { unsafe { sys::h2b_destroy(self.ctx) } }
}
fn last_error(ctx: *const sys::h2b_ctx) -> String {
    unsafe { std::ffi::CStr::from_ptr(sys::h2b_last_error(ctx)).to_string_lossy().into_owned() }
}

pub struct B200BaseDesc<'b> { handle: u64, len: usize, _raw: &'b [G1Affine] }
pub struct B200CoeffsDesc<'c> { raw: &'c [Fr] }

impl MsmAccel<G1Affine> for B200Engine {
    fn msm(&self, coeffs: &[Fr], bases: &[G1Affine]) -> G1 {
        assert_eq!(coeffs.len(), bases.len()); // best_multiexp asserts the same
        // ParamsKZG::commit{,_lagrange} pass `&self.g[0..size]` / `&self.g_lagrange[0..size]` on every call
        // (kzg/commitment.rs:313-316,362-365): same pointer each time, so the SRS is uploaded once.
        let h = self.handle_for(bases);
        let mut out = core::mem::MaybeUninit::<G1>::uninit();
        self.check(unsafe { sys::h2b_msm(self.ctx, h, coeffs.as_ptr() as *const c_void, coeffs.len(), out.as_mut_ptr() as *mut c_void) });
        unsafe { out.assume_init() }
    }
    type CoeffsDescriptor<'c> = B200CoeffsDesc<'c>;
    type BaseDescriptor<'b> = B200BaseDesc<'b>;
    fn get_coeffs_descriptor<'c>(&self, coeffs: &'c [Fr]) -> B200CoeffsDesc<'c> { B200CoeffsDesc { raw: coeffs } }
    fn get_base_descriptor<'b>(&self, base: &'b [G1Affine]) -> B200BaseDesc<'b> {
        B200BaseDesc { handle: self.handle_for(base), len: base.len(), _raw: base }
    }
    fn msm_with_cached_scalars(&self, coeffs: &B200CoeffsDesc<'_>, base: &[G1Affine]) -> G1 { self.msm(coeffs.raw, base) }
    fn msm_with_cached_base(&self, coeffs: &[Fr], base: &B200BaseDesc<'_>) -> G1 {
        assert!(coeffs.len() <= base.len);
        let mut out = core::mem::MaybeUninit::<G1>::uninit();
        self.check(unsafe { sys::h2b_msm(self.ctx, base.handle, coeffs.as_ptr() as *const c_void, coeffs.len(), out.as_mut_ptr() as *mut c_void) });
        unsafe { out.assume_init() }
    }
    fn msm_with_cached_inputs(&self, coeffs: &B200CoeffsDesc<'_>, base: &B200BaseDesc<'_>) -> G1 {
        self.msm_with_cached_base(coeffs.raw, base)
    }
}

/// The EvaluationDomain constants the C ABI needs, copied out of the reference struct (domain.rs:18-33).
pub struct DomainConsts {
    pub k: u32, pub extended_k: u32, pub quotient_poly_degree: u64,
    pub omega_inv: Fr, pub ifft_divisor: Fr, pub extended_omega: Fr, pub extended_omega_inv: Fr,
    pub extended_ifft_divisor: Fr, pub g_coset: Fr, pub g_coset_inv: Fr,
}

/// domain.rs:216-226
pub fn lagrange_to_coeff(e: &B200Engine, d: &DomainConsts, a: &mut [Fr]) {
    assert_eq!(a.len(), 1 << d.k);
    e.check(unsafe { sys::h2b_ntt(e.ctx, a.as_mut_ptr() as *mut c_void, d.k, &d.omega_inv as *const Fr as *const c_void,
                                  &d.ifft_divisor as *const Fr as *const c_void) });
}
/// domain.rs:230-244 (`a.values.resize(extended_len)` + best_fft become one call writing `out`)
pub fn coeff_to_extended(e: &B200Engine, d: &DomainConsts, a: &[Fr], out: &mut Vec<Fr>) {
    assert_eq!(a.len(), 1 << d.k);
    out.clear();
    out.reserve_exact(1 << d.extended_k);
    e.check(unsafe { sys::h2b_coeff_to_extended(e.ctx, a.as_ptr() as *const c_void, out.as_mut_ptr() as *mut c_void, d.k, d.extended_k,
                                                &d.extended_omega as *const Fr as *const c_void, &d.g_coset as *const Fr as *const c_void,
                                                &d.g_coset_inv as *const Fr as *const c_void) });
    unsafe { out.set_len(1 << d.extended_k) };
}
/// domain.rs:271-293; `t_evals` = Some(&domain.t_evaluations) fuses divide_by_vanishing_poly (domain.rs:297-316)
pub fn extended_to_coeff(e: &B200Engine, d: &DomainConsts, a: &[Fr], t_evals: Option<&[Fr]>) -> Vec<Fr> {
    assert_eq!(a.len(), 1 << d.extended_k);
    let out_len = ((1u64 << d.k) * d.quotient_poly_degree) as usize;
    let mut out: Vec<Fr> = Vec::with_capacity(out_len);
    let (tp, tn) = t_evals.map_or((core::ptr::null(), 0), |t| (t.as_ptr() as *const c_void, t.len()));
    e.check(unsafe { sys::h2b_extended_to_coeff(e.ctx, a.as_ptr() as *const c_void, out.as_mut_ptr() as *mut c_void, d.extended_k,
                                                &d.extended_omega_inv as *const Fr as *const c_void,
                                                &d.extended_ifft_divisor as *const Fr as *const c_void,
                                                &d.g_coset as *const Fr as *const c_void, &d.g_coset_inv as *const Fr as *const c_void,
                                                out_len, tp, tn) });
    unsafe { out.set_len(out_len) };
    out
}

// ---- the O(n) helpers between the accelerated steps -------------------------------------------------
fn fr_ptr(x: &Fr) -> *const c_void { x as *const Fr as *const c_void }

/// arithmetic.rs:56-82
pub fn eval_polynomial(e: &B200Engine, poly: &[Fr], point: Fr) -> Fr {
    let mut out = core::mem::MaybeUninit::<Fr>::uninit();
    e.check(unsafe { sys::h2b_eval_polynomial(e.ctx, poly.as_ptr() as *const c_void, poly.len(), fr_ptr(&point), out.as_mut_ptr() as *mut c_void) });
    unsafe { out.assume_init() }
}
/// arithmetic.rs:99-120
pub fn kate_division(e: &B200Engine, a: &[Fr], b: Fr) -> Vec<Fr> {
    let mut q: Vec<Fr> = Vec::with_capacity(a.len() - 1);           // panics on an empty slice, as the reference does
    e.check(unsafe { sys::h2b_kate_division(e.ctx, a.as_ptr() as *const c_void, a.len(), fr_ptr(&b), q.as_mut_ptr() as *mut c_void) });
    unsafe { q.set_len(a.len() - 1) };
    q
}
/// `values.iter_mut().batch_invert()` (permutation/prover.rs:124, lookup/prover.rs:225, shuffle/prover.rs:150)
pub fn batch_invert(e: &B200Engine, a: &mut [Fr]) {
    e.check(unsafe { sys::h2b_batch_invert(e.ctx, a.as_mut_ptr() as *mut c_void, a.len()) });
}
/// z[0] = init; z[i] = z[i-1] * v[i-1]  (permutation/prover.rs:160-166, lookup/prover.rs:254-259, shuffle/prover.rs:161-166)
pub fn grand_product(e: &B200Engine, v: &[Fr], init: Fr, out_len: usize) -> Vec<Fr> {
    assert!(out_len <= v.len() + 1);
    let mut z: Vec<Fr> = Vec::with_capacity(out_len);
    e.check(unsafe { sys::h2b_grand_product(e.ctx, v.as_ptr() as *const c_void, v.len(), fr_ptr(&init), z.as_mut_ptr() as *mut c_void, out_len) });
    unsafe { z.set_len(out_len) };
    z
}

// ---- quotient evaluation --------------------------------------------------------------------------------
/// Flat encoding of `GraphEvaluator` (plonk/evaluation.rs:187-196) as `h2b_graph` wants it.  In the reference fork this is built
/// inside halo2_backend (the struct's fields are private): one `push_source` per `ValueSource`, one record per `CalculationInfo`.
#[derive(Default)]
pub struct GraphProgram { pub words: Vec<u32>, pub constants: Vec<Fr>, pub rotations: Vec<i32>, pub n_intermediates: u32 }
impl GraphProgram {
    /// kind: 0 Constant, 1 Intermediate, 2 Fixed, 3 Advice, 4 Instance, 5 Challenge, 6 Beta, 7 Gamma, 8 Theta, 9 Y, 10 PreviousValue
    pub fn push_source(&mut self, kind: u32, index: u32, rotation_index: u32) { self.words.push(kind | (rotation_index << 8)); self.words.push(index); }
    /// op: 0 Add, 1 Sub, 2 Mul, 3 Square, 4 Double, 5 Negate, 7 Store (two sources follow; unary ops repeat Constant(0)); 6 Horner
    /// (start, factor, n_parts, parts...).
    pub fn begin(&mut self, op: u32, target: u32) { self.words.push(op); self.words.push(target); }
    pub fn as_raw(&self) -> sys::h2b_graph {
        sys::h2b_graph { program: self.words.as_ptr(), n_words: self.words.len(), constants32: self.constants.as_ptr() as *const c_void,
                         n_constants: self.constants.len(), rotations: self.rotations.as_ptr(), n_rotations: self.rotations.len(),
                         n_intermediates: self.n_intermediates }
    }
}
#[allow(dead_code)]
fn _layout_witness(_: Fq) {}
