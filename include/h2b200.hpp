// h2b200.hpp -- C++17 host side of libh2b200.so: the reference's operator interface for the MSM / NTT path, over the C ABI.
//
// The reference's host is Rust (no toolchain in this image); rust/h2b200-engine is that binding as source.  This header is the
// same interface for a compiled host that CAN be built here, so that the boundary is exercised by native code and the parity
// tests (tests/cpp/host_mirror_test.cpp) read like the reference's own:
//
//   B200Engine                 = impl MsmAccel<G1Affine>         halo2_middleware/src/zal.rs:57-103 (the seven trait items),
//                                                                 reference implementation H2cEngine zal.rs:113-172
//   ParamsKZG::commit*         = ParamsKZG::commit_lagrange / commit   halo2_backend/src/poly/kzg/commitment.rs:305-317, 354-366
//   EvaluationDomain           = halo2_backend/src/poly/domain.rs:36-144 (new), :216 lagrange_to_coeff, :230 coeff_to_extended,
//                                :271 extended_to_coeff, :297 divide_by_vanishing_poly, :380-396 rotate_omega
//   fr::                       = what the Rust host gets from halo2curves (Fr arithmetic, ROOT_OF_UNITY, ZETA, S): only O(1)-per-call
//                                constant work is done with it; every vector operation goes to the GPU.
//
// Error behaviour follows the reference: where it `assert!`s (slice lengths, commitment.rs:315, domain.rs:217,234,272) these
// throw std::invalid_argument before anything is launched; a failure inside the library throws h2b200::Error carrying the
// H2B_E_* code and h2b_last_error.  There is no CPU fallback: constructing a Context without a usable sm_100 device throws.
//
// Header-only; link with -lh2b200.  Plain data crosses the boundary: Fr = 32 B, G1Affine = 64 B, G1 = 96 B (h2b200.h).
#pragma once

#include <array>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "h2b200.h"

namespace h2b200 {

struct Fr { uint64_t l[4]; };                       // Montgomery form, R = 2^256, fully reduced (halo2curves bn256::Fr layout)
struct Fq { uint64_t l[4]; };
struct G1Affine { Fq x, y; };                       // identity = (0, 0)
struct G1 { Fq x, y, z; };                          // Jacobian; the library returns it normalised (z = 1, or (0, 1, 0))
static_assert(sizeof(Fr) == 32 && sizeof(G1Affine) == 64 && sizeof(G1) == 96, "layouts of h2b200.h");

inline bool operator==(const Fr& a, const Fr& b) { return std::memcmp(&a, &b, 32) == 0; }
inline bool operator!=(const Fr& a, const Fr& b) { return !(a == b); }
inline bool operator==(const G1& a, const G1& b) { return std::memcmp(&a, &b, 96) == 0; }
inline bool operator==(const G1Affine& a, const G1Affine& b) { return std::memcmp(&a, &b, 64) == 0; }

class Error : public std::runtime_error {
public:
    Error(int code, const std::string& what) : std::runtime_error(what), code_(code) {}
    int code() const { return code_; }
private:
    int code_;
};

// ---- Fr on the host: constants of an EvaluationDomain, nothing per element -----------------------------------------------------
namespace fr {
constexpr uint64_t MODULUS[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
constexpr uint64_t INV = 0xc2e1f593efffffffull;    // -r^-1 mod 2^64
constexpr uint32_t S = 28;                           // 2-adicity: r - 1 = 2^28 * odd

inline bool geq_modulus(const uint64_t a[4]) {
    for (int i = 3; i >= 0; i--) {
        if (a[i] > MODULUS[i]) return true;
        if (a[i] < MODULUS[i]) return false;
    }
    return true;
}
inline void sub_modulus(uint64_t a[4]) {
    unsigned __int128 borrow = 0;
    for (int i = 0; i < 4; i++) {
        unsigned __int128 d = (unsigned __int128)a[i] - MODULUS[i] - borrow;
        a[i] = (uint64_t)d;
        borrow = (d >> 64) & 1;
    }
}
inline Fr add(const Fr& a, const Fr& b) {
    Fr r;
    unsigned __int128 c = 0;
    for (int i = 0; i < 4; i++) { c += (unsigned __int128)a.l[i] + b.l[i]; r.l[i] = (uint64_t)c; c >>= 64; }
    if (c || geq_modulus(r.l)) sub_modulus(r.l);    // a + b < 2r < 2^255: no carry out, kept for clarity
    return r;
}
inline Fr neg(const Fr& a) {
    if ((a.l[0] | a.l[1] | a.l[2] | a.l[3]) == 0) return a;
    Fr r;
    unsigned __int128 borrow = 0;
    for (int i = 0; i < 4; i++) {
        unsigned __int128 d = (unsigned __int128)MODULUS[i] - a.l[i] - borrow;
        r.l[i] = (uint64_t)d;
        borrow = (d >> 64) & 1;
    }
    return r;
}
inline Fr sub(const Fr& a, const Fr& b) { return add(a, neg(b)); }
// Montgomery product a * b / 2^256 mod r (coarsely integrated operand scanning)
inline Fr mul(const Fr& a, const Fr& b) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        unsigned __int128 c = 0;
        for (int j = 0; j < 4; j++) { c += (unsigned __int128)a.l[j] * b.l[i] + t[j]; t[j] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[4] = (uint64_t)c; t[5] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * INV;
        c = (unsigned __int128)m * MODULUS[0] + t[0];
        c >>= 64;
        for (int j = 1; j < 4; j++) { c += (unsigned __int128)m * MODULUS[j] + t[j]; t[j - 1] = (uint64_t)c; c >>= 64; }
        c += t[4]; t[3] = (uint64_t)c; t[4] = t[5] + (uint64_t)(c >> 64);
    }
    Fr r = {{t[0], t[1], t[2], t[3]}};
    if (t[4] || geq_modulus(r.l)) sub_modulus(r.l);
    return r;
}
inline Fr square(const Fr& a) { return mul(a, a); }
inline Fr zero() { return Fr{{0, 0, 0, 0}}; }
// R^2 mod r, so that mul(x_plain, R2) = x * R = the Montgomery form of x
inline Fr r2() { return Fr{{0x1bb8e645ae216da7ull, 0x53fe3ab1e35c59e3ull, 0x8c49833d53bb8085ull, 0x0216d0b17f4e44a5ull}}; }
inline Fr from_u64(uint64_t v) { return mul(Fr{{v, 0, 0, 0}}, r2()); }
inline Fr one() { return from_u64(1); }
inline Fr from_canonical(const uint64_t limbs_le[4]) { return mul(Fr{{limbs_le[0], limbs_le[1], limbs_le[2], limbs_le[3]}}, r2()); }
inline std::array<uint64_t, 4> to_canonical(const Fr& a) {
    Fr c = mul(a, Fr{{1, 0, 0, 0}});
    return {c.l[0], c.l[1], c.l[2], c.l[3]};
}
inline Fr pow(const Fr& a, const uint64_t e_le[4]) {    // pow_vartime
    Fr acc = one();
    for (int i = 255; i >= 0; i--) {
        acc = square(acc);
        if ((e_le[i >> 6] >> (i & 63)) & 1) acc = mul(acc, a);
    }
    return acc;
}
inline Fr pow(const Fr& a, uint64_t e) { const uint64_t le[4] = {e, 0, 0, 0}; return pow(a, le); }
inline Fr invert(const Fr& a) {                           // Fermat; the caller guarantees a != 0 (the reference unwraps)
    const uint64_t e[4] = {MODULUS[0] - 2, MODULUS[1], MODULUS[2], MODULUS[3]};
    return pow(a, e);
}
// Fr::ROOT_OF_UNITY = MULTIPLICATIVE_GENERATOR ^ ((r - 1) / 2^S) with generator 7, and Fr::ZETA, a primitive cube root of unity
// (halo2curves 0.6 bn256/fr.rs; the same values as yet-another-halo2-fork_b200/fields.py, checked there numerically).
inline Fr root_of_unity() {
    const uint64_t m[4] = {MODULUS[0] - 1, MODULUS[1], MODULUS[2], MODULUS[3]};
    uint64_t e[4];                                       // (r - 1) >> S
    for (int i = 0; i < 4; i++) e[i] = (m[i] >> S) | (i < 3 ? m[i + 1] << (64 - S) : 0);
    return pow(from_u64(7), e);
}
inline Fr zeta() {
    const uint64_t z[4] = {0xb8ca0b2d36636f23ull, 0xcc37a73fec2bc5e9ull, 0x048b6e193fd84104ull, 0x30644e72e131a029ull};
    return from_canonical(z);
}
}  // namespace fr

// ---- context: the engine object of zal.rs:182-243 (PlonkEngine { msm_backend }) ------------------------------------------------------
class Context {
public:
    // devices: CUDA device indices (empty = device 0).  Several devices shard an MSM by point range and deal batched NTT columns.
    explicit Context(const std::vector<int>& devices = {}) {
        int rc = h2b_init(&ctx_, devices.empty() ? nullptr : devices.data(), (int)devices.size());
        if (rc != H2B_OK) {
            std::string why = ctx_ ? h2b_last_error(ctx_) : "";
            if (ctx_) h2b_destroy(ctx_);
            ctx_ = nullptr;
            throw Error(rc, "h2b_init failed (" + std::to_string(rc) + (why.empty() ? ")" : "): " + why) +
                                "; this library has no CPU fallback");
        }
    }
    ~Context() { if (ctx_) h2b_destroy(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    h2b_ctx* raw() const { return ctx_; }
    void check(int rc, const char* what) const {
        if (rc != H2B_OK) throw Error(rc, std::string(what) + ": " + h2b_last_error(ctx_));
    }
    void set_option(const std::string& name, const std::string& value) { check(h2b_set_option(ctx_, name.c_str(), value.c_str()), "h2b_set_option"); }
    std::string last_notice() const { return h2b_last_notice(ctx_); }
    uint64_t launch_count() const { return h2b_launch_count(ctx_); }
    static int device_count() { return h2b_device_count(); }
private:
    h2b_ctx* ctx_ = nullptr;
};

// A borrowed slice (the reference's `&'c [Fr]` / `&'b [G1Affine]`): pointer + length, no ownership.
template <class T>
struct Slice {
    const T* ptr = nullptr;
    size_t len = 0;
    Slice() = default;
    Slice(const T* p, size_t n) : ptr(p), len(n) {}
    Slice(const std::vector<T>& v) : ptr(v.data()), len(v.size()) {}          // NOLINT: implicit on purpose, like &vec[..]
    const T* data() const { return ptr; }
    size_t size() const { return len; }
    Slice first(size_t n) const {
        if (n > len) throw std::invalid_argument("slice index out of range");
        return Slice(ptr, n);
    }
};

// MsmAccel::CoeffsDescriptor (zal.rs:63, 78-80): H2cEngine keeps the borrowed slice (zal.rs:115-118); so does this one -- the
// scalars change with every commit, there is nothing to precompute for them.
struct CoeffsDescriptor {
    Slice<Fr> raw;
};

// MsmAccel::BaseDescriptor (zal.rs:67-77, 82): the bases live on the device(s) from construction to destruction, with the window
// multiples the "expensive preprocessing" comment allows for.  Drop = h2b_bases_release (zal.rs:47-50).
class BaseDescriptor {
public:
    BaseDescriptor(std::shared_ptr<Context> ctx, Slice<G1Affine> bases) : ctx_(std::move(ctx)), len_(bases.size()) {
        ctx_->check(h2b_bases_register(ctx_->raw(), bases.data(), bases.size(), &handle_), "h2b_bases_register");
    }
    ~BaseDescriptor() { if (ctx_) h2b_bases_release(ctx_->raw(), handle_); }
    BaseDescriptor(BaseDescriptor&& o) noexcept : ctx_(std::move(o.ctx_)), handle_(o.handle_), len_(o.len_) { o.ctx_.reset(); }
    BaseDescriptor(const BaseDescriptor&) = delete;
    BaseDescriptor& operator=(const BaseDescriptor&) = delete;
    uint64_t handle() const { return handle_; }
    size_t size() const { return len_; }
    // (window bits of the precomputed multiples or 0, device bytes held)
    std::pair<uint32_t, uint64_t> info() const {
        uint32_t c = 0; uint64_t bytes = 0;
        ctx_->check(h2b_bases_info(ctx_->raw(), handle_, &c, &bytes), "h2b_bases_info");
        return {c, bytes};
    }
private:
    std::shared_ptr<Context> ctx_;
    uint64_t handle_ = 0;
    size_t len_ = 0;
};

// impl MsmAccel<G1Affine> for B200Engine -- zal.rs:57-103, argument meaning and panics of H2cEngine (zal.rs:124-171)
class B200Engine {
public:
    explicit B200Engine(std::shared_ptr<Context> ctx = std::make_shared<Context>()) : ctx_(std::move(ctx)) {}
    const std::shared_ptr<Context>& context() const { return ctx_; }

    // fn msm(&self, coeffs: &[Fr], base: &[G1Affine]) -> G1                                   zal.rs:58, H2cEngine :124-126
    // (best_multiexp asserts coeffs.len() == bases.len())
    G1 msm(Slice<Fr> coeffs, Slice<G1Affine> base) const {
        if (coeffs.size() != base.size()) throw std::invalid_argument("msm: coeffs.len() != bases.len()");
        G1 out;
        ctx_->check(h2b_msm_raw(ctx_->raw(), base.data(), coeffs.data(), coeffs.size(), &out), "h2b_msm_raw");
        return out;
    }
    // fn get_coeffs_descriptor<'c>(&self, coeffs: &'c [Fr]) -> CoeffsDescriptor<'c>            zal.rs:78
    CoeffsDescriptor get_coeffs_descriptor(Slice<Fr> coeffs) const { return CoeffsDescriptor{coeffs}; }
    // fn get_base_descriptor<'b>(&self, base: &'b [G1Affine]) -> BaseDescriptor<'b>            zal.rs:82
    BaseDescriptor get_base_descriptor(Slice<G1Affine> base) const { return BaseDescriptor(ctx_, base); }
    // fn msm_with_cached_scalars(&self, coeffs: &CoeffsDescriptor, base: &[G1Affine]) -> G1     zal.rs:84
    G1 msm_with_cached_scalars(const CoeffsDescriptor& coeffs, Slice<G1Affine> base) const { return msm(coeffs.raw, base); }
    // fn msm_with_cached_base(&self, coeffs: &[Fr], base: &BaseDescriptor) -> G1                zal.rs:90
    // bases[0..coeffs.len()) of the descriptor are used; more coefficients than bases is the reference's assert
    G1 msm_with_cached_base(Slice<Fr> coeffs, const BaseDescriptor& base) const {
        if (coeffs.size() > base.size()) throw std::invalid_argument("msm_with_cached_base: more coefficients than registered bases");
        G1 out;
        ctx_->check(h2b_msm(ctx_->raw(), base.handle(), coeffs.data(), coeffs.size(), &out), "h2b_msm");
        return out;
    }
    // fn msm_with_cached_inputs(&self, coeffs: &CoeffsDescriptor, base: &BaseDescriptor) -> G1   zal.rs:96
    G1 msm_with_cached_inputs(const CoeffsDescriptor& coeffs, const BaseDescriptor& base) const {
        return msm_with_cached_base(coeffs.raw, base);
    }
    // extension: the per-column commits of one prover phase in one pipeline (h2b_msm_batch)
    std::vector<G1> msm_batch(const std::vector<Slice<Fr>>& columns, const BaseDescriptor& base) const {
        std::vector<G1> out(columns.size());
        if (columns.empty()) return out;
        std::vector<const void*> ptrs;
        for (const auto& c : columns) {
            if (c.size() != columns[0].size()) throw std::invalid_argument("msm_batch: columns of different lengths");
            ptrs.push_back(c.data());
        }
        if (columns[0].size() > base.size()) throw std::invalid_argument("msm_batch: more coefficients than registered bases");
        ctx_->check(h2b_msm_batch(ctx_->raw(), base.handle(), ptrs.data(), ptrs.size(), columns[0].size(), out.data()), "h2b_msm_batch");
        return out;
    }
private:
    std::shared_ptr<Context> ctx_;
};

// ---- ParamsKZG: the commit path of halo2_backend/src/poly/kzg/commitment.rs ---------------------------------------------------------------
// Holds g and g_lagrange (kzg/commitment.rs:26-34) and, once an engine has been seen, their resident descriptors -- the caching
// half of ZAL (zal.rs:78-100) the reference defines but does not use in commit*.
class ParamsKZG {
public:
    ParamsKZG(uint32_t k, std::vector<G1Affine> g, std::vector<G1Affine> g_lagrange)
        : k_(k), n_((uint64_t)1 << k), g_(std::move(g)), g_lagrange_(std::move(g_lagrange)) {
        if (g_.size() != n_ || g_lagrange_.size() != n_) throw std::invalid_argument("ParamsKZG: g / g_lagrange must hold 2^k points");
    }
    uint32_t k() const { return k_; }
    uint64_t n() const { return n_; }
    const std::vector<G1Affine>& get_g() const { return g_; }
    const std::vector<G1Affine>& get_g_lagrange() const { return g_lagrange_; }

    // fn commit_lagrange(&self, engine, poly: &Polynomial<Fr, LagrangeCoeff>, _: Blind) -> G1     commitment.rs:305-317
    G1 commit_lagrange(const B200Engine& engine, Slice<Fr> poly) {
        if (g_lagrange_.size() < poly.size()) throw std::invalid_argument("commit_lagrange: assert!(bases.len() >= size)");
        if (!d_lagrange_) d_lagrange_ = std::make_unique<BaseDescriptor>(engine.get_base_descriptor(g_lagrange_));
        return engine.msm_with_cached_base(poly, *d_lagrange_);
    }
    // fn commit(&self, engine, poly: &Polynomial<Fr, Coeff>, _: Blind) -> G1                       commitment.rs:354-366
    G1 commit(const B200Engine& engine, Slice<Fr> poly) {
        if (g_.size() < poly.size()) throw std::invalid_argument("commit: assert!(bases.len() >= size)");
        if (!d_g_) d_g_ = std::make_unique<BaseDescriptor>(engine.get_base_descriptor(g_));
        return engine.msm_with_cached_base(poly, *d_g_);
    }
private:
    uint32_t k_;
    uint64_t n_;
    std::vector<G1Affine> g_, g_lagrange_;
    std::unique_ptr<BaseDescriptor> d_g_, d_lagrange_;
};

// ---- EvaluationDomain: halo2_backend/src/poly/domain.rs ----------------------------------------------------------------------------
// The fields of the reference's EvaluationDomain (domain.rs:18-33), computed as `new` computes them; needs no device.
struct DomainConstants {
    uint64_t n = 0;
    uint32_t k = 0, extended_k = 0;
    uint64_t quotient_poly_degree = 0;
    Fr omega, omega_inv, extended_omega, extended_omega_inv, g_coset, g_coset_inv, ifft_divisor, extended_ifft_divisor,
        barycentric_weight;
    std::vector<Fr> t_evaluations;

    // pub fn new(j: u32, k: u32) -> Self                                                          domain.rs:38-144
    static DomainConstants compute(uint32_t j, uint32_t k) {
        if (j < 1) throw std::invalid_argument("EvaluationDomain::new: j must be at least 1");
        DomainConstants d;
        d.k = k;
        d.quotient_poly_degree = (uint64_t)j - 1;                                                    // :40
        d.n = (uint64_t)1 << k;                                                                      // :43
        d.extended_k = k;
        while (((uint64_t)1 << d.extended_k) < d.n * d.quotient_poly_degree) d.extended_k += 1;      // :48-51
        if (d.extended_k > fr::S) throw std::invalid_argument("EvaluationDomain::new: assert!(extended_k <= F::S)");   // :54
        d.extended_omega = fr::root_of_unity();
        for (uint32_t i = d.extended_k; i < fr::S; i++) d.extended_omega = fr::square(d.extended_omega);   // :61-63
        d.omega = d.extended_omega;
        for (uint32_t i = k; i < d.extended_k; i++) d.omega = fr::square(d.omega);                   // :72-74
        d.g_coset = fr::zeta();                                                                      // :83
        d.g_coset_inv = fr::square(d.g_coset);                                                       // :84
        {                                                                                            // :86-108
            const Fr orig = fr::pow(fr::zeta(), d.n), step = fr::pow(d.extended_omega, d.n);
            Fr cur = orig;
            do { d.t_evaluations.push_back(cur); cur = fr::mul(cur, step); } while (cur != orig);
            if (d.t_evaluations.size() != (size_t)1 << (d.extended_k - k)) throw std::logic_error("t_evaluations length");
            for (Fr& c : d.t_evaluations) c = fr::invert(fr::sub(c, fr::one()));
        }
        d.ifft_divisor = fr::invert(fr::from_u64((uint64_t)1 << k));                                 // :110-126 (batch_invert)
        d.extended_ifft_divisor = fr::invert(fr::from_u64((uint64_t)1 << d.extended_k));
        d.barycentric_weight = fr::invert(fr::from_u64(d.n));
        d.extended_omega_inv = fr::invert(d.extended_omega);
        d.omega_inv = fr::invert(d.omega);
        return d;
    }
};

// The transforms run on the GPU; polynomials are std::vector<Fr> (the reference's Polynomial<F, B> is a Vec<F> plus a basis marker).
class EvaluationDomain {
public:
    // pub fn new(j: u32, k: u32) -> Self                                                          domain.rs:38
    EvaluationDomain(uint32_t j, uint32_t k, std::shared_ptr<Context> ctx) : ctx_(std::move(ctx)), c_(DomainConstants::compute(j, k)) {}
    const DomainConstants& constants() const { return c_; }
    uint32_t k() const { return c_.k; }
    uint32_t extended_k() const { return c_.extended_k; }
    uint64_t extended_len() const { return (uint64_t)1 << c_.extended_k; }
    Fr get_omega() const { return c_.omega; }
    Fr get_omega_inv() const { return c_.omega_inv; }
    Fr get_extended_omega() const { return c_.extended_omega; }
    uint64_t get_quotient_poly_degree() const { return c_.quotient_poly_degree; }

    // pub fn rotate_omega(&self, value: F, rotation: Rotation) -> F                               domain.rs:383-394
    Fr rotate_omega(const Fr& value, int32_t rotation) const {
        if (rotation >= 0) return fr::mul(value, fr::pow(c_.omega, (uint64_t)rotation));
        return fr::mul(value, fr::pow(c_.omega_inv, (uint64_t)(-(int64_t)rotation)));
    }

    // pub fn lagrange_to_coeff(&self, a: Polynomial<F, LagrangeCoeff>) -> Polynomial<F, Coeff>    domain.rs:216-226
    std::vector<Fr> lagrange_to_coeff(std::vector<Fr> a) const {
        if (a.size() != c_.n) throw std::invalid_argument("lagrange_to_coeff: assert_eq!(a.values.len(), 1 << self.k)");
        ctx_->check(h2b_ntt(ctx_->raw(), a.data(), c_.k, &c_.omega_inv, &c_.ifft_divisor), "h2b_ntt");
        return a;
    }
    // best_fft(a, omega, k): coefficients -> evaluations on the size-n domain
    std::vector<Fr> coeff_to_lagrange(std::vector<Fr> a) const {
        if (a.size() != c_.n) throw std::invalid_argument("coeff_to_lagrange: wrong length");
        ctx_->check(h2b_ntt(ctx_->raw(), a.data(), c_.k, &c_.omega, nullptr), "h2b_ntt");
        return a;
    }
    // pub fn coeff_to_extended(&self, a: Polynomial<F, Coeff>) -> Polynomial<F, ExtendedLagrangeCoeff>   domain.rs:230-244
    std::vector<Fr> coeff_to_extended(const std::vector<Fr>& a) const {
        if (a.size() != c_.n) throw std::invalid_argument("coeff_to_extended: assert_eq!(a.values.len(), 1 << self.k)");
        std::vector<Fr> out(extended_len());
        ctx_->check(h2b_coeff_to_extended(ctx_->raw(), a.data(), out.data(), c_.k, c_.extended_k, &c_.extended_omega, &c_.g_coset,
                                          &c_.g_coset_inv), "h2b_coeff_to_extended");
        return out;
    }
    // the independent per-column transforms of evaluate_h (plonk/evaluation.rs:344-361) in one call, dealt over the devices
    std::vector<std::vector<Fr>> coeff_to_extended_batch(const std::vector<std::vector<Fr>>& cols) const {
        std::vector<std::vector<Fr>> outs(cols.size());
        std::vector<const void*> in;
        std::vector<void*> out;
        for (size_t i = 0; i < cols.size(); i++) {
            if (cols[i].size() != c_.n) throw std::invalid_argument("coeff_to_extended_batch: wrong length");
            outs[i].resize(extended_len());
            in.push_back(cols[i].data());
            out.push_back(outs[i].data());
        }
        if (!cols.empty())
            ctx_->check(h2b_coeff_to_extended_batch(ctx_->raw(), in.data(), out.data(), cols.size(), c_.k, c_.extended_k,
                                                    &c_.extended_omega, &c_.g_coset, &c_.g_coset_inv), "h2b_coeff_to_extended_batch");
        return outs;
    }
    // pub fn extended_to_coeff(&self, a: Polynomial<F, ExtendedLagrangeCoeff>) -> Vec<F>           domain.rs:271-293
    // (truncated to n * quotient_poly_degree coefficients, :289-291).  divide_by_vanishing = true fuses
    // divide_by_vanishing_poly (domain.rs:297-316) in front, as vanishing/prover.rs:107-111 calls the two back to back.
    std::vector<Fr> extended_to_coeff(const std::vector<Fr>& a, bool divide_by_vanishing = false) const {
        if (a.size() != extended_len()) throw std::invalid_argument("extended_to_coeff: assert_eq!(a.values.len(), self.extended_len())");
        std::vector<Fr> out((size_t)(c_.n * c_.quotient_poly_degree));
        ctx_->check(h2b_extended_to_coeff(ctx_->raw(), a.data(), out.data(), c_.extended_k, &c_.extended_omega_inv,
                                          &c_.extended_ifft_divisor, &c_.g_coset, &c_.g_coset_inv, out.size(),
                                          divide_by_vanishing ? c_.t_evaluations.data() : nullptr,
                                          divide_by_vanishing ? c_.t_evaluations.size() : 0), "h2b_extended_to_coeff");
        return out;
    }
private:
    std::shared_ptr<Context> ctx_;
    DomainConstants c_;
};

}  // namespace h2b200
