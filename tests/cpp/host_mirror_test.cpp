// host_mirror_test.cpp -- the C++ host mirror (include/h2b200.hpp) against the CPU oracle (oracle/libh2ref.so).
//
// Built and run by tests/test_cpp_host.py.  `host_mirror_test host` needs no GPU: it checks the host-side Fr arithmetic and the
// EvaluationDomain constants.  `host_mirror_test gpu` drives the library through the mirror exactly as the reference's tests drive
// the Rust API and compares every result bit for bit with the oracle:
//   * MsmAccel: engine.msm == best_multiexp, and all cached variants agree           (halo2_middleware/src/zal.rs:262-349)
//   * commit(lagrange_to_coeff(a)) == commit_lagrange(a)                              (poly/kzg/commitment.rs:381-408)
//   * lagrange_to_coeff / coeff_to_extended / extended_to_coeff (+ divide_by_vanishing_poly) (poly/domain.rs:478-560 tests)
//   * the reference's panics: unequal slice lengths, polynomials longer than the SRS, wrong domain lengths
// This file is test infrastructure: it is the only C++ that links the oracle.
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>

#include "h2b200.hpp"

extern "C" {
void h2ref_best_multiexp(const void* coeffs, const void* bases, size_t n, void* out96);
void h2ref_to_affine(const void* in96, void* out64);
void h2ref_lagrange_to_coeff(void* a, uint32_t k, const void* omega_inv, const void* ifft_divisor);
void h2ref_best_fft(void* a, const void* omega, uint32_t log_n);
void h2ref_coeff_to_extended(const void* in_n, void* out_ne, uint32_t k, uint32_t ext_k, const void* ext_omega, const void* g_coset,
                             const void* g_coset_inv);
void h2ref_extended_to_coeff(void* a_ne, uint32_t ext_k, const void* ext_omega_inv, const void* ext_ifft_divisor, const void* g_coset,
                             const void* g_coset_inv);
void h2ref_divide_by_vanishing_poly(void* a, size_t len, const void* t_evals, size_t e);
void h2ref_fr_mul(const void* a, const void* b, void* o);
void h2ref_fr_add(const void* a, const void* b, void* o);
void h2ref_fr_sub(const void* a, const void* b, void* o);
void h2ref_fr_inv(const void* a, void* o);
void h2ref_random_fr(void* out, size_t n, uint64_t seed);
void h2ref_gen_bases(void* out, size_t n, uint64_t t0);
void h2ref_kzg_setup(uint32_t k, const void* s, const void* omega, void* g_out, void* g_lagrange_out);
}

using namespace h2b200;

static int failures = 0;
static void expect(bool ok, const std::string& what) {
    std::printf("%s %s\n", ok ? "ok  " : "FAIL", what.c_str());
    if (!ok) failures++;
}
template <class E>
static void expect_throws(const std::function<void()>& f, const std::string& what) {
    bool thrown = false;
    try { f(); } catch (const E&) { thrown = true; } catch (...) {}
    expect(thrown, what);
}
static std::vector<Fr> random_fr(size_t n, uint64_t seed) {
    std::vector<Fr> v(n);
    if (n) h2ref_random_fr(v.data(), n, seed);
    return v;
}
static std::vector<G1Affine> gen_bases(size_t n, uint64_t t0) {
    std::vector<G1Affine> v(n);
    if (n) h2ref_gen_bases(v.data(), n, t0);
    return v;
}
// the oracle's best_multiexp, normalised the way the library returns points
static G1 oracle_msm(const std::vector<Fr>& s, const std::vector<G1Affine>& b, size_t n) {
    G1 jac;
    G1Affine aff;
    h2ref_best_multiexp(s.data(), b.data(), n, &jac);
    h2ref_to_affine(&jac, &aff);
    G1 out;
    const Fq zero = {{0, 0, 0, 0}};
    Fq one_m;                                                  // Montgomery one of Fq = 2^256 mod q
    const uint64_t r_q[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull, 0x0e0a77c19a07df2full};
    std::memcpy(&one_m, r_q, 32);
    const bool identity = std::memcmp(&aff.x, &zero, 32) == 0 && std::memcmp(&aff.y, &zero, 32) == 0;
    if (identity) { out.x = zero; out.y = one_m; out.z = zero; }
    else { out.x = aff.x; out.y = aff.y; out.z = one_m; }
    return out;
}

static void test_host() {
    // Fr arithmetic against the oracle's
    std::vector<Fr> a = random_fr(64, 11), b = random_fr(64, 12);
    bool ok_mul = true, ok_add = true, ok_sub = true, ok_inv = true;
    for (size_t i = 0; i < a.size(); i++) {
        Fr e;
        h2ref_fr_mul(&a[i], &b[i], &e); ok_mul &= fr::mul(a[i], b[i]) == e;
        h2ref_fr_add(&a[i], &b[i], &e); ok_add &= fr::add(a[i], b[i]) == e;
        h2ref_fr_sub(&a[i], &b[i], &e); ok_sub &= fr::sub(a[i], b[i]) == e;
        h2ref_fr_inv(&a[i], &e); ok_inv &= fr::invert(a[i]) == e;
    }
    expect(ok_mul, "fr::mul == oracle");
    expect(ok_add, "fr::add == oracle");
    expect(ok_sub, "fr::sub == oracle");
    expect(ok_inv, "fr::invert == oracle");
    expect(fr::mul(fr::one(), a[0]) == a[0] && fr::add(a[0], fr::neg(a[0])) == fr::zero(), "one and neg");
    // ROOT_OF_UNITY has order exactly 2^28, ZETA order 3
    Fr w = fr::root_of_unity();
    for (uint32_t i = 0; i + 1 < fr::S; i++) w = fr::square(w);
    expect(w != fr::one() && fr::square(w) == fr::one(), "ROOT_OF_UNITY has order 2^28");
    const Fr z = fr::zeta();
    expect(z != fr::one() && fr::mul(fr::square(z), z) == fr::one(), "ZETA^3 == 1");
    // EvaluationDomain::new (domain.rs:38-144): the identities its fields satisfy, for the (j, k) pairs the tests use
    const uint32_t jk[][2] = {{2, 4}, {3, 5}, {5, 8}, {6, 10}, {9, 3}, {5, 20}, {6, 20}, {5, 24}};
    for (const auto& p : jk) {
        DomainConstants d = DomainConstants::compute(p[0], p[1]);
        const std::string tag = "domain(j=" + std::to_string(p[0]) + ", k=" + std::to_string(p[1]) + ")";
        bool ok = ((uint64_t)1 << d.extended_k) >= d.n * d.quotient_poly_degree &&
                  (d.extended_k == d.k || ((uint64_t)1 << (d.extended_k - 1)) < d.n * d.quotient_poly_degree);
        ok &= fr::pow(d.omega, d.n) == fr::one() && fr::pow(d.omega, d.n / 2) != fr::one();
        ok &= fr::pow(d.extended_omega, (uint64_t)1 << d.extended_k) == fr::one();
        ok &= fr::pow(d.extended_omega, (uint64_t)1 << (d.extended_k - d.k)) == d.omega;
        ok &= fr::mul(d.omega, d.omega_inv) == fr::one() && fr::mul(d.extended_omega, d.extended_omega_inv) == fr::one();
        ok &= fr::mul(d.g_coset, d.g_coset_inv) == fr::one();
        ok &= fr::mul(d.ifft_divisor, fr::from_u64(d.n)) == fr::one() && d.barycentric_weight == d.ifft_divisor;
        ok &= fr::mul(d.extended_ifft_divisor, fr::from_u64((uint64_t)1 << d.extended_k)) == fr::one();
        ok &= d.t_evaluations.size() == (size_t)1 << (d.extended_k - d.k);
        // t_evaluations[i] = 1 / ((zeta * extended_omega^i)^n - 1)
        for (size_t i = 0; i < d.t_evaluations.size() && i < 8; i++) {
            Fr x = fr::mul(d.g_coset, fr::pow(d.extended_omega, (uint64_t)i));
            ok &= fr::mul(d.t_evaluations[i], fr::sub(fr::pow(x, d.n), fr::one())) == fr::one();
        }
        expect(ok, tag);
    }
    expect_throws<std::invalid_argument>([] { DomainConstants::compute(9, 27); }, "extended_k > S is refused (domain.rs:54)");
}

static void test_gpu() {
    auto ctx = std::make_shared<Context>();
    B200Engine engine(ctx);

    // ---- MsmAccel ----------------------------------------------------------------------------------------------------------
    for (size_t n : {(size_t)0, (size_t)1, (size_t)2, (size_t)33, (size_t)1000, (size_t)(1 << 14) + 5}) {
        std::vector<Fr> s = random_fr(n, 100 + n);
        std::vector<G1Affine> b = gen_bases(n, 7);
        if (n > 3) { s[1] = fr::zero(); s[2] = fr::one(); s[3] = fr::neg(fr::one()); }
        if (n > 40) { std::memset(&b[5], 0, 64); b[7] = b[6]; }                   // an identity base, a repeated base
        G1 want = oracle_msm(s, b, n);
        expect(engine.msm(s, b) == want, "engine.msm == best_multiexp, n = " + std::to_string(n));
        if (n == 0) continue;
        BaseDescriptor bd = engine.get_base_descriptor(b);
        CoeffsDescriptor cd = engine.get_coeffs_descriptor(s);
        bool same = engine.msm_with_cached_base(s, bd) == want && engine.msm_with_cached_scalars(cd, b) == want &&
                    engine.msm_with_cached_inputs(cd, bd) == want;
        expect(same, "msm_with_cached_{base, scalars, inputs} agree, n = " + std::to_string(n));
        if (n > 10) {
            const size_t m = n / 3;
            expect(engine.msm_with_cached_base(Slice<Fr>(s).first(m), bd) == oracle_msm(s, b, m), "prefix of a descriptor, n = " + std::to_string(m));
        }
    }
    {
        std::vector<Fr> s = random_fr(10, 1);
        std::vector<G1Affine> b = gen_bases(9, 1);
        expect_throws<std::invalid_argument>([&] { engine.msm(s, b); }, "msm: unequal lengths are refused");
        BaseDescriptor bd = engine.get_base_descriptor(b);
        expect_throws<std::invalid_argument>([&] { engine.msm_with_cached_base(s, bd); }, "msm_with_cached_base: more coefficients than bases is refused");
    }
    {   // one pipeline for the per-column commits of a phase == the same commits one by one
        const size_t n = 1 << 12;
        std::vector<G1Affine> b = gen_bases(n, 3);
        BaseDescriptor bd = engine.get_base_descriptor(b);
        std::vector<std::vector<Fr>> cols;
        for (int c = 0; c < 3; c++) cols.push_back(random_fr(n, 40 + c));
        std::vector<G1> got = engine.msm_batch({cols[0], cols[1], cols[2]}, bd);
        bool ok = true;
        for (int c = 0; c < 3; c++) ok &= got[c] == oracle_msm(cols[c], b, n);
        expect(ok, "msm_batch == best_multiexp per column");
    }

    // ---- ParamsKZG: commit(lagrange_to_coeff(a)) == commit_lagrange(a) (commitment.rs:381-408) --------------------------------
    {
        const uint32_t k = 9;
        const size_t n = (size_t)1 << k;
        EvaluationDomain domain(1, k, ctx);
        const Fr secret = fr::from_u64(0x1234567);
        const Fr omega = domain.get_omega();
        std::vector<G1Affine> g(n), gl(n);
        h2ref_kzg_setup(k, &secret, &omega, g.data(), gl.data());
        ParamsKZG params(k, g, gl);
        std::vector<Fr> a = random_fr(n, 77);
        G1 via_lagrange = params.commit_lagrange(engine, a);
        G1 via_coeff = params.commit(engine, domain.lagrange_to_coeff(a));
        expect(via_lagrange == via_coeff, "commit(lagrange_to_coeff(a)) == commit_lagrange(a)");
        expect(via_lagrange == oracle_msm(a, gl, n), "commit_lagrange == best_multiexp over g_lagrange");
        std::vector<Fr> shorter(a.begin(), a.begin() + 100);
        expect(params.commit(engine, shorter) == oracle_msm(shorter, g, 100), "commit of a shorter polynomial uses bases[0..size)");
        std::vector<Fr> longer = random_fr(n + 1, 5);
        expect_throws<std::invalid_argument>([&] { params.commit(engine, longer); }, "commit: polynomial longer than the SRS is refused");
    }

    // ---- EvaluationDomain ------------------------------------------------------------------------------------------------------
    const uint32_t jk[][2] = {{2, 4}, {3, 5}, {5, 8}, {6, 10}, {9, 3}, {5, 13}};
    for (const auto& p : jk) {
        EvaluationDomain domain(p[0], p[1], ctx);
        const DomainConstants& d = domain.constants();
        const std::string tag = "(j=" + std::to_string(p[0]) + ", k=" + std::to_string(p[1]) + ")";
        std::vector<Fr> a = random_fr(d.n, 900 + p[1]);
        std::vector<Fr> want = a;
        h2ref_lagrange_to_coeff(want.data(), d.k, &d.omega_inv, &d.ifft_divisor);
        std::vector<Fr> coeff = domain.lagrange_to_coeff(a);
        expect(coeff == want, "lagrange_to_coeff == oracle " + tag);
        expect(domain.coeff_to_lagrange(coeff) == a, "coeff_to_lagrange inverts it " + tag);
        std::vector<Fr> ext_want(domain.extended_len());
        h2ref_coeff_to_extended(coeff.data(), ext_want.data(), d.k, d.extended_k, &d.extended_omega, &d.g_coset, &d.g_coset_inv);
        std::vector<Fr> ext = domain.coeff_to_extended(coeff);
        expect(ext == ext_want, "coeff_to_extended == oracle " + tag);
        std::vector<std::vector<Fr>> many = domain.coeff_to_extended_batch({coeff, a});
        expect(many.size() == 2 && many[0] == ext_want, "coeff_to_extended_batch == single " + tag);
        // extended_to_coeff on arbitrary extended values, plain and with the vanishing division fused
        std::vector<Fr> e = random_fr(domain.extended_len(), 300 + p[0]);
        for (int divide = 0; divide < 2; divide++) {
            std::vector<Fr> back = e;
            if (divide) h2ref_divide_by_vanishing_poly(back.data(), back.size(), d.t_evaluations.data(), d.t_evaluations.size());
            h2ref_extended_to_coeff(back.data(), d.extended_k, &d.extended_omega_inv, &d.extended_ifft_divisor, &d.g_coset, &d.g_coset_inv);
            back.resize((size_t)(d.n * d.quotient_poly_degree));
            expect(domain.extended_to_coeff(e, divide != 0) == back, std::string("extended_to_coeff") + (divide ? " after divide_by_vanishing_poly" : "") + " == oracle " + tag);
        }
        // round trip: a polynomial of degree < n survives coeff_to_extended -> extended_to_coeff (quotient degree >= 1)
        if (d.quotient_poly_degree >= 1) {
            std::vector<Fr> rt = domain.extended_to_coeff(ext);
            bool ok = rt.size() == d.n * d.quotient_poly_degree;
            for (size_t i = 0; ok && i < rt.size(); i++) ok = rt[i] == (i < d.n ? coeff[i] : fr::zero());
            expect(ok, "extended_to_coeff(coeff_to_extended(p)) == p " + tag);
        }
        std::vector<Fr> wrong(d.n + 1);
        expect_throws<std::invalid_argument>([&] { domain.lagrange_to_coeff(wrong); }, "lagrange_to_coeff: wrong length is refused " + tag);
        expect_throws<std::invalid_argument>([&] { domain.coeff_to_extended(wrong); }, "coeff_to_extended: wrong length is refused " + tag);
        expect_throws<std::invalid_argument>([&] { domain.extended_to_coeff(wrong); }, "extended_to_coeff: wrong length is refused " + tag);
        // rotate_omega (domain.rs:383-394)
        const Fr x = a[0];
        expect(domain.rotate_omega(domain.rotate_omega(x, 3), -3) == x && domain.rotate_omega(x, 1) == fr::mul(x, d.omega), "rotate_omega " + tag);
    }
    std::printf("launches %llu\n", (unsigned long long)ctx->launch_count());
    expect(ctx->launch_count() > 0, "the CUDA library did the work");
}

int main(int argc, char** argv) {
    const std::string mode = argc > 1 ? argv[1] : "host";
    try {
        if (mode == "host") test_host();
        else if (mode == "gpu") { test_host(); test_gpu(); }
        else if (mode == "consts" && argc == 4) {
            // the fields of EvaluationDomain::new(j, k) as canonical big-endian hex, for the python-int oracle to compare
            DomainConstants d = DomainConstants::compute((uint32_t)std::atoi(argv[2]), (uint32_t)std::atoi(argv[3]));
            auto show = [](const char* name, const Fr& v) {
                auto c = fr::to_canonical(v);
                std::printf("%s 0x%016llx%016llx%016llx%016llx\n", name, (unsigned long long)c[3], (unsigned long long)c[2],
                            (unsigned long long)c[1], (unsigned long long)c[0]);
            };
            std::printf("extended_k %u\nquotient_poly_degree %llu\n", d.extended_k, (unsigned long long)d.quotient_poly_degree);
            show("omega", d.omega); show("omega_inv", d.omega_inv); show("extended_omega", d.extended_omega);
            show("extended_omega_inv", d.extended_omega_inv); show("g_coset", d.g_coset); show("g_coset_inv", d.g_coset_inv);
            show("ifft_divisor", d.ifft_divisor); show("extended_ifft_divisor", d.extended_ifft_divisor);
            show("barycentric_weight", d.barycentric_weight);
            for (const Fr& t : d.t_evaluations) show("t_evaluation", t);
            return 0;
        } else if (mode == "nodevice") {
            // without a GPU the product path must fail loudly (no CPU fallback)
            expect_throws<Error>([] { Context c; }, "Context without a device throws h2b200::Error");
        } else { std::fprintf(stderr, "usage: host_mirror_test host | gpu | nodevice | consts j k\n"); return 2; }
    } catch (const std::exception& e) {
        std::printf("FAIL uncaught exception: %s\n", e.what());
        failures++;
    }
    std::printf("%s (%d failures)\n", failures ? "FAILED" : "PASSED", failures);
    return failures ? 1 : 0;
}
