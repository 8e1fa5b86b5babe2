// curve.cuh -- BN254 G1 (y^2 = x^3 + 3 over Fq) point arithmetic for the bucket MSM.
//
// Bases arrive in the reference's G1Affine layout: (x, y) Montgomery, 64 B, identity = (0,0)
// (SURVEY.md §A.2).  Accumulators use extended Jacobian "XYZZ" coordinates (x = X/ZZ, y = Y/ZZZ,
// ZZ^3 = ZZZ^2, identity ZZ = 0): the mixed addition is 8M + 2S with no inversion and no Z.
// Formulas: EFD shortw/xyzz madd-2008-s, add-2008-s, dbl-2008-s-1, mdbl-2008-s-1 (a = 0).
#pragma once
#include "field.cuh"
#include "field_wide.cuh"

namespace h2b {

struct alignas(16) Affine {
    Fq x, y;
    H2B_HD bool is_identity() const { return x.is_zero() && y.is_zero(); }
};

struct alignas(16) Xyzz {
    Fq x, y, zz, zzz;
    H2B_HD static Xyzz identity() {
        Xyzz r;
        r.x = Fq::zero(); r.y = Fq::zero(); r.zz = Fq::zero(); r.zzz = Fq::zero();
        return r;
    }
    H2B_HD bool is_identity() const { return zz.is_zero(); }
    H2B_HD static Xyzz from_affine(const Affine& p) {
        Xyzz r;
        if (p.is_identity()) return identity();
        r.x = p.x; r.y = p.y; r.zz = Fq::one(); r.zzz = Fq::one();
        return r;
    }
};

H2B_HD Affine affine_load(const Affine* p) {
    Affine r;
    r.x = fe_load(&p->x);
    r.y = fe_load(&p->y);
    return r;
}
H2B_HD void xyzz_store(Xyzz* p, const Xyzz& v) {
    fe_store(&p->x, v.x); fe_store(&p->y, v.y); fe_store(&p->zz, v.zz); fe_store(&p->zzz, v.zzz);
}
H2B_HD Xyzz xyzz_load(const Xyzz* p) {
    Xyzz r;
    r.x = fe_load(&p->x); r.y = fe_load(&p->y); r.zz = fe_load(&p->zz); r.zzz = fe_load(&p->zzz);
    return r;
}

// 2*(x,y) for an affine point, result XYZZ   (mdbl-2008-s-1, a = 0)
H2B_HD Xyzz xyzz_double_affine(const Fq& x1, const Fq& y1) {
    Xyzz r;
    Fq u = fe_dbl(y1);
    Fq v = fe_sqr_w(u);
    Fq w = fe_mul(u, v);
    Fq s = fe_mul(x1, v);
    Fq xx = fe_sqr_w(x1);
    Fq m = fe_add(fe_dbl(xx), xx);
    r.x = fe_sub(fe_sqr_w(m), fe_dbl(s));
    r.y = fe_mul_sub_mul(m, fe_sub(s, r.x), w, y1);
    r.zz = v;
    r.zzz = w;
    return r;
}

// 2*P, P in XYZZ   (dbl-2008-s-1, a = 0).  A point with y = 0 does not exist on this curve (order is odd).
H2B_HD Xyzz xyzz_double(const Xyzz& p) {
    if (p.is_identity()) return p;
    Xyzz r;
    Fq u = fe_dbl(p.y);
    Fq v = fe_sqr_w(u);
    Fq w = fe_mul(u, v);
    Fq s = fe_mul(p.x, v);
    Fq xx = fe_sqr_w(p.x);
    Fq m = fe_add(fe_dbl(xx), xx);
    r.x = fe_sub(fe_sqr_w(m), fe_dbl(s));
    r.y = fe_mul_sub_mul(m, fe_sub(s, r.x), w, p.y);
    r.zz = fe_mul(v, p.zz);
    r.zzz = fe_mul(w, p.zzz);
    return r;
}

// acc += (x2, y2) affine, never the identity; `neg` negates the affine point first   (madd-2008-s)
H2B_HD void xyzz_madd(Xyzz& acc, const Fq& x2, const Fq& y2_in, bool neg) {
    Fq y2 = neg ? fe_neg(y2_in) : y2_in;
    if (acc.is_identity()) {
        acc.x = x2; acc.y = y2; acc.zz = Fq::one(); acc.zzz = Fq::one();
        return;
    }
    Fq u2 = fe_mul(x2, acc.zz);
    Fq s2 = fe_mul(y2, acc.zzz);
    Fq p = fe_sub(u2, acc.x);
    Fq r = fe_sub(s2, acc.y);
    if (p.is_zero()) {
        if (r.is_zero()) acc = xyzz_double_affine(x2, y2);
        else acc = Xyzz::identity();
        return;
    }
    Fq pp = fe_sqr_w(p);
    Fq ppp = fe_mul(p, pp);
    Fq q = fe_mul(acc.x, pp);
    Fq x3 = fe_sub(fe_sub(fe_sqr_w(r), ppp), fe_dbl(q));
    Fq y3 = fe_mul_sub_mul(r, fe_sub(q, x3), acc.y, ppp);     // one reduction for r*(q-x3) - y1*ppp
    acc.x = x3;
    acc.y = y3;
    acc.zz = fe_mul(acc.zz, pp);
    acc.zzz = fe_mul(acc.zzz, ppp);
}

// a + b, both XYZZ   (add-2008-s)
H2B_HD Xyzz xyzz_add(const Xyzz& a, const Xyzz& b) {
    if (a.is_identity()) return b;
    if (b.is_identity()) return a;
    Fq u1 = fe_mul(a.x, b.zz);
    Fq u2 = fe_mul(b.x, a.zz);
    Fq s1 = fe_mul(a.y, b.zzz);
    Fq s2 = fe_mul(b.y, a.zzz);
    Fq p = fe_sub(u2, u1);
    Fq r = fe_sub(s2, s1);
    if (p.is_zero()) {
        if (r.is_zero()) return xyzz_double(a);
        return Xyzz::identity();
    }
    Xyzz o;
    Fq pp = fe_sqr_w(p);
    Fq ppp = fe_mul(p, pp);
    Fq q = fe_mul(u1, pp);
    o.x = fe_sub(fe_sub(fe_sqr_w(r), ppp), fe_dbl(q));
    o.y = fe_mul_sub_mul(r, fe_sub(q, o.x), s1, ppp);
    o.zz = fe_mul(fe_mul(a.zz, b.zz), pp);
    o.zzz = fe_mul(fe_mul(a.zzz, b.zzz), ppp);
    return o;
}

H2B_HD Xyzz xyzz_neg(const Xyzz& a) {
    Xyzz r = a;
    r.y = fe_neg(a.y);
    return r;
}

// XYZZ -> affine (one inversion); identity -> (0,0)
H2B_HD Affine xyzz_to_affine(const Xyzz& p) {
    Affine r;
    if (p.is_identity()) { r.x = Fq::zero(); r.y = Fq::zero(); return r; }
    // 1/ZZ and 1/ZZZ from one inversion: i = 1/(ZZ*ZZZ); 1/ZZ = i*ZZZ; 1/ZZZ = i*ZZ
    Fq i = fe_inv(fe_mul(p.zz, p.zzz));
    r.x = fe_mul(p.x, fe_mul(i, p.zzz));
    r.y = fe_mul(p.y, fe_mul(i, p.zz));
    return r;
}

}  // namespace h2b
