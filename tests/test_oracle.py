"""CPU tests of the oracle itself: constants, the public EIP-196 vector, python-vs-C agreement, the reference's
algebraic test identities, and the committed golden vectors."""
import json
import os

import numpy as np
import pytest

from oracle import cref as C
from oracle import pyref as R

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "vectors.json")))


def m(x):
    return C.mont_from_int(x, R.R_MOD)


def test_constants():
    # SURVEY.md §A.1
    assert R.R_MOD.bit_length() == 254 and R.Q_MOD.bit_length() == 254
    assert (R.R_MOD - 1) % (1 << 28) == 0 and ((R.R_MOD - 1) >> 28) % 2 == 1
    assert R.ROOT_OF_UNITY == 0x03DDB9F5166D18B798865EA93DD31F743215CF6DD39329C8D34F1ED960C37C9C
    assert pow(R.ROOT_OF_UNITY, 1 << 28, R.R_MOD) == 1 and pow(R.ROOT_OF_UNITY, 1 << 27, R.R_MOD) != 1
    assert pow(R.FR_ZETA, 3, R.R_MOD) == 1 and R.FR_ZETA != 1
    assert pow(7, 1 << 28, R.R_MOD) == 0x09226B6E22C6F0CA64EC26AAD4C86E715B5F898E5E963F25870E56BBE533E9A2   # DELTA
    assert (-pow(R.R_MOD, -1, 1 << 64)) % (1 << 64) == 0xC2E1F593EFFFFFFF
    assert (-pow(R.Q_MOD, -1, 1 << 64)) % (1 << 64) == 0x87D20782E4866389
    assert R.g1_is_on_curve(R.G1_GEN)


def test_eip196_known_answer():
    """alt_bn128 2*G from the EIP-196 test vectors: the one public BN254 known answer we can anchor on."""
    g2 = R.g1_mul(R.G1_GEN, 2)
    assert g2 == (0x030644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD3,
                  0x15ED738C0E0A7C92E7845F96B2AE9C0A68A6A449E3538FC7FF3EBF7A5A18A2C4)
    assert R.g1_add(R.G1_GEN, R.G1_GEN) == g2
    assert R.g1_mul(R.G1_GEN, R.R_MOD) is None and R.g1_mul(R.G1_GEN, R.R_MOD - 1) == R.g1_neg(R.G1_GEN)
    got = C.affine_to_ints(C.affine_mul(C.affine_from_ints(R.G1_GEN, R.Q_MOD), m(2)), R.Q_MOD)
    assert got == g2


def test_golden_field_mul():
    for a, b, c in GOLD["fr_mul"]:
        a, b, c = int(a, 16), int(b, 16), int(c, 16)
        assert C.int_from_mont(C.fr_mul(m(a), m(b)), R.R_MOD) == c
    for a, b, c in GOLD["fq_mul"]:
        a, b, c = int(a, 16), int(b, 16), int(c, 16)
        assert C.int_from_mont(C.fq_mul(C.mont_from_int(a, R.Q_MOD), C.mont_from_int(b, R.Q_MOD)), R.Q_MOD) == c


def test_golden_ntt():
    for v in GOLD["ntt"]:
        k = v["k"]
        a = C.random_fr(1 << k, v["seed"])
        out = C.best_fft(a, m(int(v["omega"], 16)), k)
        assert [hex(x) for x in C.ints_from_fr_array(out, R.R_MOD)] == v["out"]


def test_golden_domain():
    g = GOLD["domain_5_4"]
    d = R.EvaluationDomain(5, 4)
    assert d.extended_k == g["extended_k"] and hex(d.omega) == g["omega"] and hex(d.extended_omega) == g["extended_omega"]
    assert [hex(x) for x in d.t_evaluations] == g["t_evaluations"]
    a = C.random_fr(16, g["seed"])
    coeff = C.lagrange_to_coeff(a, 4, m(d.omega_inv), m(d.ifft_divisor))
    assert [hex(x) for x in C.ints_from_fr_array(coeff, R.R_MOD)] == g["coeff"]
    ext = C.coeff_to_extended(coeff, 4, d.extended_k, m(d.extended_omega), m(d.g_coset), m(d.g_coset_inv))
    assert [hex(x) for x in C.ints_from_fr_array(ext, R.R_MOD)] == g["extended"]
    back = C.extended_to_coeff(ext, d.extended_k, m(d.extended_omega_inv), m(d.extended_ifft_divisor), m(d.g_coset),
                               m(d.g_coset_inv), 16 * 4)
    assert [hex(x) for x in C.ints_from_fr_array(back, R.R_MOD)] == g["back"]


def test_golden_msm():
    for v in GOLD["msm"]:
        n = v["n"]
        s = C.random_fr(n, v["seed"])
        b = C.gen_bases(n, v["t0"])
        for fn in (C.best_multiexp, C.msm_naive):
            got = C.affine_to_ints(C.to_affine(fn(s, b)), R.Q_MOD)
            assert (hex(got[0]), hex(got[1])) == (v["x"], v["y"])
    for k, (x, y) in GOLD["g1_multiples"].items():
        assert R.g1_mul(R.G1_GEN, int(k)) == (int(x, 16), int(y, 16))


def test_random_streams_agree():
    assert C.ints_from_fr_array(C.random_fr(64, 12345), R.R_MOD) == R.random_fr(64, 12345)
    b = C.gen_bases(70, 9)
    for i in (0, 1, 33, 69):
        assert C.affine_to_ints(b[i], R.Q_MOD) == R.g1_mul(R.G1_GEN, 9 + i)
        assert C.is_on_curve(b[i])


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 31, 32, 33, 257, 3000])
def test_best_multiexp_equals_naive(n):
    """zal.rs:345-349 (engine.msm == best_multiexp) restated on the oracle: Booth Pippenger == double-and-add."""
    s = C.random_fr(n, 50 + n)
    b = C.gen_bases(n, 2)
    if n > 8:
        s[1] = 0
        s[2] = m(R.R_MOD - 1)
        b[3] = 0
        b[5] = b[4]
    assert np.array_equal(C.to_affine(C.best_multiexp(s, b)), C.to_affine(C.msm_naive(s, b)))


@pytest.mark.parametrize("k", range(0, 12))
def test_best_fft_c_vs_python(k):
    a = C.random_fr(1 << k, k)
    w = pow(R.ROOT_OF_UNITY, 1 << (28 - k), R.R_MOD)
    out = C.best_fft(a, m(w), k)
    ai = C.ints_from_fr_array(a, R.R_MOD)
    assert C.ints_from_fr_array(out, R.R_MOD) == R.best_fft(ai, w, k)
    if k <= 6:
        assert R.best_fft(ai, w, k) == R.ntt_naive(ai, w)


def test_best_fft_parallel_path_and_round_trip():
    k = 16      # log_n > log2(threads): recursive parallel butterflies
    a = C.random_fr(1 << k, 3)
    w = pow(R.ROOT_OF_UNITY, 1 << (28 - k), R.R_MOD)
    f = C.best_fft(a, m(w), k)
    back = C.lagrange_to_coeff(f, k, m(pow(w, -1, R.R_MOD)), m(pow(1 << k, -1, R.R_MOD)))
    assert np.array_equal(back, a)
    # X[0] = sum of inputs, X[n/2] = alternating sum
    ai = C.ints_from_fr_array(a, R.R_MOD)
    assert C.int_from_mont(f[0], R.R_MOD) == sum(ai) % R.R_MOD
    assert C.int_from_mont(f[1 << (k - 1)], R.R_MOD) == (sum(ai[0::2]) - sum(ai[1::2])) % R.R_MOD


def test_commit_lagrange_identity():
    """kzg/commitment.rs:381-408 restated: commit(lagrange_to_coeff(a)) == commit_lagrange(a) == [p(s)]G."""
    k, s = 5, 0xDEADBEEFCAFEBABE1234
    dom = R.EvaluationDomain(1, k)
    g, gl = C.kzg_setup(k, m(s), m(dom.omega))
    py = R.ParamsKZG(3, s)
    g3, gl3 = C.kzg_setup(3, m(s), m(R.EvaluationDomain(1, 3).omega))
    assert [C.affine_to_ints(x, R.Q_MOD) for x in g3] == py.g and [C.affine_to_ints(x, R.Q_MOD) for x in gl3] == py.g_lagrange
    a = C.fr_array_from_ints(list(range(1 << k)), R.R_MOD)
    coeff = C.lagrange_to_coeff(a, k, m(dom.omega_inv), m(dom.ifft_divisor))
    c1 = C.to_affine(C.best_multiexp(coeff, g))
    c2 = C.to_affine(C.best_multiexp(a, gl))
    assert np.array_equal(c1, c2)
    p_at_s = R.eval_polynomial(C.ints_from_fr_array(coeff, R.R_MOD), s)
    assert C.affine_to_ints(c1, R.Q_MOD) == R.g1_mul(R.G1_GEN, p_at_s)


def test_rotation_orientation():
    """domain.rs:478-517 test_rotate restated on the oracle."""
    k = 3
    d = R.EvaluationDomain(1, k)
    vals = R.random_fr(1 << k, 5)
    poly = d.lagrange_to_coeff(vals)
    poly_rot = d.lagrange_to_coeff(vals[1:] + vals[:1])
    x = 0x9876543
    assert R.eval_polynomial(poly, d.rotate_omega(x, 1)) == R.eval_polynomial(poly_rot, x)
    poly_rot_back = d.lagrange_to_coeff(vals[-1:] + vals[:-1])
    assert R.eval_polynomial(poly, d.rotate_omega(x, -1)) == R.eval_polynomial(poly_rot_back, x)


@pytest.mark.parametrize("j,k", [(2, 3), (5, 3), (3, 5), (9, 4), (5, 8)])
def test_coset_extension_is_evaluation(j, k):
    """coeff_to_extended output[i] == p(zeta * ext_omega^i); extended_to_coeff inverts it (domain.rs:230-293)."""
    d = R.EvaluationDomain(j, k)
    coeff = R.random_fr(1 << k, j * 100 + k)
    ext = C.ints_from_fr_array(C.coeff_to_extended(C.fr_array_from_ints(coeff, R.R_MOD), k, d.extended_k, m(d.extended_omega),
                                                   m(d.g_coset), m(d.g_coset_inv)), R.R_MOD)
    assert ext == d.coeff_to_extended(coeff)
    for i in (0, 1, 7, (1 << d.extended_k) - 1):
        assert ext[i] == R.eval_polynomial(coeff, R.FR_ZETA * pow(d.extended_omega, i, R.R_MOD) % R.R_MOD)
    back = d.extended_to_coeff(ext)
    assert len(back) == (1 << k) * (j - 1)
    keep = min(len(back), 1 << k)
    assert back[:keep] == coeff[:keep] and all(x == 0 for x in back[1 << k:])
    # t_evaluations are 1 / (x^n - 1) on the coset
    for i in range(len(d.t_evaluations)):
        x = R.FR_ZETA * pow(d.extended_omega, i, R.R_MOD) % R.R_MOD
        assert d.t_evaluations[i] * (pow(x, 1 << k, R.R_MOD) - 1) % R.R_MOD == 1


def test_poly_helpers_c_port_matches_python():
    """SURVEY.md 8f N3 helpers: the C restatement (arithmetic.rs:56-120, ff::BatchInvert, the Z running products) against
    python big ints, including the reference's own identity for kate_division: (X - b) q + a(b) = a."""
    for n in (1, 2, 5, 64, 257):
        a = C.random_fr(n, 900 + n)
        ai = C.ints_from_fr_array(a, R.R_MOD)
        for b in (0, 1, 5, R.R_MOD - 1, 0xABCDEF0123456789):
            bm = C.mont_from_int(b, R.R_MOD)
            assert C.int_from_mont(C.eval_polynomial(a, bm), R.R_MOD) == R.eval_polynomial(ai, b)
            q = C.ints_from_fr_array(C.kate_division(a, bm), R.R_MOD)
            assert q == R.kate_division(ai, b)
            rec = [0] * n
            for i, c in enumerate(q):
                rec[i + 1] = (rec[i + 1] + c) % R.R_MOD
                rec[i] = (rec[i] - b * c) % R.R_MOD
            rec[0] = (rec[0] + R.eval_polynomial(ai, b)) % R.R_MOD
            assert rec == ai
        z = a.copy()
        z[::3] = 0
        assert C.ints_from_fr_array(C.batch_invert(z), R.R_MOD) == R.batch_invert(C.ints_from_fr_array(z, R.R_MOD))
        for init, out_len in ((1, n), (7, n + 1), (3, max(n - 1, 0))):
            got = C.ints_from_fr_array(C.grand_product(a, C.mont_from_int(init, R.R_MOD), out_len), R.R_MOD) if out_len else []
            assert got == R.grand_product(ai, init, out_len)
