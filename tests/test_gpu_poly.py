"""GPU parity of the polynomial helpers (eval_polynomial, kate_division, batch inversion, grand products)
through the C ABI / PolyOps mirror against the CPU oracle.  Bit-exact: every output limb must be equal."""
import numpy as np
import pytest

from oracle import cref as C
from oracle import pyref as R

pytestmark = pytest.mark.gpu

SIZES = [1, 2, 7, 8, 9, 255, 256, 257, 2047, 2048, 2049, 4096 + 5, (1 << 16) + 3, 1 << 19, (1 << 19) + 4097]


def m(x):
    return C.mont_from_int(x, R.R_MOD)


@pytest.fixture(scope="module")
def ops(ctx):
    import h2b200
    return h2b200.PolyOps(ctx)


@pytest.mark.parametrize("n", SIZES)
def test_eval_polynomial(ops, n):
    a = C.random_fr(n, 10 + n)
    for pt in (0, 1, R.R_MOD - 1, 0x1234567890ABCDEF1234567890ABCDEF % R.R_MOD):
        exp = C.int_from_mont(C.eval_polynomial(a, m(pt)), R.R_MOD)
        assert ops.eval_polynomial(a, pt) == exp, (n, pt)


def test_eval_polynomial_empty_and_pyref(ops):
    assert ops.eval_polynomial(np.zeros((0, 4), dtype=np.uint64), 5) == 0
    a = C.random_fr(300, 3)
    assert ops.eval_polynomial(a, 77) == R.eval_polynomial(C.ints_from_fr_array(a, R.R_MOD), 77)


@pytest.mark.parametrize("n", SIZES)
def test_kate_division(ops, n):
    a = C.random_fr(n, 20 + n)
    for b in (0, 1, 0xDEADBEEFCAFEBABE1234 % R.R_MOD):
        got = ops.kate_division(a, b)
        assert got.shape[0] == n - 1
        assert np.array_equal(got, C.kate_division(a, m(b))), (n, b)


def test_kate_division_is_exact_division(ops):
    """(X - b) * q + a(b) == a : the identity the SHPLONK / GWC openers rely on (shplonk/prover.rs, gwc/prover.rs)."""
    n, b = 1000, 987654321
    a = C.random_fr(n, 5)
    ai = C.ints_from_fr_array(a, R.R_MOD)
    q = C.ints_from_fr_array(ops.kate_division(a, b), R.R_MOD)
    rem = ops.eval_polynomial(a, b)
    rec = [0] * n
    for i, c in enumerate(q):
        rec[i + 1] = (rec[i + 1] + c) % R.R_MOD
        rec[i] = (rec[i] - b * c) % R.R_MOD
    rec[0] = (rec[0] + rem) % R.R_MOD
    assert rec == ai
    with pytest.raises(AssertionError):
        ops.kate_division(np.zeros((0, 4), dtype=np.uint64), 1)


@pytest.mark.parametrize("n", SIZES)
def test_batch_invert(ops, n):
    a = C.random_fr(n, 30 + n)
    a[:: 5] = 0                      # ff::BatchInvert leaves zeros untouched
    if n > 3:
        a[3] = m(1)
    got = ops.batch_invert(a.copy())
    assert np.array_equal(got, C.batch_invert(a)), n


def test_batch_invert_all_zero_and_involution(ops):
    z = np.zeros((5000, 4), dtype=np.uint64)
    assert not ops.batch_invert(z.copy()).any()
    a = C.random_fr(70000, 4)
    assert np.array_equal(ops.batch_invert(ops.batch_invert(a.copy())), a)


@pytest.mark.parametrize("n", SIZES)
def test_grand_product(ops, n):
    v = C.random_fr(n, 40 + n)
    for init, out_len in ((1, n), (12345, n + 1), (R.R_MOD - 1, max(n - 6, 0))):
        got = ops.grand_product(v, init, out_len)
        assert np.array_equal(got, C.grand_product(v, m(init), out_len)), (n, init, out_len)


def test_grand_product_with_zero_and_empty(ops):
    v = C.random_fr(5000, 8)
    v[2500] = 0
    got = ops.grand_product(v, 1, 5001)
    assert np.array_equal(got, C.grand_product(v, m(1), 5001))
    assert not got[2501:].any()
    assert np.array_equal(ops.grand_product(np.zeros((0, 4), dtype=np.uint64), 9, 1), m(9).reshape(1, 4))


def test_device_resident_variants(ops):
    """torch CUDA tensors in, results stay on the device (the *_device entry points)."""
    import torch
    n = (1 << 17) + 11
    a = C.random_fr(n, 77)
    d = torch.from_numpy(a.view(np.int64)).cuda()
    assert ops.eval_polynomial(d, 4242) == C.int_from_mont(C.eval_polynomial(a, m(4242)), R.R_MOD)
    q = ops.kate_division(d, 4242)
    assert np.array_equal(q.cpu().numpy().view(np.uint64), C.kate_division(a, m(4242)))
    z = ops.grand_product(d, 3, n)
    assert np.array_equal(z.cpu().numpy().view(np.uint64), C.grand_product(a, m(3), n))
    inv = ops.batch_invert(d.clone())
    assert np.array_equal(inv.cpu().numpy().view(np.uint64), C.batch_invert(a))


@pytest.mark.parametrize("n", [1, 7, 2048, 2049, (1 << 16) + 3, (1 << 19) + 4097])
def test_eval_polynomial_batch(ops, n):
    """h2b_eval_polynomial_batch_device: many (polynomial, point) pairs in one call == the oracle, pair by pair."""
    import torch
    polys = [C.random_fr(n, 900 + i) for i in range(3)]
    devs = [torch.from_numpy(a.view(np.int64)).cuda() for a in polys]
    points = [0, 1, R.R_MOD - 1, 5, 0x1234567890ABCDEF1234567890ABCDEF % R.R_MOD, 77, 78]
    pairs = [(i % 3, pt) for i, pt in enumerate(points)] + [(0, 5), (2, 5)]           # repeated polynomials and points
    got = ops.eval_polynomials([(devs[i], pt) for i, pt in pairs])
    exp = [C.int_from_mont(C.eval_polynomial(polys[i], m(pt)), R.R_MOD) for i, pt in pairs]
    assert got == exp
    assert got == [ops.eval_polynomial(devs[i], pt) for i, pt in pairs]


def test_eval_polynomial_batch_edges(ops):
    import torch
    assert ops.eval_polynomials([]) == []
    a = C.random_fr(300, 3)
    d = torch.from_numpy(a.view(np.int64)).cuda()
    # host arrays and mixed lengths fall back to single evaluations, same values
    b = C.random_fr(17, 4)
    got = ops.eval_polynomials([(d, 9), (b, 9), (a, 11)])
    assert got == [R.eval_polynomial(C.ints_from_fr_array(a, R.R_MOD), 9), R.eval_polynomial(C.ints_from_fr_array(b, R.R_MOD), 9),
                   R.eval_polynomial(C.ints_from_fr_array(a, R.R_MOD), 11)]
    many = ops.eval_polynomials([(d, 1000 + i) for i in range(130)])                  # more jobs than one wave of blocks per tile
    ints = C.ints_from_fr_array(a, R.R_MOD)
    assert many == [R.eval_polynomial(ints, 1000 + i) for i in range(130)]


def test_permutation_product_shape(ops):
    """The sequence of permutation/prover.rs:118-166 on the GPU: denominators -> batch_invert -> numerators -> running product."""
    n = 1 << 12
    num, den = C.random_fr(n, 1), C.random_fr(n, 2)
    inv = ops.batch_invert(den.copy())
    frac = np.stack([C.fr_mul(num[i], inv[i]) for i in range(n)])
    z = ops.grand_product(frac, 1, n)
    ni, di = C.ints_from_fr_array(num, R.R_MOD), C.ints_from_fr_array(den, R.R_MOD)
    exp = R.grand_product([a * pow(b, -1, R.R_MOD) % R.R_MOD for a, b in zip(ni, di)], 1, n)
    assert C.ints_from_fr_array(z, R.R_MOD) == exp


@pytest.mark.parametrize("n", [1, 33, 4096 + 5, (1 << 18) + 1])
def test_pointwise_mul_and_coset_scaling(ops, ctx, n):
    """h2b_fr_mul_device: out = a * b (* scale), bit-exact against python ints, in place and out of place; with b = the power table
    of a coset shift it is distribute_powers (domain.rs:325-341) for that shift: a size-n transform of the scaled coefficients gives
    the polynomial's values on the coset (checked against direct evaluation at a few rows)."""
    import ctypes
    import torch
    a, b = C.random_fr(n, 500 + n), C.random_fr(n, 900 + n)
    ai = [C.int_from_mont(x, R.R_MOD) for x in a[:64]]
    bi = [C.int_from_mont(x, R.R_MOD) for x in b[:64]]
    da, db = (torch.from_numpy(x.view(np.int64)).cuda() for x in (a, b))
    out = ops.mul(da, db).cpu().numpy().view(np.uint64)
    assert [C.int_from_mont(x, R.R_MOD) for x in out[:64]] == [x * y % R.R_MOD for x, y in zip(ai, bi)]
    tail = C.int_from_mont(out[n - 1], R.R_MOD)
    assert tail == C.int_from_mont(a[n - 1], R.R_MOD) * C.int_from_mont(b[n - 1], R.R_MOD) % R.R_MOD
    scaled = ops.mul(da, db, scale=7, out=da).cpu().numpy().view(np.uint64)          # in place, with the constant
    assert [C.int_from_mont(x, R.R_MOD) for x in scaled[:64]] == [7 * x * y % R.R_MOD for x, y in zip(ai, bi)]
    assert ctx.lib.h2b_fr_mul_device(ctx.handle, None, ctypes.c_void_p(db.data_ptr()), None, ctypes.c_void_p(da.data_ptr()), n) == -1
    assert ctx.lib.h2b_fr_mul_device(ctx.handle, None, None, None, None, 0) == 0


def test_coset_transform_by_scaling(ops, ctx):
    import h2b200
    import torch
    k = 10
    n = 1 << k
    dom = h2b200.EvaluationDomain(1, k, ctx)
    coeffs = C.random_fr(n, 77)
    ci = [C.int_from_mont(x, R.R_MOD) for x in coeffs]
    shift = R.FR_ZETA * pow(R.ROOT_OF_UNITY, 12345, R.R_MOD) % R.R_MOD
    powers = torch.from_numpy(np.stack([C.mont_from_int(pow(shift, t, R.R_MOD), R.R_MOD) for t in range(n)]).view(np.int64)).cuda()
    vals = dom.coeff_to_lagrange(ops.mul(torch.from_numpy(coeffs.view(np.int64)).cuda(), powers)).cpu().numpy().view(np.uint64)
    for r in (0, 1, 17, n - 1):
        x = shift * pow(dom.omega, r, R.R_MOD) % R.R_MOD
        assert C.int_from_mont(vals[r], R.R_MOD) == R.eval_polynomial(ci, x)
