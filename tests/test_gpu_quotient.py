"""GPU parity of the quotient-evaluation kernels (GraphEvaluator::evaluate and the permutation / lookup / shuffle
folds of Evaluator::evaluate_h) against the python-int restatement of plonk/evaluation.rs.  Bit-exact."""
import random

import numpy as np
import pytest

from oracle import cref as C
from oracle import pyref as R
from test_evaluation_host import random_expr

pytestmark = pytest.mark.gpu


def to_dev(col):
    import torch
    return torch.from_numpy(C.fr_array_from_ints(col, R.R_MOD).view(np.int64)).cuda()


def to_ints(t):
    return C.ints_from_fr_array(t.cpu().numpy().view(np.uint64), R.R_MOD)


def rand_cols(rnd, ncols, size):
    return [[rnd.randrange(R.R_MOD) for _ in range(size)] for _ in range(ncols)]


@pytest.mark.parametrize("k,ext_k", [(3, 5), (4, 6), (5, 5), (6, 9)])
def test_graph_evaluate_random_programs(ctx, k, ext_k):
    from h2b200.evaluation import GraphEvaluator
    rnd = random.Random(100 * k + ext_k)
    size, rot_scale = 1 << ext_k, 1 << (ext_k - k)
    fixed, advice, instance = rand_cols(rnd, 2, size), rand_cols(rnd, 3, size), rand_cols(rnd, 1, size)
    d_fixed, d_advice, d_instance = [to_dev(c) for c in fixed], [to_dev(c) for c in advice], [to_dev(c) for c in instance]
    challenges = [rnd.randrange(R.R_MOD) for _ in range(2)]
    beta, gamma, theta, y = (rnd.randrange(R.R_MOD) for _ in range(4))
    for trial in range(6):
        g = GraphEvaluator()
        exprs = [random_expr(rnd, 4, 2, 3, 1, 2) for _ in range(rnd.randrange(1, 5))]
        parts = [g.add_expression(e) for e in exprs]
        if trial % 2 == 0:      # custom gates: Horner(PreviousValue, parts, Y) (evaluation.rs:223-227)
            g.add_calculation(("horner", (10,), tuple(parts), (9,)))
        else:                   # lookup shape (evaluation.rs:233-262): Horner(Constant(0), parts, Theta), + Gamma, + Beta, Mul
            lc = g.add_calculation(("horner", (0, 0), tuple(parts), (8,)))
            right = g.add_calculation(("add", lc, (7,)))
            left = g.add_calculation(("add", lc, (6,)))
            g.add_calculation(("mul", left, right))
        prev = [rnd.randrange(R.R_MOD) for _ in range(size)]
        exp = R.graph_evaluate(g.calculations, g.constants, g.rotations, fixed, advice, instance, challenges, beta, gamma, theta, y,
                               prev, size, rot_scale)
        d_prev = to_dev(prev)
        out = g.evaluate(ctx, d_fixed, d_advice, d_instance, challenges, beta, gamma, theta, y, k, ext_k, previous=d_prev)
        assert to_ints(out) == exp, trial
        # in place over the previous values (how evaluate_h folds the custom gates into `values`)
        g.evaluate(ctx, d_fixed, d_advice, d_instance, challenges, beta, gamma, theta, y, k, ext_k, previous=d_prev, out=d_prev)
        assert to_ints(d_prev) == exp


def test_graph_edge_cases(ctx):
    import h2b200
    from h2b200.evaluation import GraphEvaluator
    k, ext_k = 3, 4
    size = 1 << ext_k
    col = [i + 1 for i in range(size)]
    d = to_dev(col)
    g = GraphEvaluator()          # no calculations at all -> zero (evaluation.rs:829-833)
    assert to_ints(g.evaluate(ctx, [], [d], [], [], 0, 0, 0, 0, k, ext_k)) == [0] * size
    g = GraphEvaluator()          # negative rotation wraps (rem_euclid, evaluation.rs:22-24)
    g.add_expression(("query", "advice", 0, -5))
    assert to_ints(g.evaluate(ctx, [], [d], [], [], 0, 0, 0, 0, k, ext_k)) == [col[(i - 10) % size] for i in range(size)]
    g = GraphEvaluator()          # a column index the call does not supply is rejected on the host
    g.add_expression(("query", "fixed", 3, 0))
    with pytest.raises(h2b200.H2bError):
        g.evaluate(ctx, [d], [], [], [], 0, 0, 0, 0, k, ext_k)


@pytest.mark.parametrize("k,ext_k,n_cols,chunk_len", [(4, 6, 3, 3), (4, 6, 5, 2), (5, 7, 4, 4), (3, 6, 1, 3)])
def test_permutation_fold(ctx, k, ext_k, n_cols, chunk_len):
    from h2b200 import evaluation as E
    rnd = random.Random(k * 1000 + n_cols)
    size, rot_scale = 1 << ext_k, 1 << (ext_k - k)
    n_sets = (n_cols + chunk_len - 1) // chunk_len
    sets, cv, cp = rand_cols(rnd, n_sets, size), rand_cols(rnd, n_cols, size), rand_cols(rnd, n_cols, size)
    l0, l_last, l_active, values = rand_cols(rnd, 4, size)
    beta, gamma, y = (rnd.randrange(R.R_MOD) for _ in range(3))
    ext_omega = pow(R.ROOT_OF_UNITY, 1 << (28 - ext_k), R.R_MOD)
    delta = pow(7, 1 << 28, R.R_MOD)                        # Fr::DELTA = GENERATOR^(2^S)
    blinding = 5
    exp = R.permutation_fold(values, sets, cv, cp, chunk_len, l0, l_last, l_active, beta, gamma, y, R.FR_ZETA, delta, ext_omega,
                             blinding, size, rot_scale)
    d_values = to_dev(values)
    E.permutation_fold(ctx, d_values, [to_dev(c) for c in sets], [to_dev(c) for c in cv], [to_dev(c) for c in cp], chunk_len,
                       to_dev(l0), to_dev(l_last), to_dev(l_active), beta, gamma, y, R.FR_ZETA, delta, ext_omega, blinding, k, ext_k)
    assert to_ints(d_values) == exp


def test_permutation_fold_no_sets_is_a_no_op(ctx):
    from h2b200 import evaluation as E
    values = [3, 4, 5, 6]
    d = to_dev(values)
    E.permutation_fold(ctx, d, [], [], [], 3, d, d, d, 1, 2, 3, R.FR_ZETA, 5, 7, 5, 1, 2)
    assert to_ints(d) == values


@pytest.mark.parametrize("k,ext_k", [(3, 5), (5, 8), (6, 6)])
def test_lookup_and_shuffle_folds(ctx, k, ext_k):
    from h2b200 import evaluation as E
    rnd = random.Random(k * 77 + ext_k)
    size, rot_scale = 1 << ext_k, 1 << (ext_k - k)
    table_values, product, p_in, p_tab, l0, l_last, l_active, values = rand_cols(rnd, 8, size)
    beta, gamma, y = (rnd.randrange(R.R_MOD) for _ in range(3))
    exp = R.lookup_fold(values, table_values, product, p_in, p_tab, l0, l_last, l_active, beta, gamma, y, size, rot_scale)
    d_values = to_dev(values)
    dl0, dll, dla = to_dev(l0), to_dev(l_last), to_dev(l_active)
    E.lookup_fold(ctx, d_values, to_dev(table_values), to_dev(product), to_dev(p_in), to_dev(p_tab), dl0, dll, dla, beta, gamma, y,
                  k, ext_k)
    assert to_ints(d_values) == exp
    input_values, shuffle_values = rand_cols(rnd, 2, size)
    exp = R.shuffle_fold(values, input_values, shuffle_values, product, l0, l_last, l_active, y, size, rot_scale)
    d_values = to_dev(values)
    E.shuffle_fold(ctx, d_values, to_dev(input_values), to_dev(shuffle_values), to_dev(product), dl0, dll, dla, y, k, ext_k)
    assert to_ints(d_values) == exp


def test_evaluate_h_shape_end_to_end(ctx):
    """A miniature evaluate_h (evaluation.rs:317-623): coset NTTs of the columns on the device, custom gate + permutation + lookup
    folds, then extended_to_coeff with the vanishing division -- all device resident -- against the same sequence in the oracle."""
    import h2b200
    import torch
    from h2b200 import evaluation as E
    rnd = random.Random(5)
    k, j = 4, 4                                    # degree 4 -> quotient degree 3 -> extended_k = 6
    dom, ref = h2b200.EvaluationDomain(j, k, ctx), R.EvaluationDomain(j, k)
    ext_k, size, n = ref.extended_k, 1 << ref.extended_k, 1 << k
    rot_scale = 1 << (ext_k - k)
    coeff_cols = rand_cols(rnd, 6, n)              # a0, a1, f0, z, l0, l_active as coefficient-form polynomials
    ext_ref = [ref.coeff_to_extended(list(c)) for c in coeff_cols]
    ext_dev = [dom.coeff_to_extended(to_dev(c)) for c in coeff_cols]
    for a, b in zip(ext_dev, ext_ref):
        assert to_ints(a) == b
    a0, a1, f0, z, l0, l_act = ext_ref
    d_a0, d_a1, d_f0, d_z, d_l0, d_lact = ext_dev
    beta, gamma, theta, y = (rnd.randrange(R.R_MOD) for _ in range(4))
    gate = ("prod", ("query", "fixed", 0, 0), ("sum", ("prod", ("query", "advice", 0, 0), ("query", "advice", 1, 1)),
                                               ("neg", ("query", "advice", 0, -1))))
    g = E.GraphEvaluator()
    g.add_calculation(("horner", (10,), (g.add_expression(gate),), (9,)))
    values = torch.zeros((size, 4), dtype=torch.int64, device="cuda")
    g.evaluate(ctx, [d_f0], [d_a0, d_a1], [], [], beta, gamma, theta, y, k, ext_k, previous=values, out=values)
    exp = R.graph_evaluate(g.calculations, g.constants, g.rotations, [f0], [a0, a1], [], [], beta, gamma, theta, y, None, size, rot_scale)
    assert to_ints(values) == exp
    delta, ext_omega = pow(7, 1 << 28, R.R_MOD), ref.extended_omega
    E.permutation_fold(ctx, values, [d_z], [d_a0, d_a1], [d_f0, d_l0], 2, d_l0, d_lact, d_lact, beta, gamma, y, R.FR_ZETA, delta,
                       ext_omega, 5, k, ext_k)
    exp = R.permutation_fold(exp, [z], [a0, a1], [f0, l0], 2, l0, l_act, l_act, beta, gamma, y, R.FR_ZETA, delta, ext_omega, 5, size, rot_scale)
    assert to_ints(values) == exp
    h = dom.extended_to_coeff(values, divide_by_vanishing=True)
    exp_h = ref.extended_to_coeff(ref.divide_by_vanishing_poly(exp))
    assert to_ints(h) == exp_h
