"""Host logic of the quotient-evaluation mirror (no GPU): the GraphEvaluator compiler restated from
plonk/evaluation.rs:644-777 must produce calculation graphs whose value (oracle interpreter, evaluation.rs:789-834) equals the
direct evaluation of the expression (evaluation.rs:837-880), and the wire encoding must round-trip."""
import random

import numpy as np

import h2b200
from h2b200.evaluation import GraphEvaluator
from oracle import pyref as R


def random_expr(rnd, depth, n_fixed, n_advice, n_instance, n_chal):
    if depth == 0 or rnd.random() < 0.25:
        c = rnd.random()
        if c < 0.25:
            return ("const", rnd.choice([0, 1, 2, 3, R.R_MOD - 1, rnd.randrange(R.R_MOD)]))
        if c < 0.9:
            kind = rnd.choice(["fixed"] * (n_fixed > 0) + ["advice"] * (n_advice > 0) + ["instance"] * (n_instance > 0))
            ncol = {"fixed": n_fixed, "advice": n_advice, "instance": n_instance}[kind]
            return ("query", kind, rnd.randrange(ncol), rnd.choice([0, 0, 1, -1, 2, -3]))
        return ("challenge", rnd.randrange(n_chal))
    c = rnd.random()
    if c < 0.15:
        return ("neg", random_expr(rnd, depth - 1, n_fixed, n_advice, n_instance, n_chal))
    if c < 0.55:
        return ("sum", random_expr(rnd, depth - 1, n_fixed, n_advice, n_instance, n_chal),
                random_expr(rnd, depth - 1, n_fixed, n_advice, n_instance, n_chal))
    a = random_expr(rnd, depth - 1, n_fixed, n_advice, n_instance, n_chal)
    b = a if rnd.random() < 0.15 else random_expr(rnd, depth - 1, n_fixed, n_advice, n_instance, n_chal)
    return ("prod", a, b)


def random_columns(rnd, ncols, size):
    return [[rnd.randrange(R.R_MOD) for _ in range(size)] for _ in range(ncols)]


def test_graph_compiler_matches_direct_evaluation():
    rnd = random.Random(7)
    size, rot_scale = 32, 4
    fixed, advice, instance = random_columns(rnd, 2, size), random_columns(rnd, 3, size), random_columns(rnd, 1, size)
    challenges = [rnd.randrange(R.R_MOD) for _ in range(2)]
    y = rnd.randrange(R.R_MOD)
    for trial in range(40):
        g = GraphEvaluator()
        exprs = [random_expr(rnd, 4, 2, 3, 1, 2) for _ in range(rnd.randrange(1, 4))]
        parts = [g.add_expression(e) for e in exprs]
        # custom gates fold as Evaluator::new does (evaluation.rs:223-227): Horner(PreviousValue, parts, Y)
        g.add_calculation(("horner", (10,), tuple(parts), (9,)))
        prev = [rnd.randrange(R.R_MOD) for _ in range(size)]
        got = R.graph_evaluate(g.calculations, g.constants, g.rotations, fixed, advice, instance, challenges, 0, 0, 0, y, prev,
                               size, rot_scale)
        for idx in range(size):
            exp = prev[idx]
            for e in exprs:
                exp = (exp * y + R.evaluate_expression(e, idx, size, rot_scale, fixed, advice, instance, challenges)) % R.R_MOD
            assert got[idx] == exp, (trial, idx)


def test_compiler_simplifications_and_dedup():
    """The special cases of add_expression (evaluation.rs:716-775)."""
    g = GraphEvaluator()
    a = ("query", "advice", 0, 0)
    assert g.add_expression(("prod", ("const", 0), a)) == (0, 0)
    ra = g.add_expression(a)
    assert g.add_expression(("prod", ("const", 1), a)) == ra
    n = g.num_intermediates
    assert g.add_expression(a) == ra and g.num_intermediates == n              # identical calculation reused (:672-690)
    d = g.add_expression(("prod", ("const", 2), a))
    assert g.calculations[d[1]][0] == ("double", ra)
    s = g.add_expression(("prod", a, a))
    assert g.calculations[s[1]][0] == ("square", ra)
    sub = g.add_expression(("sum", a, ("neg", ("query", "fixed", 1, 1))))
    assert g.calculations[sub[1]][0][0] == "sub"
    assert g.add_expression(("neg", ("const", 5))) == (0, g.constants.index(R.R_MOD - 5))
    assert g.add_expression(("sum", ("const", 0), a)) == ra
    assert g.rotations == [0, 1]
    # operands are ordered by the derived PartialOrd (variant index first): Fixed < Advice
    m = g.add_expression(("prod", a, ("query", "fixed", 0, 0)))
    calc = g.calculations[m[1]][0]
    assert calc[0] == "mul" and calc[1] <= calc[2]


def test_wire_encoding():
    g = GraphEvaluator()
    p0 = g.add_expression(("prod", ("query", "advice", 1, -1), ("challenge", 0)))
    g.add_calculation(("horner", (0, 0), (p0, (6,)), (8,)))
    prog, consts, rots = g.encode()
    assert prog.dtype == np.uint32 and consts.shape == (3, 4) and list(rots) == [-1]
    # store advice(1, rot 0): [7, 0, 3 | 0 << 8, 1, 0, 0]
    assert list(prog[:6]) == [7, 0, 3, 1, 0, 0]
    # horner: [6, target, start(Constant 0), factor(Theta), n_parts = 2, parts...]
    h = list(prog[-11:])
    assert h[0] == 6 and h[2:4] == [0, 0] and h[4:6] == [8, 0] and h[6] == 2 and h[7:9] == [1, p0[1]] and h[9:11] == [6, 0]
    assert h2b200.fields.fr_int(consts[2]) == 2
