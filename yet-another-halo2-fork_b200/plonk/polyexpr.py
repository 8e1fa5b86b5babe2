"""Pointwise polynomial arithmetic expressed as GraphEvaluator programs, so that the one interpreter kernel
(h2b_quotient_graph_device) serves every element-wise step of the prover: `acc * theta + expr` (lookup/prover.rs:96-98),
`(beta + a) * (gamma + s)` (lookup/prover.rs:196-206), `poly * scalar + poly` (shplonk/prover.rs:152-158, 214-219) ...
Inputs are referenced as ("query", "advice", i, rotation) = cols[i]."""
from __future__ import annotations

from typing import Sequence, Tuple

from ..evaluation import CONSTANT, INTERMEDIATE, GraphEvaluator
from . import circuit as C


def compile_expr(expr) -> GraphEvaluator:
    g = GraphEvaluator()
    res = g.add_expression(expr)
    if not g.calculations or res != (INTERMEDIATE, g.calculations[-1][1]):
        g.calculations.append((("store", res), g.num_intermediates))      # make the result the last calculation
        g.num_intermediates += 1
    return g


def pointwise(ops, expr, cols: Sequence, extended: bool = False, challenges: Sequence[int] = ()):
    return ops.graph(compile_expr(expr), [], list(cols), [], list(challenges), 0, 0, 0, 0, extended)


def col(i: int, rotation: int = 0):
    return C.query("advice", i, rotation)


def lincomb(ops, terms: Sequence[Tuple[int, object]], extended: bool = False):
    """sum_i c_i * P_i for equal-length polynomials."""
    assert terms
    expr = None
    for i, (c, _) in enumerate(terms):
        t = C.mul(C.const(c), col(i))
        expr = t if expr is None else C.add(expr, t)
    return pointwise(ops, expr, [p for _, p in terms], extended)


def horner(ops, polys: Sequence, factor: int, extended: bool = False):
    """fold(0, |acc, p| acc * factor + p) over `polys` in order (the compress_expressions fold, lookup/prover.rs:96-98)."""
    g = GraphEvaluator()
    parts = tuple(g.add_calculation(("store", (3, i, g.add_rotation(0)))) for i in range(len(polys)))
    g.add_calculation(("horner", (CONSTANT, 0), parts, (5, 0)))
    return ops.graph(g, [], list(polys), [], [factor], 0, 0, 0, 0, extended)
