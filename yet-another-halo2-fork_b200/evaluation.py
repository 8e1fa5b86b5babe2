"""Host mirror of halo2_backend/src/plonk/evaluation.rs with the per-row work on the GPU.

    GraphEvaluator            evaluation.rs:187-196, 626-834   the expression -> calculation graph compiler (host logic, restated
                                                               so that tests build graphs exactly as the reference does) and
                                                               `evaluate` over every row of the extended domain on the device
    permutation_fold          evaluation.rs:397-493
    lookup_fold               evaluation.rs:517-565
    shuffle_fold              evaluation.rs:573-620

Expressions are the reference's `ExpressionBack` as nested tuples:
    ("const", int) | ("query", "fixed"|"advice"|"instance", column_index, rotation) | ("challenge", index)
    | ("neg", e) | ("sum", a, b) | ("prod", a, b)
Value sources are tuples whose ordering matches the derived `PartialOrd` of the reference enum (variant index first):
    (0, i) Constant  (1, i) Intermediate  (2, col, rot) Fixed  (3, col, rot) Advice  (4, col, rot) Instance  (5, i) Challenge
    (6,) Beta  (7,) Gamma  (8,) Theta  (9,) Y  (10,) PreviousValue
Columns are torch CUDA tensors of shape (2^extended_k, 4) int64 (Montgomery Fr); results stay on the device.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from . import fields as F
from .engine import Context

CONSTANT, INTERMEDIATE, FIXED, ADVICE, INSTANCE, CHALLENGE, BETA, GAMMA, THETA, Y, PREVIOUS = range(11)
_OPS = {"add": 0, "sub": 1, "mul": 2, "square": 3, "double": 4, "negate": 5, "horner": 6, "store": 7}
_COLUMN_KIND = {"fixed": FIXED, "advice": ADVICE, "instance": INSTANCE}


def _vp(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _src_words(v: Tuple[int, ...]) -> Tuple[int, int]:
    kind = v[0]
    if kind in (FIXED, ADVICE, INSTANCE):
        return kind | (v[2] << 8), v[1]
    if kind in (CONSTANT, INTERMEDIATE, CHALLENGE):
        return kind, v[1]
    return kind, 0


class GraphEvaluator:
    """evaluation.rs:187-196 (+ Default at :626-640: constants start as [0, 1, 2])."""

    def __init__(self):
        self.constants: List[int] = [0, 1, 2]
        self.rotations: List[int] = []
        self.calculations: List[Tuple[tuple, int]] = []       # (calculation, target)
        self.num_intermediates = 0

    # evaluation.rs:644-654
    def add_rotation(self, rotation: int) -> int:
        if rotation in self.rotations:
            return self.rotations.index(rotation)
        self.rotations.append(rotation)
        return len(self.rotations) - 1

    # evaluation.rs:657-666
    def add_constant(self, constant: int):
        constant %= F.R_MOD
        if constant in self.constants:
            return (CONSTANT, self.constants.index(constant))
        self.constants.append(constant)
        return (CONSTANT, len(self.constants) - 1)

    # evaluation.rs:672-690
    def add_calculation(self, calculation: tuple):
        for calc, target in self.calculations:
            if calc == calculation:
                return (INTERMEDIATE, target)
        target = self.num_intermediates
        self.calculations.append((calculation, target))
        self.num_intermediates += 1
        return (INTERMEDIATE, target)

    # evaluation.rs:693-777
    def add_expression(self, expr: tuple):
        tag = expr[0]
        if tag == "const":
            return self.add_constant(expr[1])
        if tag == "query":
            rot_idx = self.add_rotation(expr[3])
            return self.add_calculation(("store", (_COLUMN_KIND[expr[1]], expr[2], rot_idx)))
        if tag == "challenge":
            return self.add_calculation(("store", (CHALLENGE, expr[1])))
        if tag == "neg":
            a = expr[1]
            if a[0] == "const":
                return self.add_constant(-a[1])
            ra = self.add_expression(a)
            return ra if ra == (CONSTANT, 0) else self.add_calculation(("negate", ra))
        if tag == "sum":
            a, b = expr[1], expr[2]
            if b[0] == "neg":
                ra, rb = self.add_expression(a), self.add_expression(b[1])
                if ra == (CONSTANT, 0):
                    return self.add_calculation(("negate", rb))
                if rb == (CONSTANT, 0):
                    return ra
                return self.add_calculation(("sub", ra, rb))
            ra, rb = self.add_expression(a), self.add_expression(b)
            if ra == (CONSTANT, 0):
                return rb
            if rb == (CONSTANT, 0):
                return ra
            return self.add_calculation(("add", ra, rb) if ra <= rb else ("add", rb, ra))
        if tag == "prod":
            ra, rb = self.add_expression(expr[1]), self.add_expression(expr[2])
            if ra == (CONSTANT, 0) or rb == (CONSTANT, 0):
                return (CONSTANT, 0)
            if ra == (CONSTANT, 1):
                return rb
            if rb == (CONSTANT, 1):
                return ra
            if ra == (CONSTANT, 2):
                return self.add_calculation(("double", rb))
            if rb == (CONSTANT, 2):
                return self.add_calculation(("double", ra))
            if ra == rb:
                return self.add_calculation(("square", ra))
            return self.add_calculation(("mul", ra, rb) if ra <= rb else ("mul", rb, ra))
        raise ValueError(f"unknown expression node {tag!r}")

    # ---- the wire format of include/h2b200.h `h2b_graph` ----------------------------------------------------------
    def encode(self):
        words: List[int] = []
        for calc, target in self.calculations:
            op = _OPS[calc[0]]
            if calc[0] == "horner":
                start, parts, factor = calc[1], calc[2], calc[3]
                words += [op, target, *_src_words(start), *_src_words(factor), len(parts)]
                for part in parts:
                    words += list(_src_words(part))
            else:
                a = calc[1]
                b = calc[2] if len(calc) > 2 else (CONSTANT, 0)
                words += [op, target, *_src_words(a), *_src_words(b)]
        return (np.array(words, dtype=np.uint32), F.fr_array(self.constants) if self.constants else np.zeros((0, 4), np.uint64),
                np.array(self.rotations, dtype=np.int32))

    # evaluation.rs:789-834 for every row
    def evaluate(self, ctx: Context, fixed: Sequence, advice: Sequence, instance: Sequence, challenges: Sequence[int], beta: int,
                 gamma: int, theta: int, y: int, k: int, extended_k: int, previous=None, out=None):
        import torch
        prog, consts, rots = self.encode()
        g = _lib.Graph()
        g.program = prog.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)) if prog.size else None
        g.n_words = prog.size
        g.constants32 = _vp(consts) if consts.size else None
        g.n_constants = consts.shape[0]
        g.rotations = rots.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) if rots.size else None
        g.n_rotations = rots.size
        g.n_intermediates = self.num_intermediates

        def table(cols):
            arr = (ctypes.c_void_p * max(len(cols), 1))(*[c.data_ptr() for c in cols])
            return arr, len(cols)

        size = 1 << extended_k
        for c in list(fixed) + list(advice) + list(instance):
            assert c.is_cuda and c.numel() * c.element_size() == size * 32, "columns must be extended-domain device tensors"
        if out is None:
            out = torch.empty((size, 4), dtype=torch.int64, device="cuda")
        (ft, nf), (at, na), (it, ni) = table(fixed), table(advice), table(instance)
        ch = F.fr_array(list(challenges)) if len(challenges) else np.zeros((0, 4), np.uint64)
        sc = [F.fr_mont(v) for v in (beta, gamma, theta, y)]
        ctx.check(ctx.lib.h2b_quotient_graph_device(
            ctx.handle, ctypes.byref(g), ft, nf, at, na, it, ni, _vp(ch) if ch.size else None, ch.shape[0], _vp(sc[0]), _vp(sc[1]),
            _vp(sc[2]), _vp(sc[3]), ctypes.c_void_p(previous.data_ptr()) if previous is not None else None,
            ctypes.c_void_p(out.data_ptr()), k, extended_k))
        return out


def permutation_fold(ctx: Context, values, sets: Sequence, col_values: Sequence, col_perms: Sequence, chunk_len: int, l0, l_last,
                     l_active, beta: int, gamma: int, y: int, zeta: int, delta: int, extended_omega: int, blinding_factors: int,
                     k: int, extended_k: int):
    """evaluation.rs:397-493, in place on `values`.  `sets` = permutation_product_coset per set; `col_values` / `col_perms` the
    permuted columns' cosets and pk.permutation.cosets in p.columns order; chunk_len = cs.degree() - 2."""
    assert len(col_values) == len(col_perms)
    st = (ctypes.c_void_p * max(len(sets), 1))(*[s.data_ptr() for s in sets])
    cv = (ctypes.c_void_p * max(len(col_values), 1))(*[c.data_ptr() for c in col_values])
    cp = (ctypes.c_void_p * max(len(col_perms), 1))(*[c.data_ptr() for c in col_perms])
    m = [F.fr_mont(v) for v in (beta, gamma, y, beta * zeta % F.R_MOD, delta, extended_omega)]
    ctx.check(ctx.lib.h2b_quotient_permutation_device(
        ctx.handle, ctypes.c_void_p(values.data_ptr()), st, len(sets), cv, cp, len(col_values), chunk_len,
        ctypes.c_void_p(l0.data_ptr()), ctypes.c_void_p(l_last.data_ptr()), ctypes.c_void_p(l_active.data_ptr()),
        _vp(m[0]), _vp(m[1]), _vp(m[2]), _vp(m[3]), _vp(m[4]), _vp(m[5]), -(blinding_factors + 1), k, extended_k))
    return values


def lookup_fold(ctx: Context, values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta: int,
                gamma: int, y: int, k: int, extended_k: int):
    """evaluation.rs:517-565, in place on `values`."""
    m = [F.fr_mont(v) for v in (beta, gamma, y)]
    ptr = lambda t: ctypes.c_void_p(t.data_ptr())
    ctx.check(ctx.lib.h2b_quotient_lookup_device(ctx.handle, ptr(values), ptr(table_values), ptr(product), ptr(permuted_input),
                                                 ptr(permuted_table), ptr(l0), ptr(l_last), ptr(l_active), _vp(m[0]), _vp(m[1]),
                                                 _vp(m[2]), k, extended_k))
    return values


def shuffle_fold(ctx: Context, values, input_values, shuffle_values, product, l0, l_last, l_active, y: int, k: int,
                 extended_k: int):
    """evaluation.rs:573-620, in place on `values`."""
    ym = F.fr_mont(y)
    ptr = lambda t: ctypes.c_void_p(t.data_ptr())
    ctx.check(ctx.lib.h2b_quotient_shuffle_device(ctx.handle, ptr(values), ptr(input_values), ptr(shuffle_values), ptr(product),
                                                  ptr(l0), ptr(l_last), ptr(l_active), _vp(ym), k, extended_k))
    return values
