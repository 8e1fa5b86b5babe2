"""keygen_vk / keygen_pk: halo2_backend/src/plonk/keygen.rs:43-187, permutation/keygen.rs and the Evaluator constructor
(plonk/evaluation.rs:174-255), with every size-n / size-n*e vector produced on the device through `ops`.

Not a performance target (keygen is outside the proving hot path) but it uses the same kernels: commit_lagrange for
the fixed and sigma columns, lagrange_to_coeff / coeff_to_extended for the cosets the quotient evaluation reads.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from .. import fields as F
from ..evaluation import ADVICE, BETA, CONSTANT, GAMMA, PREVIOUS, THETA, Y, GraphEvaluator
from . import circuit as C
from . import pinned
from . import polyexpr as PX
from .transcript import fr_from_uniform_bytes

FR_DELTA = pow(7, 1 << F.FR_S, F.R_MOD)          # Fr::DELTA = g^(2^S), g = 7


class Assembly:
    """permutation/keygen.rs:20-125: union-find over cells with cycle splicing."""

    def __init__(self, n: int, columns: List[C.Column]):
        self.n = n
        self.columns = list(columns)
        m = len(columns)
        self.map_col = np.repeat(np.arange(m, dtype=np.int64)[:, None], n, axis=1)
        self.map_row = np.repeat(np.arange(n, dtype=np.int64)[None, :], m, axis=0)
        self.aux_col = self.map_col.copy()
        self.aux_row = self.map_row.copy()
        self.sizes = np.ones((m, n), dtype=np.int64)

    @staticmethod
    def new_from_assembly_mid(n: int, columns: List[C.Column], copies) -> "Assembly":
        a = Assembly(n, columns)
        for (lc, lr), (rc, rr) in copies:
            a.copy(lc, lr, rc, rr)
        return a

    def copy(self, left_column: C.Column, left_row: int, right_column: C.Column, right_row: int) -> None:
        try:
            lc, rc = self.columns.index(tuple(left_column)), self.columns.index(tuple(right_column))
        except ValueError as e:
            raise KeyError(f"ColumnNotInPermutation: {e}")
        if left_row >= self.n or right_row >= self.n:
            raise IndexError("BoundsFailure")
        left = (int(self.aux_col[lc, left_row]), int(self.aux_row[lc, left_row]))
        right = (int(self.aux_col[rc, right_row]), int(self.aux_row[rc, right_row]))
        if left == right:
            return
        if self.sizes[left] < self.sizes[right]:
            left, right = right, left
        self.sizes[left] += self.sizes[right]
        i = right
        while True:
            self.aux_col[i], self.aux_row[i] = left
            i = (int(self.map_col[i]), int(self.map_row[i]))
            if i == right:
                break
        tmp = (self.map_col[lc, left_row], self.map_row[lc, left_row])
        self.map_col[lc, left_row], self.map_row[lc, left_row] = self.map_col[rc, right_row], self.map_row[rc, right_row]
        self.map_col[rc, right_row], self.map_row[rc, right_row] = tmp

    def permutations(self, ops, omega: int):
        """The sigma columns in Lagrange form: permutations[i][j] = delta^i' * omega^j' for mapping(i, j) = (i', j')
        (permutation/keygen.rs:139-176, 228-262)."""
        m, n = len(self.columns), self.n
        if m == 0:
            return []
        omega_powers = ops.grand_product(ops.full(n, omega), 1, n)
        deltaomega = [omega_powers if i == 0 else PX.lincomb(ops, [(pow(FR_DELTA, i, F.R_MOD), omega_powers)]) for i in range(m)]
        table = ops.concat(deltaomega)
        return [ops.gather_rows(table, self.map_col[i] * n + self.map_row[i]) for i in range(m)]


class Evaluator:
    """plonk/evaluation.rs:174-255: the calculation graphs of the custom gates, each lookup and each shuffle."""

    def __init__(self, cs: C.ConstraintSystemBack):
        self.custom_gates = GraphEvaluator()
        parts = tuple(self.custom_gates.add_expression(poly) for _, poly in cs.gates)
        self.custom_gates.add_calculation(("horner", (PREVIOUS,), parts, (Y,)))
        self.lookups: List[GraphEvaluator] = []
        for lookup in cs.lookups:
            g = GraphEvaluator()
            lc = lambda exprs: g.add_calculation(("horner", (CONSTANT, 0), tuple(g.add_expression(e) for e in exprs), (THETA,)))
            compressed_input_coset = lc(lookup.input_expressions)
            compressed_table_coset = lc(lookup.table_expressions)
            right_gamma = g.add_calculation(("add", compressed_table_coset, (GAMMA,)))
            left = g.add_calculation(("add", compressed_input_coset, (BETA,)))
            g.add_calculation(("mul", left, right_gamma))
            self.lookups.append(g)
        self.shuffles: List[GraphEvaluator] = []
        for shuffle in cs.shuffles:
            for exprs in (shuffle.input_expressions, shuffle.shuffle_expressions):
                g = GraphEvaluator()
                compressed = g.add_calculation(("horner", (CONSTANT, 0), tuple(g.add_expression(e) for e in exprs), (THETA,)))
                g.add_calculation(("add", compressed, (GAMMA,)))
                self.shuffles.append(g)


@dataclass
class VerifyingKey:                                       # plonk.rs:43-50
    domain: object
    fixed_commitments: list
    permutation_commitments: list                         # permutation::VerifyingKey.commitments
    cs: C.ConstraintSystemBack
    cs_degree: int
    transcript_repr: int = 0

    def hash_into(self, transcript) -> None:              # plonk.rs:205-213
        transcript.common_scalar(self.transcript_repr)


@dataclass
class PermutationProvingKey:                              # permutation.rs ProvingKey
    permutations: list
    polys: list
    cosets: list


@dataclass
class ProvingKey:                                         # plonk.rs:242-252
    vk: VerifyingKey
    l0: object
    l_last: object
    l_active_row: object
    fixed_values: list
    fixed_polys: list
    fixed_cosets: list
    permutation: PermutationProvingKey
    ev: Evaluator
    l0_poly: object = None                                # coefficient forms of l0 / l_last / l_active_row: the coset prover
    l_last_poly: object = None                            # (cosets.py) evaluates them coset by coset
    l_active_poly: object = None

    def get_vk(self) -> VerifyingKey:
        return self.vk


def _fixed_column(ops, col):
    if isinstance(col, np.ndarray) and col.ndim == 2:
        return ops.upload(col)
    return ops.from_ints(list(col))


def _transcript_repr(vk: VerifyingKey) -> int:
    """plonk.rs:189-203: BLAKE2b-512, personal "Halo2-Verify-Key", over len(s) as u64 LE || s, s = `{:?}` of the pinned key
    (pinned.py; the field / point formatting inside it is recalled from halo2curves, see there)."""
    s = pinned.verification_key(vk).encode()
    h = hashlib.blake2b(digest_size=64, person=b"Halo2-Verify-Key")
    h.update(len(s).to_bytes(8, "little"))
    h.update(s)
    return fr_from_uniform_bytes(h.digest())


def keygen_vk(ops, circuit: C.CompiledCircuit, transcript_repr: Optional[int] = None) -> VerifyingKey:
    """keygen.rs:43-100.  `ops` is bound to a domain of the right k; the quotient degree is re-derived from the circuit.
    `transcript_repr`: the value `VerifyingKey::transcript_repr()` returned in a Rust run of the same circuit, used verbatim instead
    of the restated hash (the escape hatch while halo2curves' formatters cannot be checked here)."""
    cs = C.ConstraintSystemBack(circuit.cs)
    n = ops.n
    if n < cs.minimum_rows():
        raise ValueError(f"not enough rows available (k = {ops.k})")
    assert ops.domain.quotient_poly_degree == cs.degree() - 1, "ops.domain must be EvaluationDomain::new(cs.degree(), k)"
    assembly = Assembly.new_from_assembly_mid(n, cs.permutation_columns, circuit.preprocessing.copies)
    sigma = assembly.permutations(ops, ops.domain.omega)
    permutation_commitments = [ops.commit_lagrange(p) for p in sigma]
    fixed_commitments = [ops.commit_lagrange(_fixed_column(ops, col)) for col in circuit.preprocessing.fixed]
    vk = VerifyingKey(ops.domain, fixed_commitments, permutation_commitments, cs, cs.degree())
    vk.transcript_repr = _transcript_repr(vk) if transcript_repr is None else transcript_repr % F.R_MOD
    return vk


def indicator_polys(ops, cs):
    """keygen.rs:134-167 in coefficient form: l_0, l_blind (the last `blinding_factors` rows), l_last, and l_active = 1 - (l_last +
    l_blind).  They depend on the constraint system and the domain only, so a key that was read back (serialization.pk_read) gets
    them the same way a generated one does."""
    n, bf = ops.n, cs.blinding_factors()

    def indicator(rows):
        p = ops.zeros(n)
        for r in rows:
            ops.write_ints(p, r, [1])
        return ops.lagrange_to_coeff(p)

    l0_poly, l_blind_poly, l_last_poly = indicator([0]), indicator(range(n - bf, n)), indicator([n - bf - 1])
    # 1 - (l_last + l_blind) in coefficient form: negate every coefficient, add one to the constant term
    l_active_poly = PX.pointwise(ops, C.neg(C.add(PX.col(0), PX.col(1))), [l_last_poly, l_blind_poly])
    ops.write_ints(l_active_poly, 0, [(ops.read_ints(l_active_poly, 0, 1)[0] + 1) % F.R_MOD])
    return l0_poly, l_blind_poly, l_last_poly, l_active_poly


def _l_active_row(ops, l_last_ext, l_blind_ext):
    one_minus = C.sub(C.const(1), C.add(PX.col(0), PX.col(1)))
    return PX.pointwise(ops, one_minus, [l_last_ext, l_blind_ext], extended=True)


def extended_parts(ops, pk):
    """(l0, l_last, l_active_row, fixed_cosets, permutation cosets) in the extended domain -- what ProvingKey::write stores
    (plonk.rs:276-287).  A key made with keygen_pk(extended=False) does not hold them; they are rebuilt here from its coefficient
    forms (nothing is cached: such a key exists to avoid keeping n * e rows per column)."""
    if pk.l0 is not None:
        return pk.l0, pk.l_last, pk.l_active_row, pk.fixed_cosets, pk.permutation.cosets
    ext = ops.coeff_to_extended
    _, l_blind_poly, _, _ = indicator_polys(ops, pk.vk.cs)
    l_last = ext(pk.l_last_poly)
    return (ext(pk.l0_poly), l_last, _l_active_row(ops, l_last, ext(l_blind_poly)), [ext(p) for p in pk.fixed_polys],
            [ext(p) for p in pk.permutation.polys])


def keygen_pk(ops, vk: VerifyingKey, circuit: C.CompiledCircuit, extended: bool = True) -> ProvingKey:
    """keygen.rs:103-187."""
    cs, n = vk.cs, ops.n
    if n < cs.minimum_rows():
        raise ValueError(f"not enough rows available (k = {ops.k})")
    fixed_values = [_fixed_column(ops, col) for col in circuit.preprocessing.fixed]
    fixed_polys = [ops.lagrange_to_coeff(ops.clone(v)) for v in fixed_values]
    # extended=False: a key for the coset prover (cosets.py) -- nothing of the extended length n * e is built or kept
    ext = (lambda p: ops.coeff_to_extended(p)) if extended else (lambda p: None)
    fixed_cosets = [ext(p) for p in fixed_polys] if extended else []
    l0_poly, l_blind_poly, l_last_poly, l_active_poly = indicator_polys(ops, cs)
    l0, l_last = ext(l0_poly), ext(l_last_poly)
    l_active_row = _l_active_row(ops, l_last, ext(l_blind_poly)) if extended else None
    assembly = Assembly.new_from_assembly_mid(n, cs.permutation_columns, circuit.preprocessing.copies)
    permutations = assembly.permutations(ops, ops.domain.omega)
    polys = [ops.lagrange_to_coeff(ops.clone(p)) for p in permutations]
    cosets = [ext(p) for p in polys] if extended else []
    return ProvingKey(vk, l0, l_last, l_active_row, fixed_values, fixed_polys, fixed_cosets,
                      PermutationProvingKey(permutations, polys, cosets), Evaluator(cs), l0_poly, l_last_poly, l_active_poly)
