"""The backend's view of a circuit: halo2_middleware/src/circuit.rs + halo2_backend/src/plonk/circuit.rs + the query collection of
halo2_backend/src/plonk/keygen.rs:192-368.

Expressions are nested tuples (the same encoding the GraphEvaluator mirror in ../evaluation.py consumes):
    ("const", int) | ("query", "fixed"|"advice"|"instance", column_index, rotation) | ("challenge", index)
    | ("neg", e) | ("sum", a, b) | ("prod", a, b)
A column is (kind, index) with kind in {"advice", "fixed", "instance"} (`ColumnMid`, middleware circuit.rs:172-176).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

from ..fields import R_MOD

Expr = tuple
Column = Tuple[str, int]


# ---- small expression helpers (middleware expression.rs) -------------------------------------------------------------
def const(v: int) -> Expr:
    return ("const", v % R_MOD)


def query(kind: str, index: int, rotation: int = 0) -> Expr:
    return ("query", kind, index, rotation)


def challenge(index: int) -> Expr:
    return ("challenge", index)


def neg(a: Expr) -> Expr:
    return ("neg", a)


def add(a: Expr, b: Expr) -> Expr:
    return ("sum", a, b)


def sub(a: Expr, b: Expr) -> Expr:
    return ("sum", a, ("neg", b))                     # `a - b` builds Sum(a, Negated(b)), expression.rs Sub impl


def mul(a: Expr, b: Expr) -> Expr:
    return ("prod", a, b)


def degree(e: Expr) -> int:
    """Expression::degree (middleware expression.rs): queries have degree 1, challenges and constants 0."""
    tag = e[0]
    if tag == "const" or tag == "challenge":
        return 0
    if tag == "query":
        return 1
    if tag == "neg":
        return degree(e[1])
    if tag == "sum":
        return max(degree(e[1]), degree(e[2]))
    if tag == "prod":
        return degree(e[1]) + degree(e[2])
    raise ValueError(tag)


def evaluate_at(e: Expr, fixed_evals, advice_evals, instance_evals, challenges, query_index) -> int:
    """Expression::evaluate with the verifier's closures (plonk/verifier.rs:349-365): a query reads the eval at its query index."""
    tag = e[0]
    if tag == "const":
        return e[1] % R_MOD
    if tag == "query":
        evals = {"fixed": fixed_evals, "advice": advice_evals, "instance": instance_evals}[e[1]]
        return evals[query_index[(e[1], e[2], e[3])]]
    if tag == "challenge":
        return challenges[e[1]]
    ev = lambda x: evaluate_at(x, fixed_evals, advice_evals, instance_evals, challenges, query_index)
    if tag == "neg":
        return (-ev(e[1])) % R_MOD
    if tag == "sum":
        return (ev(e[1]) + ev(e[2])) % R_MOD
    if tag == "prod":
        return ev(e[1]) * ev(e[2]) % R_MOD
    raise ValueError(tag)


# ---- middleware types ------------------------------------------------------------------------------------------------------
@dataclass
class LookupArgument:                                   # middleware lookup.rs
    name: str
    input_expressions: List[Expr]
    table_expressions: List[Expr]


@dataclass
class ShuffleArgument:                                  # middleware shuffle.rs
    name: str
    input_expressions: List[Expr]
    shuffle_expressions: List[Expr]


@dataclass
class ConstraintSystemMid:                              # middleware circuit.rs:85-113
    num_fixed_columns: int
    num_advice_columns: int
    num_instance_columns: int
    num_challenges: int = 0
    unblinded_advice_columns: List[int] = field(default_factory=list)
    advice_column_phase: List[int] = field(default_factory=list)
    challenge_phase: List[int] = field(default_factory=list)
    gates: List[Tuple[str, Expr]] = field(default_factory=list)          # (name, poly)
    permutation_columns: List[Column] = field(default_factory=list)      # permutation::ArgumentMid.columns
    lookups: List[LookupArgument] = field(default_factory=list)
    shuffles: List[ShuffleArgument] = field(default_factory=list)
    minimum_degree: Optional[int] = None


@dataclass
class Preprocessing:                                    # middleware circuit.rs:115-121
    copies: List[Tuple[Tuple[Column, int], Tuple[Column, int]]]          # AssemblyMid.copies: ((column, row), (column, row))
    fixed: Sequence                                                      # per fixed column: n canonical ints, or an (n, 4) uint64
                                                                         # Montgomery array


@dataclass
class CompiledCircuit:                                  # middleware circuit.rs:123-127
    preprocessing: Preprocessing
    cs: ConstraintSystemMid


# ---- backend constraint system -------------------------------------------------------------------------------------------
class ConstraintSystemBack:
    """plonk/circuit.rs:47-77 built by `From<ConstraintSystemMid>` (keygen.rs:343-368): the same expressions plus the ordered query
    lists `collect_queries` assigns (gates, then lookups, then shuffles, then one Rotation::cur() query per permutation column)."""

    def __init__(self, mid: ConstraintSystemMid):
        self.num_fixed_columns = mid.num_fixed_columns
        self.num_advice_columns = mid.num_advice_columns
        self.num_instance_columns = mid.num_instance_columns
        self.num_challenges = mid.num_challenges
        self.unblinded_advice_columns = list(mid.unblinded_advice_columns)
        self.advice_column_phase = list(mid.advice_column_phase) or [0] * mid.num_advice_columns
        self.challenge_phase = list(mid.challenge_phase)
        self.permutation_columns = list(mid.permutation_columns)
        self.minimum_degree = mid.minimum_degree
        self.advice_queries: List[Tuple[int, int]] = []       # (column index, rotation)
        self.instance_queries: List[Tuple[int, int]] = []
        self.fixed_queries: List[Tuple[int, int]] = []
        self.query_index: Dict[Tuple[str, int, int], int] = {}

        def collect(e: Expr):
            tag = e[0]
            if tag == "query":
                self._add_query(e[1], e[2], e[3])
            elif tag == "neg":
                collect(e[1])
            elif tag in ("sum", "prod"):
                collect(e[1])
                collect(e[2])

        self.gates = list(mid.gates)
        for _, poly in self.gates:
            collect(poly)
        self.lookups = list(mid.lookups)
        for l in self.lookups:
            for e in l.input_expressions:
                collect(e)
            for e in l.table_expressions:
                collect(e)
        self.shuffles = list(mid.shuffles)
        for s in self.shuffles:
            for e in s.input_expressions:
                collect(e)
            for e in s.shuffle_expressions:
                collect(e)
        for kind, index in self.permutation_columns:
            self._add_query(kind, index, 0)
        self.num_advice_queries = [0] * self.num_advice_columns
        for col, _ in self.advice_queries:
            self.num_advice_queries[col] += 1

    def _add_query(self, kind: str, index: int, rotation: int) -> int:
        key = (kind, index, rotation)
        if key not in self.query_index:
            lst = {"advice": self.advice_queries, "instance": self.instance_queries, "fixed": self.fixed_queries}[kind]
            lst.append((index, rotation))
            self.query_index[key] = len(lst) - 1
        return self.query_index[key]

    # plonk/circuit.rs:80-117
    def degree(self) -> int:
        deg = 3                                               # permutation_argument_required_degree
        for l in self.lookups:
            assert len(l.input_expressions) == len(l.table_expressions)
            i = max([1] + [degree(e) for e in l.input_expressions])
            t = max([1] + [degree(e) for e in l.table_expressions])
            deg = max(deg, max(4, 2 + i + t))
        for s in self.shuffles:
            assert len(s.input_expressions) == len(s.shuffle_expressions)
            i = max([1] + [degree(e) for e in s.input_expressions])
            t = max([1] + [degree(e) for e in s.shuffle_expressions])
            deg = max(deg, max(2 + t, 2 + i))
        for _, poly in self.gates:
            deg = max(deg, degree(poly))
        return max(deg, self.minimum_degree or 1)

    # plonk/circuit.rs:121-152
    def blinding_factors(self) -> int:
        factors = max(self.num_advice_queries, default=1)
        return max(3, factors) + 2

    def minimum_rows(self) -> int:                            # plonk/circuit.rs:156-163
        return self.blinding_factors() + 3

    def get_any_query_index(self, kind: str, index: int, rotation: int = 0) -> int:   # plonk/circuit.rs:165-178
        key = (kind, index, rotation)
        if key not in self.query_index:
            raise KeyError("get_any_query_index called for non-existent query")
        return self.query_index[key]

    def phases(self) -> List[int]:                            # plonk/circuit.rs:181-189
        return list(range(max(self.advice_column_phase, default=0) + 1))
