"""`format!("{:?}", vk.pinned())`: the string `VerifyingKey::from_parts` hashes into `transcript_repr`
(halo2_backend/src/plonk.rs:189-203, 238-246).

Written as the chain of `Formatter::debug_struct / debug_tuple / debug_list` calls that `#[derive(Debug)]` expands to in its compact
({:?}) form: `Name { field: value, .. }`, `Name(value, ..)`, `[a, b]`, unit variants bare.  Who prints what:
    PinnedVerificationKey                    plonk.rs:238-246 (derive; field order of the struct)
    PinnedEvaluationDomain                   poly/domain.rs:469-476
    PinnedConstraintSystem                   plonk/circuit.rs:257-285 (hand-written: multi-phase fields only when num_challenges > 0,
                                             `shuffles` only when non-empty); PinnedGates prints the gate polynomials alone (:230-238)
    Expression / VarBack / QueryBack         halo2_middleware/src/expression.rs:22-34, backend circuit.rs:8-26
    ColumnMid, Any, ChallengeMid, Rotation   halo2_middleware/src/circuit.rs:9-13, 166-207, poly.rs:9
    lookup / shuffle Argument, ArgumentMid   halo2_middleware/src/{lookup,shuffle}.rs:6-11, permutation.rs ArgumentMid
    permutation::VerifyingKey                halo2_backend/src/plonk/permutation.rs (derive: `commitments`)
The structure is pinned by the reference's own golden text (halo2_proofs/tests/plonk_api.rs:664-1141, through the oracle's
formatter, tests/test_oracle_prover.py + tests/test_plonk_host.py).  RECALLED, not checkable in this tree (halo2curves 0.6 is not
vendored): `{:?}` of Fr / Fq is "0x" + 64 lower-case hex digits of the canonical value, of G1Affine "(x, y)" or "Infinity", and
`MODULUS` is the lower-case "0x.." string.  A caller holding a value computed by the Rust code passes it to
`keygen_vk(..., transcript_repr=)` instead.
"""
from __future__ import annotations

from .. import fields as F

_ANY = {"advice": "Advice", "fixed": "Fixed", "instance": "Instance"}


def _field(v: int) -> str:
    return "0x" + format(v, "064x")


def _point(p) -> str:
    return "Infinity" if p is None else "(" + _field(p[0]) + ", " + _field(p[1]) + ")"


def _list(items) -> str:
    return "[" + ", ".join(items) + "]"


def _string(s: str) -> str:
    return '"' + s.replace("\\", "\\\\").replace('"', '\\"') + '"'


def _column(col) -> str:
    return "ColumnMid { column_type: " + _ANY[col[0]] + ", index: " + str(col[1]) + " }"


def _queries(kind: str, queries) -> str:
    return _list("(" + _column((kind, c)) + ", Rotation(" + str(r) + "))" for c, r in queries)


def expression(cs, e) -> str:
    tag = e[0]
    if tag == "const":
        return "Constant(" + _field(e[1] % F.R_MOD) + ")"
    if tag == "query":
        return ("Var(Query(QueryBack { index: " + str(cs.query_index[(e[1], e[2], e[3])]) + ", column_index: " + str(e[2])
                + ", column_type: " + _ANY[e[1]] + ", rotation: Rotation(" + str(e[3]) + ") }))")
    if tag == "challenge":
        return "Var(Challenge(ChallengeMid { index: " + str(e[1]) + ", phase: " + str(cs.challenge_phase[e[1]]) + " }))"
    if tag == "neg":
        return "Negated(" + expression(cs, e[1]) + ")"
    if tag == "sum":
        return "Sum(" + expression(cs, e[1]) + ", " + expression(cs, e[2]) + ")"
    if tag == "prod":
        return "Product(" + expression(cs, e[1]) + ", " + expression(cs, e[2]) + ")"
    raise ValueError(tag)


def constraint_system(cs) -> str:
    ex = lambda es: _list(expression(cs, e) for e in es)
    parts = ["num_fixed_columns: %d" % cs.num_fixed_columns, "num_advice_columns: %d" % cs.num_advice_columns,
             "num_instance_columns: %d" % cs.num_instance_columns]
    if cs.num_challenges > 0:
        parts += ["num_challenges: %d" % cs.num_challenges, "advice_column_phase: " + _list(map(str, cs.advice_column_phase)),
                  "challenge_phase: " + _list(map(str, cs.challenge_phase))]
    parts += ["gates: " + ex(poly for _, poly in cs.gates),
              "advice_queries: " + _queries("advice", cs.advice_queries),
              "instance_queries: " + _queries("instance", cs.instance_queries),
              "fixed_queries: " + _queries("fixed", cs.fixed_queries),
              "permutation: ArgumentMid { columns: " + _list(_column(c) for c in cs.permutation_columns) + " }",
              "lookups: " + _list("Argument { name: " + _string(l.name) + ", input_expressions: " + ex(l.input_expressions)
                                  + ", table_expressions: " + ex(l.table_expressions) + " }" for l in cs.lookups)]
    if cs.shuffles:
        parts.append("shuffles: " + _list("Argument { name: " + _string(s.name) + ", input_expressions: " + ex(s.input_expressions)
                                          + ", shuffle_expressions: " + ex(s.shuffle_expressions) + " }" for s in cs.shuffles))
    parts.append("minimum_degree: " + ("None" if cs.minimum_degree is None else "Some(%d)" % cs.minimum_degree))
    return "PinnedConstraintSystem { " + ", ".join(parts) + " }"


def verification_key(vk) -> str:
    d = vk.domain
    return ("PinnedVerificationKey { base_modulus: " + _string(_field(F.Q_MOD)) + ", scalar_modulus: " + _string(_field(F.R_MOD))
            + ", domain: PinnedEvaluationDomain { k: %d, extended_k: %d, omega: %s }" % (d.k, d.extended_k, _field(d.omega))
            + ", cs: " + constraint_system(vk.cs)
            + ", fixed_commitments: " + _list(_point(c) for c in vk.fixed_commitments)
            + ", permutation: VerifyingKey { commitments: " + _list(_point(c) for c in vk.permutation_commitments) + " } }")
