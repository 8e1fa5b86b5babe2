"""The independent prover oracle (oracle/plonk_prover.py) against what the reference itself holds:

* the `{:#?}` text of a pinned verifying key in halo2_proofs/tests/plonk_api.rs:664-1141 -- the only golden the reference has for
  the string the transcript is seeded with.  The oracle's Debug model must re-render it exactly, and `pinned_vk_node` must rebuild
  it from nothing but the circuit's expressions (so the QueryBack indices, i.e. the query collection order of keygen.rs:311-341,
  are pinned by a real reference output);
* the RNG streams against the committed known answers (tests/golden/vectors.json, rand_chacha / rand_core).
No GPU needed."""
import hashlib
import json
import os
import re
import types

import pytest

from oracle import plonk_prover as PP
from oracle import pyref as R

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_GOLDEN = "/root/reference/halo2_proofs/tests/plonk_api.rs"
# sha256 of the golden string as extracted below: the committed structural digest (tests/golden/pinned_vk_debug.json) was made from it
GOLDEN_SHA256 = None


def reference_golden():
    if not os.path.exists(REFERENCE_GOLDEN):
        pytest.skip("the reference tree is not on this machine")
    src = open(REFERENCE_GOLDEN).read()
    return re.search(r'r#"(PinnedVerificationKey \{.*?\n\})"#', src, re.S).group(1)


def field(node, name):
    assert node[0] == "struct"
    return dict(node[2])[name]


def expr_of(node):
    """ExpressionBack Debug tree -> the oracle's expression tuples."""
    kind, name, items = node
    if name == "Var":
        inner = items[0]
        if inner[1] == "Query":
            q = inner[2][0]
            return ("query", field(q, "column_type")[1].lower(), int(field(q, "column_index")[1]), int(field(q, "rotation")[2][0][1]))
        ch = inner[2][0]
        return ("challenge", int(field(ch, "index")[1]))
    if name == "Constant":
        return ("const", int(items[0][1], 16))
    if name == "Negated":
        return ("neg", expr_of(items[0]))
    return ({"Sum": "sum", "Product": "prod"}[name], expr_of(items[0]), expr_of(items[1]))


def point_of(node):
    m = re.fullmatch(r"\((0x[0-9a-f]+), (0x[0-9a-f]+)\)", node[1])
    return (int(m.group(1), 16), int(m.group(2), 16))


def test_debug_model_rerenders_the_reference_golden():
    golden = reference_golden()
    tree = PP.parse_debug(golden)
    assert PP.render(tree, pretty=True) == golden
    compact = PP.render(tree)
    assert PP.parse_debug(compact) == tree and "\n" not in compact
    assert compact.startswith('PinnedVerificationKey { base_modulus: "0x4000')


def test_pinned_vk_is_rebuilt_from_the_circuit_alone():
    """Strip the golden down to what a circuit description holds (expressions without query indices, columns, lookups), run it through
    the oracle's ConstraintSystem + pinned_vk_node, and get the reference's text back -- query indices, field order and all."""
    golden = reference_golden()
    tree = PP.parse_debug(golden)
    cs_node = field(tree, "cs")
    column = lambda n: (field(n, "column_type")[1].lower(), int(field(n, "index")[1]))
    md = field(cs_node, "minimum_degree")
    cs = {
        "num_fixed_columns": int(field(cs_node, "num_fixed_columns")[1]), "num_advice_columns": int(field(cs_node, "num_advice_columns")[1]),
        "num_instance_columns": int(field(cs_node, "num_instance_columns")[1]),
        "gates": [("", expr_of(g)) for g in field(cs_node, "gates")[1]],
        "permutation_columns": [column(c) for c in field(field(cs_node, "permutation"), "columns")[1]],
        "lookups": [(json.loads(field(l, "name")[1]), [expr_of(e) for e in field(l, "input_expressions")[1]],
                     [expr_of(e) for e in field(l, "table_expressions")[1]]) for l in field(cs_node, "lookups")[1]],
        "minimum_degree": None if md[1] == "None" else int(md[2][0][1]),
    }
    dom = field(tree, "domain")
    domain = types.SimpleNamespace(k=int(field(dom, "k")[1]), extended_k=int(field(dom, "extended_k")[1]), omega=int(field(dom, "omega")[1], 16))
    system = PP.ConstraintSystem(cs)
    rebuilt = PP.pinned_vk_node(system, domain, [point_of(p) for p in field(tree, "fixed_commitments")[1]],
                                [point_of(p) for p in field(field(tree, "permutation"), "commitments")[1]],
                                base_modulus=int(json.loads(field(tree, "base_modulus")[1]), 16),
                                scalar_modulus=int(json.loads(field(tree, "scalar_modulus")[1]), 16))
    assert PP.render(rebuilt, pretty=True) == golden
    # the degree rule gives the golden's extended_k as well (circuit.rs:80-117, domain.rs:48-51)
    ext = domain.k
    while (1 << ext) < (1 << domain.k) * (system.degree() - 1):
        ext += 1
    assert ext == domain.extended_k
    # and the query lists the golden prints are the ones the oracle collected
    for kind in ("advice", "instance", "fixed"):
        listed = [(column(q[2][0])[1], int(q[2][1][2][0][1])) for q in field(cs_node, kind + "_queries")[1]]
        assert listed == system.queries[kind]


@pytest.mark.parametrize("k,degree", [(3, 3), (4, 5), (4, 6), (5, 9)])
def test_quotient_by_coset_identity(k, degree):
    """The algebra plonk/cosets.py rests on, in python ints (no GPU): the extended coset of domain.rs:230-244 is the union of
    e = 2^(extended_k - k) cosets s_j * H of the size-n domain (row e * r + j of the extended vector is row r of coset j), and the
    n-sized pieces h_q of a polynomial of degree < n * e come back from the per-coset interpolants c_j by
        h_q = zeta^(-n q) / e * sum_j eps^(-j q) c_j,      eps = w_e^n."""
    import random
    P = R.R_MOD
    dom = R.EvaluationDomain(degree, k)
    n, e = dom.n, 1 << (dom.extended_k - dom.k)
    rnd = random.Random(1000 * k + degree)
    pieces = [[rnd.randrange(P) for _ in range(n)] for _ in range(dom.quotient_poly_degree)]
    h = [c for piece in pieces for c in piece]                                    # degree < n * (d - 1) <= n * e
    ext = [R.eval_polynomial(h, R.FR_ZETA * pow(dom.extended_omega, i, P) % P) for i in range(n * e)]
    # the vanishing polynomial is constant on a coset: t_evaluations[j] = 1 / (X^n - 1) there (domain.rs:86-111)
    for i in range(n * e):
        x = R.FR_ZETA * pow(dom.extended_omega, i, P) % P
        assert (pow(x, n, P) - 1) * dom.t_evaluations[i % e] % P == 1
    eps_inv = pow(pow(dom.extended_omega, n, P), -1, P)
    zeta_n_inv = pow(pow(R.FR_ZETA, n, P), -1, P)
    c = []
    for j in range(e):
        s_j = R.FR_ZETA * pow(dom.extended_omega, j, P) % P
        values = ext[j::e]                                                        # coset j, rows r = 0 .. n-1
        assert values == [R.eval_polynomial(h, s_j * pow(dom.omega, r, P) % P) for r in range(n)]
        coeff = dom.lagrange_to_coeff(values)                                     # interpolant in X / s_j
        s_inv = pow(s_j, -1, P)
        c.append([v * pow(s_inv, t, P) % P for t, v in enumerate(coeff)])
    e_inv = pow(e, -1, P)
    for q, piece in enumerate(pieces):
        got = [pow(zeta_n_inv, q, P) * e_inv % P * sum(pow(eps_inv, j * q, P) * c[j][t] for j in range(e)) % P for t in range(n)]
        assert got == piece
    # and a column on coset j is the size-n transform of its coefficients scaled by s_j^t (rotations stay inside the coset)
    a = [rnd.randrange(P) for _ in range(n)]
    ext_a = dom.coeff_to_extended(a)
    for j in range(e):
        s_j = R.FR_ZETA * pow(dom.extended_omega, j, P) % P
        scaled = [v * pow(s_j, t, P) % P for t, v in enumerate(a)]
        assert R.best_fft(scaled, dom.omega, k) == ext_a[j::e]
