"""verify_proof for KZG / SHPLONK on python ints -- TEST INFRASTRUCTURE ONLY (the acceptance check of the prover mirror).

Restates halo2_backend/src/plonk/verifier.rs:49-511 (one circuit), the argument verifiers it calls
(permutation/verifier.rs:100-253, lookup/verifier.rs:90-212, shuffle/verifier.rs:55-142, vanishing/verifier.rs:40-137),
shplonk/verifier.rs:47-139 and the final `DualMSM::check` (kzg/msm.rs:188-206):  e(left, [s]G2) == e(right, G2).
There is no pairing library here, so the check uses the SRS trapdoor: with the test secret s known, the pairing equation holds
iff  s * left == right  in G1.
"""
from __future__ import annotations

import importlib.util
import os
import sys
from typing import List, Sequence

from . import pyref as R

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import h2b200  # noqa: E402  (host-side mirrors only: transcript, circuit model, SHPLONK set construction)
from h2b200.plonk import circuit as C  # noqa: E402
from h2b200.plonk import shplonk as SH  # noqa: E402
from h2b200.plonk.keygen import FR_DELTA  # noqa: E402
from h2b200.plonk.transcript import Blake2bRead  # noqa: E402

P = R.R_MOD


class VerifyError(Exception):
    pass


def _inv(x: int) -> int:
    return pow(x % P, -1, P)


def l_i_range(domain, x: int, xn: int, rotations: Sequence[int]) -> List[int]:
    """poly/domain.rs:425-450."""
    results = [_inv(x - domain.rotate_omega(1, rot)) for rot in rotations]
    common = (xn - 1) * domain.barycentric_weight % P
    return [domain.rotate_omega(r * common % P, rot) for rot, r in zip(rotations, results)]


class _Msm:
    """MSMKZG (kzg/msm.rs:17-100): a list of (scalar, point) terms."""

    def __init__(self, terms=None):
        self.terms = list(terms or [])

    def scale(self, f: int):
        self.terms = [(s * f % P, p) for s, p in self.terms]

    def append_term(self, s: int, p):
        self.terms.append((s % P, p))

    def add_msm(self, other: "_Msm"):
        self.terms += other.terms

    def eval(self):
        acc = None
        for s, p in self.terms:
            acc = R.g1_add(acc, R.g1_mul(p, s))
        return acc


def verify_proof(vk, instances: Sequence[Sequence[int]], proof: bytes, s: int, n: int, debug: dict = None,
                 multiopen: str = "shplonk") -> bool:
    """verify_proof_single (plonk/verifier.rs:29-47): one circuit."""
    return verify_proof_batch(vk, [instances], proof, s, n, debug, multiopen)


def verify_proof_batch(vk, circuits_instances: Sequence[Sequence[Sequence[int]]], proof: bytes, s: int, n: int, debug: dict = None,
                       multiopen: str = "shplonk") -> bool:
    """verify_proof (plonk/verifier.rs:49-511) for `num_proofs` instances of one circuit proven in one transcript."""
    cs, domain = vk.cs, vk.domain
    for instances in circuits_instances:
        if len(instances) != cs.num_instance_columns:
            raise VerifyError("InvalidInstances")
    num_proofs = len(circuits_instances)
    proofs = range(num_proofs)
    t = Blake2bRead(proof)
    vk.hash_into(t)
    for instances in circuits_instances:
        for column in instances:
            for v in column:
                t.common_scalar(v)
    advice_commitments = [[None] * cs.num_advice_columns for _ in proofs]
    challenges = [0] * cs.num_challenges
    for phase in cs.phases():                                                       # verifier.rs:151-176
        for c in proofs:
            for i, p in enumerate(cs.advice_column_phase):
                if p == phase:
                    advice_commitments[c][i] = t.read_point()
        for i, p in enumerate(cs.challenge_phase):
            if p == phase:
                challenges[i] = t.squeeze_challenge_scalar()
    theta = t.squeeze_challenge_scalar()
    lookups = [[{"permuted_input": t.read_point(), "permuted_table": t.read_point()} for _ in cs.lookups] for _ in proofs]
    beta = t.squeeze_challenge_scalar()
    gamma = t.squeeze_challenge_scalar()
    chunk_len = vk.cs_degree - 2
    n_sets = (len(cs.permutation_columns) + chunk_len - 1) // chunk_len
    perm_sets = [[{"commitment": t.read_point()} for _ in range(n_sets)] for _ in proofs]
    for c in proofs:
        for l in lookups[c]:
            l["product"] = t.read_point()
    shuffles = [[{"product": t.read_point()} for _ in cs.shuffles] for _ in proofs]
    random_poly_commitment = t.read_point()
    y = t.squeeze_challenge_scalar()
    h_commitments = [t.read_point() for _ in range(domain.quotient_poly_degree)]
    x = t.squeeze_challenge_scalar()
    xn = pow(x, n, P)
    # instance evals by Lagrange interpolation (QUERY_INSTANCE == false), verifier.rs:243-279
    min_rot = min([0] + [rot for _, rot in cs.instance_queries])
    max_rot = max([0] + [rot for _, rot in cs.instance_queries])
    max_len = max([len(col) for instances in circuits_instances for col in instances], default=0)
    l_i_s = l_i_range(domain, x, xn, list(range(-max_rot, max_len + abs(min_rot))))
    instance_evals = []
    for instances in circuits_instances:
        evals_c = []
        for column, rot in cs.instance_queries:
            vals = instances[column]
            offset = max_rot - rot
            evals_c.append(sum(a * b for a, b in zip(vals, l_i_s[offset:offset + len(vals)])) % P)
        instance_evals.append(evals_c)
    advice_evals = [[t.read_scalar() for _ in cs.advice_queries] for _ in proofs]
    fixed_evals = [t.read_scalar() for _ in cs.fixed_queries]
    random_eval = t.read_scalar()
    permutation_evals = [t.read_scalar() for _ in vk.permutation_commitments]
    for c in proofs:
        for i, st in enumerate(perm_sets[c]):                                       # permutation/verifier.rs:66-96
            st["eval"], st["next_eval"] = t.read_scalar(), t.read_scalar()
            st["last_eval"] = t.read_scalar() if i + 1 < len(perm_sets[c]) else None
    for c in proofs:
        for l in lookups[c]:                                                        # lookup/verifier.rs:62-86
            l["product_eval"], l["product_next_eval"] = t.read_scalar(), t.read_scalar()
            l["permuted_input_eval"], l["permuted_input_inv_eval"], l["permuted_table_eval"] = (t.read_scalar(), t.read_scalar(),
                                                                                                t.read_scalar())
    for c in proofs:
        for sh in shuffles[c]:
            sh["product_eval"], sh["product_next_eval"] = t.read_scalar(), t.read_scalar()

    bf = cs.blinding_factors()
    l_evals = l_i_range(domain, x, xn, list(range(-(bf + 1), 1)))
    assert len(l_evals) == 2 + bf
    l_last, l_blind, l_0 = l_evals[0], sum(l_evals[1:1 + bf]) % P, l_evals[1 + bf]
    active = (1 - (l_last + l_blind)) % P
    expressions = []
    for c in proofs:
        ev = lambda e: C.evaluate_at(e, fixed_evals, advice_evals[c], instance_evals[c], challenges, cs.query_index)
        compress = lambda exprs: _fold(theta, [ev(e) for e in exprs])
        expressions += [ev(poly) for _, poly in cs.gates]
        sets = perm_sets[c]
        if sets:                                                                    # permutation/verifier.rs:100-194
            expressions.append(l_0 * (1 - sets[0]["eval"]) % P)
            last = sets[-1]["eval"]
            expressions.append((last * last - last) * l_last % P)
            for prev, cur in zip(sets, sets[1:]):
                expressions.append((cur["eval"] - prev["last_eval"]) * l_0 % P)
            evals_of = {"advice": advice_evals[c], "fixed": fixed_evals, "instance": instance_evals[c]}
            for chunk_index, st in enumerate(sets):
                cols = cs.permutation_columns[chunk_index * chunk_len:(chunk_index + 1) * chunk_len]
                pevals = permutation_evals[chunk_index * chunk_len:(chunk_index + 1) * chunk_len]
                col_evals = [evals_of[kind][cs.get_any_query_index(kind, index, 0)] for kind, index in cols]
                left = st["next_eval"]
                for e, pe in zip(col_evals, pevals):
                    left = left * (e + beta * pe + gamma) % P
                right = st["eval"]
                current_delta = beta * x % P * pow(FR_DELTA, chunk_index * chunk_len, P) % P
                for e in col_evals:
                    right = right * (e + current_delta + gamma) % P
                    current_delta = current_delta * FR_DELTA % P
                expressions.append((left - right) * active % P)
        for l, arg in zip(lookups[c], cs.lookups):                                  # lookup/verifier.rs:90-163
            left = l["product_next_eval"] * (l["permuted_input_eval"] + beta) % P * (l["permuted_table_eval"] + gamma) % P
            right = l["product_eval"] * (compress(arg.input_expressions) + beta) % P * (compress(arg.table_expressions) + gamma) % P
            a_minus_s = (l["permuted_input_eval"] - l["permuted_table_eval"]) % P
            expressions += [l_0 * (1 - l["product_eval"]) % P,
                            l_last * (l["product_eval"] ** 2 - l["product_eval"]) % P,
                            (left - right) * active % P,
                            l_0 * a_minus_s % P,
                            a_minus_s * (l["permuted_input_eval"] - l["permuted_input_inv_eval"]) % P * active % P]
        for sh, arg in zip(shuffles[c], cs.shuffles):                               # shuffle/verifier.rs:55-115
            left = sh["product_next_eval"] * (compress(arg.shuffle_expressions) + gamma) % P
            right = sh["product_eval"] * (compress(arg.input_expressions) + gamma) % P
            expressions += [l_0 * (1 - sh["product_eval"]) % P,
                            l_last * (sh["product_eval"] ** 2 - sh["product_eval"]) % P,
                            (left - right) * active % P]
    expected_h_eval = _fold(y, expressions) * _inv(xn - 1) % P                      # vanishing/verifier.rs:88-96
    if debug is not None:
        debug.update(x=x, y=y, theta=theta, beta=beta, gamma=gamma, expected_h_eval=expected_h_eval, expressions=expressions,
                     challenges=challenges)
    h_commitment = _Msm()
    for cm in reversed(h_commitments):
        h_commitment.scale(xn)
        h_commitment.append_term(1, cm)

    # the verifier queries (commitment handle, point, eval), verifier.rs:437-499
    rot = domain.rotate_omega
    queries = []
    x_next, x_last, x_inv = rot(x, 1), rot(x, -(bf + 1)), rot(x, -1)
    for c in proofs:
        for qi, (column, at) in enumerate(cs.advice_queries):
            queries.append((rot(x, at), ("advice", c, column), advice_evals[c][qi]))
        for i, st in enumerate(perm_sets[c]):
            queries.append((x, ("perm_z", c, i), st["eval"]))
            queries.append((x_next, ("perm_z", c, i), st["next_eval"]))
        for i in reversed(range(len(perm_sets[c]) - 1)):
            queries.append((x_last, ("perm_z", c, i), perm_sets[c][i]["last_eval"]))
        for i, l in enumerate(lookups[c]):
            queries += [(x, ("lk_z", c, i), l["product_eval"]), (x, ("lk_in", c, i), l["permuted_input_eval"]),
                        (x, ("lk_tab", c, i), l["permuted_table_eval"]), (x_inv, ("lk_in", c, i), l["permuted_input_inv_eval"]),
                        (x_next, ("lk_z", c, i), l["product_next_eval"])]
        for i, sh in enumerate(shuffles[c]):
            queries += [(x, ("sh_z", c, i), sh["product_eval"]), (x_next, ("sh_z", c, i), sh["product_next_eval"])]
    for qi, (column, at) in enumerate(cs.fixed_queries):
        queries.append((rot(x, at), ("fixed", 0, column), fixed_evals[qi]))
    for i, e in enumerate(permutation_evals):
        queries.append((x, ("sigma", 0, i), e))
    queries.append((x, ("h",), expected_h_eval))
    queries.append((x, ("random",), random_eval))

    def commitment_msm(handle) -> _Msm:
        kind = handle[0]
        if kind == "h":
            return _Msm(h_commitment.terms)
        if kind == "random":
            return _Msm([(1, random_poly_commitment)])
        c, i = handle[1], handle[2]
        point = {"advice": lambda: advice_commitments[c][i], "perm_z": lambda: perm_sets[c][i]["commitment"],
                 "lk_z": lambda: lookups[c][i]["product"], "lk_in": lambda: lookups[c][i]["permuted_input"],
                 "lk_tab": lambda: lookups[c][i]["permuted_table"], "sh_z": lambda: shuffles[c][i]["product"],
                 "fixed": lambda: vk.fixed_commitments[i], "sigma": lambda: vk.permutation_commitments[i]}[kind]()
        return _Msm([(1, point)])

    if multiopen == "gwc":
        return _verify_gwc(t, queries, commitment_msm, s)
    # shplonk/verifier.rs:47-139
    rotation_sets, super_point_set = SH.construct_intermediate_sets(queries, lambda q: q[1], lambda q: q[2])
    y_s = t.squeeze_challenge_scalar()
    v = t.squeeze_challenge_scalar()
    h1 = t.read_point()
    u = t.squeeze_challenge_scalar()
    h2 = t.read_point()
    if t.remaining():
        raise VerifyError("trailing bytes in the proof")
    z_0_diff_inverse = z_0 = 0
    outer, r_outer_acc, power_of_v = _Msm(), 0, 1
    for i, rs in enumerate(rotation_sets):
        diffs = [p for p in super_point_set if p not in rs.points]
        z_diff_i = SH.evaluate_vanishing_polynomial(diffs, u)
        if i == 0:
            z_0 = SH.evaluate_vanishing_polynomial(rs.points, u)
            z_0_diff_inverse = _inv(z_diff_i)
            z_diff_i = 1
        else:
            z_diff_i = z_diff_i * z_0_diff_inverse % P
        inner, r_inner_acc, power_of_y = _Msm(), 0, 1
        for q, evals in rs.commitments:
            r_x = SH.lagrange_interpolate(rs.points, evals)
            r_inner_acc = (r_inner_acc + power_of_y * SH.eval_small(r_x, u)) % P
            m = commitment_msm(q[1])
            m.scale(power_of_y)
            inner.add_msm(m)
            power_of_y = power_of_y * y_s % P
        inner.scale(power_of_v * z_diff_i % P)
        outer.add_msm(inner)
        r_outer_acc = (r_outer_acc + power_of_v * r_inner_acc % P * z_diff_i) % P
        power_of_v = power_of_v * v % P
    outer.append_term(-r_outer_acc, R.G1_GEN)                                       # params.g[0] = G
    outer.append_term(-z_0, h1)
    outer.append_term(u, h2)
    left, right = h2, outer.eval()
    return R.g1_mul(left, s % P) == right                                           # kzg/msm.rs:188-206 via the trapdoor


def _verify_gwc(t, queries, commitment_msm, s: int) -> bool:
    """gwc/verifier.rs:40-123 + DualMSM::check through the trapdoor: s * left == right."""
    from h2b200.plonk import gwc
    v = t.squeeze_challenge_scalar()
    commitment_data = gwc.construct_intermediate_sets(queries)
    w = [t.read_point() for _ in commitment_data]
    u = t.squeeze_challenge_scalar()
    if t.remaining():
        raise VerifyError("trailing bytes in the proof")
    commitment_multi, eval_multi, witness, witness_with_aux, power_of_u = _Msm(), 0, _Msm(), _Msm(), 1
    for (z, qs), wi in zip(commitment_data, w):
        batch, eval_batch, power_of_v = _Msm(), 0, 1
        for q in qs:
            m = commitment_msm(q[1])
            m.scale(power_of_v)
            batch.add_msm(m)
            eval_batch = (eval_batch + power_of_v * q[2]) % P
            power_of_v = power_of_v * v % P
        batch.scale(power_of_u)
        commitment_multi.add_msm(batch)
        eval_multi = (eval_multi + power_of_u * eval_batch) % P
        witness_with_aux.append_term(power_of_u * z, wi)
        witness.append_term(power_of_u, wi)
        power_of_u = power_of_u * u % P
    right = _Msm()
    right.add_msm(witness_with_aux)
    right.add_msm(commitment_multi)
    right.append_term(eval_multi, R.g1_neg(R.G1_GEN))
    return R.g1_mul(witness.eval(), s % P) == right.eval()


def _fold(factor: int, values: Sequence[int]) -> int:
    acc = 0
    for v in values:
        acc = (acc * factor + v) % P
    return acc
