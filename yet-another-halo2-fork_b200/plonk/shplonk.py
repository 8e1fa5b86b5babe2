"""SHPLONK multi-open prover: halo2_backend/src/poly/kzg/multiopen/shplonk.rs:46-137 (`construct_intermediate_sets`) and
shplonk/prover.rs:104-306 (`ProverSHPLONK::create_proof_with_engine`).

Host side: grouping queries into rotation sets, the low-degree interpolants r_i(X) (at most a handful of points each) and the
scalar bookkeeping.  Device side: every size-n linear combination, the divisions by (X - point) (h2b_kate_division), the two
commitments (MSM) and the evaluations `get_eval` takes (h2b_eval_polynomial).
A query is (point, poly, blind); polynomials are compared by identity like `PolynomialPointer` (poly/query.rs:49-53).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

from .. import fields as F
from . import polyexpr as PX

R = F.R_MOD


def lagrange_interpolate(points: Sequence[int], evals: Sequence[int]) -> List[int]:
    """arithmetic.rs:177-232: coefficients (lowest first) of the unique polynomial of degree < len(points) through the pairs."""
    assert len(points) == len(evals)
    if len(points) == 1:
        return [evals[0] % R]
    final = [0] * len(points)
    for j, x_j in enumerate(points):
        tmp = [1]
        for k, x_k in enumerate(points):
            if k == j:
                continue
            denom = pow((x_j - x_k) % R, -1, R)
            a, b = tmp + [0], [0] + tmp
            tmp = [(u * (-denom * x_k) + v * denom) % R for u, v in zip(a, b)]
        for i, c in enumerate(tmp):
            final[i] = (final[i] + c * evals[j]) % R
    return final


def evaluate_vanishing_polynomial(roots: Sequence[int], z: int) -> int:
    """arithmetic.rs:234-256."""
    acc = 1
    for p in roots:
        acc = (z - p) * acc % R
    return acc


def eval_small(coeffs: Sequence[int], x: int) -> int:
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * x + c) % R
    return acc


class RotationSet:
    def __init__(self, points: List[int]):
        self.points = points                                    # ascending canonical order (BTreeSet<F>)
        self.commitments: List[Tuple[object, List[int]]] = []   # (commitment, evals at self.points)


def construct_intermediate_sets(queries: Sequence[tuple], get_commitment, get_eval):
    """shplonk.rs:46-137.  `get_commitment(query)` returns a hashable identity; `get_eval(query)` its evaluation.  Commitments keep
    first-seen order, rotation sets the order of the first commitment that has them, points inside a set ascend (BTreeSet)."""
    evals: Dict[Tuple[object, int], int] = {}
    super_point_set = set()
    commitment_points: Dict[object, set] = {}
    order: List[object] = []
    handle: Dict[object, tuple] = {}
    for q in queries:
        c, point = get_commitment(q), q[0]
        if (c, point) not in evals:                             # `find` returns the first matching query
            evals[(c, point)] = get_eval(q)
        super_point_set.add(point)
        if c not in commitment_points:
            commitment_points[c] = set()
            order.append(c)
            handle[c] = q
        commitment_points[c].add(point)
    sets: List[RotationSet] = []
    by_points: Dict[tuple, RotationSet] = {}
    for c in order:
        key = tuple(sorted(commitment_points[c]))
        if key not in by_points:
            by_points[key] = RotationSet(list(key))
            sets.append(by_points[key])
        by_points[key].commitments.append((handle[c], [evals[(c, p)] for p in key]))
    return sets, sorted(super_point_set)


def create_proof(ops, transcript, queries: Sequence[tuple], known_evals: Optional[Dict[Tuple[int, int], int]] = None) -> None:
    """shplonk/prover.rs:117-305.  Writes the two opening commitments h and h' to the transcript.  `known_evals` holds values
    the caller already has, keyed (id(polynomial), point); the remaining queries are evaluated in one batched device call."""
    n = ops.n
    cache: Dict[Tuple[int, int], int] = dict(known_evals or {})
    missing, seen = [], set()
    for q in queries:
        key = (id(q[1]), q[0])
        if key not in cache and key not in seen:
            seen.add(key)
            missing.append((q[1], q[0]))
    for (poly, point), value in zip(missing, ops.eval_polynomials(missing)):
        cache[(id(poly), point)] = value

    def get_eval(q):                                            # poly/query.rs:62-64
        return cache[(id(q[1]), q[0])]

    y = transcript.squeeze_challenge_scalar()
    rotation_sets, super_point_set = construct_intermediate_sets(queries, lambda q: id(q[1]), get_eval)
    # CommitmentExtension: the low degree equivalent r_i(X) of every opened polynomial on its rotation set (prover.rs:39-53)
    extended = [[(q[1], lagrange_interpolate(rs.points, evals)) for q, evals in rs.commitments] for rs in rotation_sets]
    v = transcript.squeeze_challenge_scalar()

    # quotient_contribution (prover.rs:133-160): sum_j y^j (p_j - r_j), divided by prod (X - point) over the set, padded to n
    quotient_polynomials = []
    for rs, commitments in zip(rotation_sets, extended):
        terms, low, power = [], [0] * len(rs.points), 1
        for poly, r in commitments:
            terms.append((power, poly))
            for i, c in enumerate(r):
                low[i] = (low[i] + power * c) % R
            power = power * y % R
        n_x = PX.lincomb(ops, terms)
        head = ops.read_ints(n_x, 0, len(low))
        ops.write_ints(n_x, 0, [(a - b) % R for a, b in zip(head, low)])
        for point in rs.points:                                 # div_by_vanishing, prover.rs:28-34
            n_x = ops.kate_division(n_x, point)
        quotient_polynomials.append(ops.resize(n_x, n))
    terms, power = [], 1
    for q_poly in quotient_polynomials:                         # prover.rs:190-195
        terms.append((power, q_poly))
        power = power * v % R
    h_x = PX.lincomb(ops, terms)
    transcript.write_point(ops.commit(h_x))
    u = transcript.squeeze_challenge_scalar()

    # linearisation (prover.rs:203-260): l(X) = sum_i v^i z_i sum_j y^j (p_ij(X) - r_ij(u)) - Z_T(u) h(X), which vanishes at u
    terms, constant, z_diffs, power_v = [], 0, [], 1
    for rs, commitments in zip(rotation_sets, extended):
        diffs = [p for p in super_point_set if p not in rs.points]
        z_i = evaluate_vanishing_polynomial(diffs, u)
        z_diffs.append(z_i)
        power_y = 1
        for poly, r in commitments:
            w = power_v * z_i % R * power_y % R
            terms.append((w, poly))
            constant = (constant + w * eval_small(r, u)) % R
            power_y = power_y * y % R
        power_v = power_v * v % R
    zt_eval = evaluate_vanishing_polynomial(super_point_set, u)
    terms.append(((-zt_eval) % R, h_x))
    z_0_diff_inv = pow(z_diffs[0], -1, R)                       # prover.rs:282-285, folded into the coefficients (exact arithmetic)
    l_x = PX.lincomb(ops, [(w * z_0_diff_inv % R, p) for w, p in terms])
    head = ops.read_ints(l_x, 0, 1)[0]
    ops.write_ints(l_x, 0, [(head - constant * z_0_diff_inv) % R])
    h_x2 = ops.kate_division(l_x, u)                            # prover.rs:280
    transcript.write_point(ops.commit(h_x2))
