"""GWC multi-open prover (the plonk paper's batched KZG opening): halo2_backend/src/poly/kzg/multiopen/gwc.rs:26-50
(`construct_intermediate_sets`: queries grouped by point, first-seen order) and gwc/prover.rs:45-90: for every distinct point z,
W_z = commit( (sum_i v^i p_i(X) - sum_i v^i p_i(z)) / (X - z) ).  One linear combination, one kate_division and one MSM per point,
all on the device."""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

from .. import fields as F
from . import polyexpr as PX

R = F.R_MOD


def construct_intermediate_sets(queries: Sequence[tuple]) -> List[Tuple[int, List[tuple]]]:
    """gwc.rs:26-50: [(point, [queries at that point])] in first-seen order."""
    out: List[Tuple[int, List[tuple]]] = []
    index: Dict[int, int] = {}
    for q in queries:
        if q[0] not in index:
            index[q[0]] = len(out)
            out.append((q[0], []))
        out[index[q[0]]][1].append(q)
    return out


def create_proof(ops, transcript, queries: Sequence[tuple], known_evals: Optional[Dict[Tuple[int, int], int]] = None) -> None:
    """gwc/prover.rs:45-90.  `known_evals`: values the caller already has, keyed (id(polynomial), point); the rest are evaluated
    in one batched device call."""
    cache: Dict[Tuple[int, int], int] = dict(known_evals or {})
    missing, seen = [], set()
    for q in queries:
        key = (id(q[1]), q[0])
        if key not in cache and key not in seen:
            seen.add(key)
            missing.append((q[1], q[0]))
    for (poly, point), value in zip(missing, ops.eval_polynomials(missing)):
        cache[(id(poly), point)] = value
    v = transcript.squeeze_challenge_scalar()
    for z, qs in construct_intermediate_sets(queries):
        terms, eval_batch, power = [], 0, 1
        for q in qs:
            assert q[0] == z
            terms.append((power, q[1]))
            eval_batch = (eval_batch + power * cache[(id(q[1]), z)]) % R
            power = power * v % R
        poly_batch = PX.lincomb(ops, terms)
        head = ops.read_ints(poly_batch, 0, 1)[0]
        ops.write_ints(poly_batch, 0, [(head - eval_batch) % R])
        transcript.write_point(ops.commit(ops.kate_division(poly_batch, z)))
