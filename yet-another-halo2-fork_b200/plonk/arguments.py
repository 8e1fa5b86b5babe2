"""The four argument provers create_proof is assembled from, each a restatement of the reference's host orchestration with the
size-n vector work issued through `ops` (device kernels):

    lookup_commit_permuted / Permuted.commit_product / evaluate / open     plonk/lookup/prover.rs
    permutation_commit / evaluate / open                                   plonk/permutation/prover.rs
    shuffle_commit_product / evaluate / open                               plonk/shuffle/prover.rs
    vanishing commit / construct / evaluate / open                         plonk/vanishing/prover.rs

RNG draws happen in exactly the reference's order: that, and nothing else, fixes the proof bytes.
A ProverQuery is (point, poly, blind) as in poly/query.rs:20-25; polynomials are compared by identity (query.rs:49-53).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from .. import fields as F
from ..evaluation import CONSTANT, THETA, GraphEvaluator
from . import circuit as C
from . import polyexpr as PX
from .keygen import FR_DELTA

R = F.R_MOD


class EvalBatch:
    """The evaluations create_proof writes after the challenge x (plonk/prover.rs:742-811, and the `evaluate` methods of the argument
    provers) depend on nothing the transcript produces in between: the `*_evaluate` functions record their (polynomial, point)
    pairs here in transcript order, `flush` evaluates them in ONE device call (h2b_eval_polynomial_batch_device) and writes the
    scalars in that order.  The values are kept by (id(polynomial), point) so that the multi-open argument, which the reference
    lets evaluate the same pairs again (poly/query.rs:62-64), can reuse them."""

    def __init__(self, ops):
        self.ops, self.pairs, self.known = ops, [], {}

    def add(self, poly, point: int) -> None:
        self.pairs.append((poly, point))

    def flush(self, transcript) -> dict:
        values = self.ops.eval_polynomials(self.pairs)
        for (poly, point), value in zip(self.pairs, values):
            transcript.write_scalar(value)
            self.known[(id(poly), point)] = value
        self.pairs = []
        return self.known


class ConstraintSystemFailure(Exception):
    """plonk::Error::ConstraintSystemFailure."""


def _draw(rng, count: int) -> List[int]:
    return [rng.random_fr() for _ in range(count)]


def compress_expressions(ops, expressions: Sequence, theta: int, pk, advice_values, instance_values, challenges):
    """lookup/prover.rs:81-100 == shuffle/prover.rs:56-75: fold(0, |acc, e| acc * theta + evaluate(e, n, 1, ...))."""
    g = GraphEvaluator()
    parts = tuple(g.add_expression(e) for e in expressions)
    g.add_calculation(("horner", (CONSTANT, 0), parts, (THETA,)))
    return ops.graph(g, pk.fixed_values, advice_values, instance_values, challenges, 0, 0, theta, 0, extended=False)


# ---------------------------------------------------------------------------------------------------------------------------
# lookup
# ---------------------------------------------------------------------------------------------------------------------------
def permute_expression_pair(inp: np.ndarray, tab: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """lookup/prover.rs:398-494 on the usable rows, vectorised.  `inp` / `tab` are (u, 4) uint64 canonical little-endian limbs
    (`Fr: Ord` compares canonical values).  Returns the sorted input column and the table column arranged so that every first
    occurrence of an input value sits beside the same table value; the left-over table values fill the repeated rows from the last
    repeated row backwards in ascending value order (`repeated_input_rows.pop()` while iterating the BTreeMap)."""
    u = inp.shape[0]
    both = np.concatenate([inp, tab], axis=0)
    order = np.lexsort((both[:, 0], both[:, 1], both[:, 2], both[:, 3]))
    srt = both[order]
    new_group = np.ones(2 * u, dtype=bool)
    if u:
        new_group[1:] = np.any(srt[1:] != srt[:-1], axis=1)
    gid_sorted = np.cumsum(new_group) - 1
    rank = np.empty(2 * u, dtype=np.int64)
    rank[order] = gid_sorted
    values = srt[new_group]                                   # value of each rank, ascending
    sorted_in = np.sort(rank[:u])
    first = np.ones(u, dtype=bool)
    if u:
        first[1:] = sorted_in[1:] != sorted_in[:-1]
    counts = np.bincount(rank[u:], minlength=values.shape[0])
    distinct = sorted_in[first]
    if np.any(counts[distinct] == 0):
        raise ConstraintSystemFailure("lookup input value missing from the table")
    counts[distinct] -= 1
    leftover = np.repeat(np.arange(values.shape[0]), counts)
    repeated_rows = np.nonzero(~first)[0]
    assert leftover.shape[0] == repeated_rows.shape[0]
    table_rank = sorted_in.copy()
    table_rank[repeated_rows[::-1]] = leftover
    return values[sorted_in], values[table_rank]


class LookupPermuted:
    pass


def lookup_commit_permuted(ops, arg: C.LookupArgument, pk, theta: int, advice_values, instance_values, challenges, rng, transcript):
    """lookup/prover.rs:58-175."""
    cs, n = pk.vk.cs, ops.n
    bf = cs.blinding_factors()
    usable = n - (bf + 1)
    p = LookupPermuted()
    p.compressed_input_expression = compress_expressions(ops, arg.input_expressions, theta, pk, advice_values, instance_values,
                                                         challenges)
    p.compressed_table_expression = compress_expressions(ops, arg.table_expressions, theta, pk, advice_values, instance_values,
                                                         challenges)
    p.permuted_input_expression, p.permuted_table_expression = ops.permute_expression_pair(
        p.compressed_input_expression, p.compressed_table_expression, usable)
    in_tail, tab_tail = _draw(rng, bf + 1), _draw(rng, bf + 1)                     # lookup/prover.rs:469-471
    ops.write_ints(p.permuted_input_expression, usable, in_tail)
    ops.write_ints(p.permuted_table_expression, usable, tab_tail)

    # commit_values (lookup/prover.rs:129-135) for both columns; the two commit_lagrange MSMs share one batched pipeline (the
    # blinds are drawn in the reference's order and commitments do not consume randomness)
    p.permuted_input_poly = ops.lagrange_to_coeff(ops.clone(p.permuted_input_expression))
    p.permuted_input_blind = rng.random_fr()
    p.permuted_table_poly = ops.lagrange_to_coeff(ops.clone(p.permuted_table_expression))
    p.permuted_table_blind = rng.random_fr()
    c_in, c_tab = ops.commit_lagrange_many([p.permuted_input_expression, p.permuted_table_expression])
    transcript.write_point(c_in)
    transcript.write_point(c_tab)
    return p


class LookupCommitted:
    pass


def lookup_commit_product(ops, p: LookupPermuted, pk, beta: int, gamma: int, rng, transcript) -> LookupCommitted:
    """lookup/prover.rs:177-327."""
    n, bf = ops.n, pk.vk.cs.blinding_factors()
    b, g, col = C.const(beta), C.const(gamma), PX.col
    denom = PX.pointwise(ops, C.mul(C.add(b, col(0)), C.add(g, col(1))), [p.permuted_input_expression, p.permuted_table_expression])
    denom = ops.batch_invert(denom)
    product = PX.pointwise(ops, C.mul(C.mul(col(0), C.add(col(1), b)), C.add(col(2), g)),
                           [denom, p.compressed_input_expression, p.compressed_table_expression])
    z = ops.resize(ops.grand_product(product, 1, n - bf), n)
    ops.write_ints(z, n - bf, _draw(rng, bf))
    c = LookupCommitted()
    c.product_blind = rng.random_fr()
    commitment = ops.commit_lagrange(z)
    c.product_poly = ops.lagrange_to_coeff(z)
    transcript.write_point(commitment)
    c.permuted_input_poly, c.permuted_input_blind = p.permuted_input_poly, p.permuted_input_blind
    c.permuted_table_poly, c.permuted_table_blind = p.permuted_table_poly, p.permuted_table_blind
    return c


def lookup_evaluate(ops, c: LookupCommitted, domain, x: int, evals) -> None:
    """lookup/prover.rs:329-359."""
    x_inv, x_next = domain.rotate_omega(x, -1), domain.rotate_omega(x, 1)
    for poly, point in ((c.product_poly, x), (c.product_poly, x_next), (c.permuted_input_poly, x), (c.permuted_input_poly, x_inv),
                        (c.permuted_table_poly, x)):
        evals.add(poly, point)


def lookup_open(c: LookupCommitted, domain, x: int):
    """lookup/prover.rs:361-394."""
    x_inv, x_next = domain.rotate_omega(x, -1), domain.rotate_omega(x, 1)
    return [(x, c.product_poly, c.product_blind), (x, c.permuted_input_poly, c.permuted_input_blind),
            (x, c.permuted_table_poly, c.permuted_table_blind), (x_inv, c.permuted_input_poly, c.permuted_input_blind),
            (x_next, c.product_poly, c.product_blind)]


# ---------------------------------------------------------------------------------------------------------------------------
# permutation
# ---------------------------------------------------------------------------------------------------------------------------
class PermutationSet:
    pass


def permutation_commit(ops, pk, advice, instance, beta: int, gamma: int, rng, transcript) -> List[PermutationSet]:
    """permutation/prover.rs:47-190.  One grand product per chunk of cs_degree - 2 columns, chained through `last_z`."""
    cs, n, domain = pk.vk.cs, ops.n, pk.vk.domain
    assert pk.vk.cs_degree >= 3
    chunk_len = pk.vk.cs_degree - 2
    bf = cs.blinding_factors()
    columns = cs.permutation_columns
    sets: List[PermutationSet] = []
    if not columns:
        return sets
    source = {"advice": advice, "fixed": pk.fixed_values, "instance": instance}
    omega_powers = ops.grand_product(ops.full(n, domain.omega), 1, n)
    b, g, col = C.const(beta), C.const(gamma), PX.col
    last_z = 1
    for start in range(0, len(columns), chunk_len):
        chunk = columns[start:start + chunk_len]
        m = len(chunk)
        values = [source[kind][index] for kind, index in chunk]
        sigmas = pk.permutation.permutations[start:start + m]
        # denominator: prod (beta * sigma + gamma + value); numerator: prod (delta^c * omega^j * beta + gamma + value)
        den = None
        for i in range(m):
            t = C.add(C.add(C.mul(b, col(m + i)), g), col(i))
            den = t if den is None else C.mul(den, t)
        modified = ops.batch_invert(PX.pointwise(ops, den, values + sigmas))
        num = col(0)
        for i in range(m):
            dw = C.const(pow(FR_DELTA, start + i, R) * beta % R)
            num = C.mul(num, C.add(C.add(C.mul(dw, col(1)), g), col(2 + i)))
        modified = PX.pointwise(ops, num, [modified, omega_powers] + values)
        z = ops.grand_product(modified, last_z, n)
        ops.write_ints(z, n - bf, _draw(rng, bf))                                   # permutation/prover.rs:169-171
        last_z = ops.read_ints(z, n - (bf + 1), 1)[0]
        s = PermutationSet()
        s.permutation_product_blind = rng.random_fr()
        commitment = ops.commit_lagrange(z)
        s.permutation_product_poly = ops.lagrange_to_coeff(z)
        # evaluate_h reads the product on the extended coset (evaluation.rs:398-401); the coset prover extends it coset by coset
        s.permutation_product_coset = None if getattr(ops, "cosets", None) else ops.coeff_to_extended(s.permutation_product_poly)
        transcript.write_point(commitment)
        sets.append(s)
    return sets


def permutation_pk_evaluate(ops, pk, x: int, evals) -> None:
    """permutation/prover.rs:219-231."""
    for poly in pk.permutation.polys:
        evals.add(poly, x)


def permutation_evaluate(ops, sets: List[PermutationSet], pk, x: int, evals) -> None:
    """permutation/prover.rs:233-278."""
    domain, bf = pk.vk.domain, pk.vk.cs.blinding_factors()
    for i, s in enumerate(sets):
        evals.add(s.permutation_product_poly, x)
        evals.add(s.permutation_product_poly, domain.rotate_omega(x, 1))
        if i + 1 < len(sets):
            evals.add(s.permutation_product_poly, domain.rotate_omega(x, -(bf + 1)))


def permutation_open(sets: List[PermutationSet], pk, x: int):
    """permutation/prover.rs:280-333."""
    domain, bf = pk.vk.domain, pk.vk.cs.blinding_factors()
    x_next, x_last = domain.rotate_omega(x, 1), domain.rotate_omega(x, -(bf + 1))
    out = []
    for s in sets:
        out.append((x, s.permutation_product_poly, s.permutation_product_blind))
        out.append((x_next, s.permutation_product_poly, s.permutation_product_blind))
    for s in list(reversed(sets))[1:]:
        out.append((x_last, s.permutation_product_poly, s.permutation_product_blind))
    return out


def permutation_pk_open(pk, x: int):
    """permutation/prover.rs:206-217."""
    return [(x, poly, 0) for poly in pk.permutation.polys]


# ---------------------------------------------------------------------------------------------------------------------------
# shuffle
# ---------------------------------------------------------------------------------------------------------------------------
class ShuffleCommitted:
    pass


def shuffle_commit_product(ops, arg: C.ShuffleArgument, pk, theta: int, gamma: int, advice_values, instance_values, challenges, rng,
                           transcript) -> ShuffleCommitted:
    """shuffle/prover.rs:40-205."""
    n, bf = ops.n, pk.vk.cs.blinding_factors()
    input_expression = compress_expressions(ops, arg.input_expressions, theta, pk, advice_values, instance_values, challenges)
    shuffle_expression = compress_expressions(ops, arg.shuffle_expressions, theta, pk, advice_values, instance_values, challenges)
    g, col = C.const(gamma), PX.col
    product = ops.batch_invert(PX.pointwise(ops, C.add(g, col(0)), [shuffle_expression]))
    product = PX.pointwise(ops, C.mul(col(0), C.add(g, col(1))), [product, input_expression])
    z = ops.resize(ops.grand_product(product, 1, n - bf), n)
    ops.write_ints(z, n - bf, _draw(rng, bf))
    c = ShuffleCommitted()
    c.product_blind = rng.random_fr()
    commitment = ops.commit_lagrange(z)
    c.product_poly = ops.lagrange_to_coeff(z)
    transcript.write_point(commitment)
    return c


def shuffle_evaluate(ops, c: ShuffleCommitted, domain, x: int, evals) -> None:
    """shuffle/prover.rs:207-228."""
    evals.add(c.product_poly, x)
    evals.add(c.product_poly, domain.rotate_omega(x, 1))


def shuffle_open(c: ShuffleCommitted, domain, x: int):
    """shuffle/prover.rs:230-255."""
    return [(x, c.product_poly, c.product_blind), (domain.rotate_omega(x, 1), c.product_poly, c.product_blind)]


# ---------------------------------------------------------------------------------------------------------------------------
# vanishing
# ---------------------------------------------------------------------------------------------------------------------------
def parallelize_chunks(total: int, num_threads: int) -> List[Tuple[int, int]]:
    """(offset, length) of the chunks `parallelize` hands out (arithmetic.rs:148-170)."""
    base, cutoff = total // num_threads, total % num_threads
    split = cutoff * (base + 1)
    chunks = [(i * (base + 1), base + 1) for i in range(cutoff)]
    if base:
        chunks += [(split + i * base, base) for i in range((total - split) // base)]
    return chunks


class VanishingCommitted:
    pass


def vanishing_commit(ops, domain, rng, transcript, num_threads: int) -> VanishingCommitted:
    """vanishing/prover.rs:44-98: a uniformly random polynomial of degree n - 1, drawn by `num_threads` ChaCha20 streams seeded from
    `rng` in chunk order (`current_num_threads()` of the reference run: it changes the proof bytes, so it is an explicit parameter)."""
    n = 1 << domain.k
    chunk_size = n // num_threads
    offsets = [i * (chunk_size + 1) for i in range(n % num_threads)]
    if chunk_size:
        start = (n % num_threads) * (chunk_size + 1)
        offsets += [start + i * chunk_size for i in range(num_threads - len(offsets))]
    offsets = offsets[:num_threads]
    seeds = {off: rng.fill_bytes(32) for off in offsets}                            # vanishing/prover.rs:70-73, in iterator order
    chunks = parallelize_chunks(n, num_threads)
    c = VanishingCommitted()
    c.random_poly = ops.random_poly_chacha20([seeds[off] for off, _ in chunks], chunks)
    c.random_blind = rng.random_fr()
    transcript.write_point(ops.commit(c.random_poly))
    return c


class VanishingConstructed:
    pass


def vanishing_construct(ops, committed: VanishingCommitted, domain, h_poly, rng, transcript) -> VanishingConstructed:
    """vanishing/prover.rs:100-152: divide by X^n - 1 on the coset, back to coefficients, cut into n-sized pieces, commit each."""
    n = 1 << domain.k
    h = ops.extended_to_coeff(h_poly, divide_by_vanishing=True)
    c = VanishingConstructed()
    c.h_pieces = [ops.slice(h, i * n, n) for i in range(ops.rows(h) // n)]
    c.h_blinds = _draw(rng, len(c.h_pieces))
    for commitment in ops.commit_many(c.h_pieces):                                  # vanishing/prover.rs:136-146
        transcript.write_point(commitment)
    c.committed = committed
    return c


def vanishing_construct_pieces(ops, committed: VanishingCommitted, pieces, rng, transcript) -> VanishingConstructed:
    """vanishing/prover.rs:117-152 from the quotient's n-sized pieces directly (cosets.py computes them without the extended
    domain): blinds, commitments, transcript -- the same draws and writes as vanishing_construct."""
    c = VanishingConstructed()
    c.h_pieces = list(pieces)
    c.h_blinds = _draw(rng, len(c.h_pieces))
    for commitment in ops.commit_many(c.h_pieces):
        transcript.write_point(commitment)
    c.committed = committed
    return c


class VanishingEvaluated:
    pass


def vanishing_evaluate(ops, c: VanishingConstructed, x: int, xn: int, evals) -> VanishingEvaluated:
    """vanishing/prover.rs:154-186."""
    e = VanishingEvaluated()
    terms, acc = [], 1
    for piece in c.h_pieces:                                                        # rev().fold(acc * xn + piece)
        terms.append((acc, piece))
        acc = acc * xn % R
    e.h_poly = PX.lincomb(ops, terms)
    e.h_blind = 0
    for blind in reversed(c.h_blinds):
        e.h_blind = (e.h_blind * xn + blind) % R
    evals.add(c.committed.random_poly, x)
    e.committed = c.committed
    return e


def vanishing_open(e: VanishingEvaluated, x: int):
    """vanishing/prover.rs:188-206."""
    return [(x, e.h_poly, e.h_blind), (x, e.committed.random_poly, e.committed.random_blind)]
