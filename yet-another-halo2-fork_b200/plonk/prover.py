"""Prover / ProverSingle: halo2_backend/src/plonk/prover.rs restated over device-resident polynomials.

    Prover.new / new_with_engine           prover.rs:181-326           vk hash, instance scalars into the transcript, instance polys
    commit_phase(phase, witness)           prover.rs:329-499           blinding rows, blinds, commit_lagrange per column, challenges
    create_proof()                         prover.rs:517-897           theta .. x, the four arguments, evaluate_h, evals, multi-open
    ProverSingle                           prover.rs:40-127            the one-circuit wrapper

`Prover` proves several instances of the same circuit in one transcript (batch proving), in the reference's order: per circuit
inside every transcript step, one vanishing argument and one multi-open for all of them.
All size-n and size-n*e vectors live on the device (`ops`); the host holds scalars, the transcript and the RNG.
KZG: `P::QUERY_INSTANCE == false` (shplonk/prover.rs:113), so instance columns enter the transcript as scalars and are not opened.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

from .. import fields as F
from . import arguments as A
from . import gwc, shplonk
from .keygen import FR_DELTA, ProvingKey

R = F.R_MOD


class ProverError(Exception):
    """plonk::Error (plonk/error.rs)."""


def evaluate_h(ops, pk: ProvingKey, advice_polys, instance_polys, challenges, y, beta, gamma, theta, lookups, shuffles, permutations):
    """Evaluator::evaluate_h (plonk/evaluation.rs:317-623): per circuit, all columns to the extended coset, the custom-gate graph
    over every extended row (continuing the y-Horner of the previous circuit through PreviousValue), then the permutation / lookup /
    shuffle folds, each `value = value * y + term`.  Arguments are lists with one entry per circuit."""
    cs, ev = pk.vk.cs, pk.ev
    fixed = pk.fixed_cosets
    values = None
    for adv_polys, inst_polys, circuit_lookups, circuit_shuffles, permutation in zip(advice_polys, instance_polys, lookups, shuffles,
                                                                                       permutations):
        # every coset extension of this circuit in one call (evaluation.rs:344-361, 497-505, 571): the columns are independent, a
        # multi-device context spreads them over its GPUs
        todo = list(adv_polys) + list(inst_polys)
        for lookup in circuit_lookups:
            todo += [lookup.product_poly, lookup.permuted_input_poly, lookup.permuted_table_poly]
        todo += [shuffle.product_poly for shuffle in circuit_shuffles]
        cosets = ops.coeff_to_extended_many(todo)
        advice, cosets = cosets[:len(adv_polys)], cosets[len(adv_polys):]
        instance, cosets = cosets[:len(inst_polys)], cosets[len(inst_polys):]
        values = ops.graph(ev.custom_gates, fixed, advice, instance, challenges, beta, gamma, theta, y, extended=True, previous=values)
        if permutation:                                                             # evaluation.rs:397-493
            source = {"advice": advice, "fixed": fixed, "instance": instance}
            col_values = [source[kind][index] for kind, index in cs.permutation_columns]
            values = ops.permutation_fold(values, [s.permutation_product_coset for s in permutation], col_values,
                                          pk.permutation.cosets, cs.degree() - 2, pk.l0, pk.l_last, pk.l_active_row, beta, gamma, y,
                                          FR_DELTA, cs.blinding_factors())
        for n_l, lookup in enumerate(circuit_lookups):                              # evaluation.rs:496-567
            product_coset, permuted_input_coset, permuted_table_coset = cosets[:3]
            cosets = cosets[3:]
            table_values = ops.graph(ev.lookups[n_l], fixed, advice, instance, challenges, beta, gamma, theta, y, extended=True)
            values = ops.lookup_fold(values, table_values, product_coset, permuted_input_coset, permuted_table_coset, pk.l0,
                                     pk.l_last, pk.l_active_row, beta, gamma, y)
            del product_coset, permuted_input_coset, permuted_table_coset, table_values
        for n_s, shuffle in enumerate(circuit_shuffles):                            # evaluation.rs:570-620
            product_coset, cosets = cosets[0], cosets[1:]
            input_values = ops.graph(ev.shuffles[2 * n_s], fixed, advice, instance, challenges, beta, gamma, theta, y, extended=True)
            shuffle_values = ops.graph(ev.shuffles[2 * n_s + 1], fixed, advice, instance, challenges, beta, gamma, theta, y,
                                       extended=True)
            values = ops.shuffle_fold(values, input_values, shuffle_values, product_coset, pk.l0, pk.l_last, pk.l_active_row, y)
            del product_coset, input_values, shuffle_values
        del advice, instance, cosets
    return values


class Prover:
    def __init__(self, ops, pk: ProvingKey, circuits_instances: Sequence[Sequence[Sequence[int]]], rng, transcript,
                 num_threads: int = 1, multiopen: str = "shplonk"):
        """prover.rs:181-326.  `circuits_instances`: per circuit, one list of canonical ints per instance column.  `num_threads` is
        the reference's `current_num_threads()` (it seeds the vanishing argument's random polynomial, vanishing/prover.rs:57-81);
        `multiopen` stands for the `P: commitment::Prover` type parameter (ProverSHPLONK / ProverGWC)."""
        cs = pk.vk.cs
        for instance in circuits_instances:
            if len(instance) != cs.num_instance_columns:
                raise ProverError("InvalidInstances")
        self.ops, self.pk, self.rng, self.transcript, self.num_threads = ops, pk, rng, transcript, num_threads
        self.multiopen = {"shplonk": shplonk, "gwc": gwc}[multiopen]
        pk.vk.hash_into(transcript)                                                 # [TRANSCRIPT-1]
        n = ops.n
        self.phases = cs.phases()
        self.instance_values, self.instance_polys = [], []
        for instance in circuits_instances:                                         # [TRANSCRIPT-2]
            values_c, polys_c = [], []
            for values in instance:
                if len(values) > n - (cs.blinding_factors() + 1):
                    raise ProverError("InstanceTooLarge")
                for v in values:
                    transcript.common_scalar(v)
                poly = ops.zeros(n)
                ops.write_ints(poly, 0, list(values))
                values_c.append(poly)
                polys_c.append(ops.lagrange_to_coeff(ops.clone(poly)))
            self.instance_values.append(values_c)
            self.instance_polys.append(polys_c)
        self.n_circuits = len(circuits_instances)
        self.advice_polys: List[List[Optional[object]]] = [[None] * cs.num_advice_columns for _ in range(self.n_circuits)]
        self.advice_blinds: List[List[int]] = [[0] * cs.num_advice_columns for _ in range(self.n_circuits)]
        self.challenges: Dict[int, int] = {}
        self.next_phase_index = 0

    # prover.rs:329-499
    def commit_phase(self, phase: int, witness: Sequence[Sequence[Optional[object]]]) -> Dict[int, int]:
        ops, cs, rng = self.ops, self.pk.vk.cs, self.rng
        if self.next_phase_index >= len(self.phases):
            raise ProverError("All phases already committed")
        current_phase = self.phases[self.next_phase_index]
        if phase != current_phase:
            raise ProverError(f"Committing invalid phase.  Expected {current_phase}, got {phase}")
        column_indices = [i for i, p in enumerate(cs.advice_column_phase) if p == current_phase]
        if len(witness) != self.n_circuits:
            raise ProverError("witness.len() != advice.len()")
        n = ops.n
        per_circuit = []
        for witness_circuit in witness:                                             # prover.rs:359-390: validate everything first
            if len(witness_circuit) != cs.num_advice_columns:
                raise ProverError(f"unexpected length in witness_circuitk.  Got {len(witness_circuit)}, "
                                  f"expected {cs.num_advice_columns}")
            columns = {}
            for column_index, advice_column in enumerate(witness_circuit):
                if column_index in column_indices:
                    if advice_column is None:
                        raise ProverError(f"expected advice column with index {column_index} at phase {current_phase}")
                    poly = ops.as_poly(advice_column)
                    if ops.rows(poly) != n:
                        raise ProverError(f"expected advice column with index {column_index} to have length {n}")
                    columns[column_index] = poly
                elif advice_column is not None:
                    raise ProverError(f"expected no advice column with index {column_index} at phase {current_phase}")
            per_circuit.append(columns)
        unusable_rows_start = n - (cs.blinding_factors() + 1)
        unblinded = set(cs.unblinded_advice_columns)
        for circuit, columns in enumerate(per_circuit):                             # commit_phase_fn, prover.rs:397-457
            for column_index in column_indices:                                     # prover.rs:406-421
                if column_index not in unblinded:
                    ops.write_ints(columns[column_index], unusable_rows_start, A._draw(rng, n - unusable_rows_start))
            blinds = [0 if i in unblinded else rng.random_fr() for i in column_indices]      # prover.rs:424-434
            commitments = ops.commit_lagrange_many([columns[i] for i in column_indices])
            for c in commitments:                                                   # [TRANSCRIPT-3]
                self.transcript.write_point(c)
            for column_index, blind in zip(column_indices, blinds):
                self.advice_polys[circuit][column_index] = columns[column_index]
                self.advice_blinds[circuit][column_index] = blind
        for index, ch_phase in enumerate(cs.challenge_phase):                       # [TRANSCRIPT-4]
            if ch_phase == current_phase:
                assert index not in self.challenges
                self.challenges[index] = self.transcript.squeeze_challenge_scalar()
        self.next_phase_index += 1
        return dict(self.challenges)

    # prover.rs:517-897
    def create_proof(self) -> None:
        ops, pk, rng, transcript = self.ops, self.pk, self.rng, self.transcript
        cs, domain = pk.vk.cs, pk.vk.domain
        assert len(self.challenges) == cs.num_challenges
        challenges = [self.challenges[i] for i in range(cs.num_challenges)]
        circuits = range(self.n_circuits)

        theta = transcript.squeeze_challenge_scalar()                               # [TRANSCRIPT-5]
        permuted_lookups = [[A.lookup_commit_permuted(ops, lookup, pk, theta, self.advice_polys[c], self.instance_values[c],
                                                      challenges, rng, transcript) for lookup in cs.lookups]
                            for c in circuits]                                      # [TRANSCRIPT-6]
        beta = transcript.squeeze_challenge_scalar()                                # [TRANSCRIPT-7]
        gamma = transcript.squeeze_challenge_scalar()                               # [TRANSCRIPT-8]
        permutations = [A.permutation_commit(ops, pk, self.advice_polys[c], self.instance_values[c], beta, gamma, rng, transcript)
                        for c in circuits]                                          # [TRANSCRIPT-9]
        lookups = [[A.lookup_commit_product(ops, p, pk, beta, gamma, rng, transcript) for p in permuted_lookups[c]]
                   for c in circuits]                                               # [TRANSCRIPT-10]
        shuffles = [[A.shuffle_commit_product(ops, s, pk, theta, gamma, self.advice_polys[c], self.instance_values[c], challenges,
                                              rng, transcript) for s in cs.shuffles] for c in circuits]   # [TRANSCRIPT-11]
        vanishing = A.vanishing_commit(ops, domain, rng, transcript, self.num_threads)                  # [TRANSCRIPT-12]
        # prover.rs:665-685 (consumes the Lagrange form)
        advice_polys = [[ops.lagrange_to_coeff(p) for p in self.advice_polys[c]] for c in circuits]
        y = transcript.squeeze_challenge_scalar()                                   # [TRANSCRIPT-13]
        if getattr(ops, "cosets", None):                # the quotient coset by coset, dealt over the GPUs (cosets.py)
            pieces = ops.cosets.evaluate_h_pieces(pk, advice_polys, self.instance_polys, challenges, y, beta, gamma, theta, lookups,
                                                  shuffles, permutations)
            vanishing = A.vanishing_construct_pieces(ops, vanishing, pieces, rng, transcript)           # [TRANSCRIPT-14]
            del pieces
        else:
            h_poly = evaluate_h(ops, pk, advice_polys, self.instance_polys, challenges, y, beta, gamma, theta, lookups, shuffles,
                                permutations)
            vanishing = A.vanishing_construct(ops, vanishing, domain, h_poly, rng, transcript)          # [TRANSCRIPT-14]
            del h_poly
        x = transcript.squeeze_challenge_scalar()                                   # [TRANSCRIPT-15]
        xn = pow(x, ops.n, R)
        evals = A.EvalBatch(ops)                                                    # one device call for [TRANSCRIPT-17..23]
        for c in circuits:                                                          # [TRANSCRIPT-17]
            for column, at in cs.advice_queries:
                evals.add(advice_polys[c][column], domain.rotate_omega(x, at))
        for column, at in cs.fixed_queries:                                         # [TRANSCRIPT-18]
            evals.add(pk.fixed_polys[column], domain.rotate_omega(x, at))
        vanishing = A.vanishing_evaluate(ops, vanishing, x, xn, evals)              # [TRANSCRIPT-19]
        A.permutation_pk_evaluate(ops, pk, x, evals)                                # [TRANSCRIPT-20]
        for c in circuits:                                                          # [TRANSCRIPT-21]
            A.permutation_evaluate(ops, permutations[c], pk, x, evals)
        for c in circuits:                                                          # [TRANSCRIPT-22]
            for committed in lookups[c]:
                A.lookup_evaluate(ops, committed, domain, x, evals)
        for c in circuits:                                                          # [TRANSCRIPT-23]
            for committed in shuffles[c]:
                A.shuffle_evaluate(ops, committed, domain, x, evals)
        known_evals = evals.flush(transcript)
        queries = []
        for c in circuits:                                                          # prover.rs:817-866
            queries += [(domain.rotate_omega(x, at), advice_polys[c][column], self.advice_blinds[c][column])
                        for column, at in cs.advice_queries]
            queries += A.permutation_open(permutations[c], pk, x)
            for committed in lookups[c]:
                queries += A.lookup_open(committed, domain, x)
            for committed in shuffles[c]:
                queries += A.shuffle_open(committed, domain, x)
        queries += [(domain.rotate_omega(x, at), pk.fixed_polys[column], 0) for column, at in cs.fixed_queries]
        queries += A.permutation_pk_open(pk, x)
        queries += A.vanishing_open(vanishing, x)
        self.multiopen.create_proof(ops, transcript, queries, known_evals)                      # prover.rs:884-890


class ProverSingle(Prover):
    """prover.rs:40-127: `ProverSingle(Prover)` for one circuit; `instance` is one list per instance column and `commit_phase` takes
    that circuit's witness columns."""

    def __init__(self, ops, pk: ProvingKey, instance: Sequence[Sequence[int]], rng, transcript, num_threads: int = 1,
                 multiopen: str = "shplonk"):
        super().__init__(ops, pk, [instance], rng, transcript, num_threads, multiopen)

    @staticmethod
    def new(ops, pk, instance, rng, transcript, num_threads: int = 1, multiopen: str = "shplonk") -> "ProverSingle":
        return ProverSingle(ops, pk, instance, rng, transcript, num_threads, multiopen)

    def commit_phase(self, phase: int, witness: Sequence[Optional[object]]) -> Dict[int, int]:
        return super().commit_phase(phase, [witness])


def create_proof(ops, pk: ProvingKey, instance, witness_fn, rng, transcript, num_threads: int = 1,
                 multiopen: str = "shplonk") -> None:
    """plonk::create_proof for one circuit (halo2_proofs/src/plonk/prover.rs:38-123): `witness_fn(phase, challenges)` plays the
    WitnessCalculator and returns the advice columns of that phase (None for the others)."""
    create_proof_batch(ops, pk, [instance], [witness_fn], rng, transcript, num_threads, multiopen)


def create_proof_batch(ops, pk: ProvingKey, circuits_instances, witness_fns, rng, transcript, num_threads: int = 1,
                       multiopen: str = "shplonk") -> None:
    """plonk::create_proof over `circuits: &[ConcreteCircuit]` (halo2_proofs/src/plonk/prover.rs:60-123): every phase is
    synthesised for all circuits, then committed in one `commit_phase`."""
    if len(circuits_instances) != len(witness_fns):
        raise ProverError("InvalidInstances")
    prover = Prover(ops, pk, circuits_instances, rng, transcript, num_threads, multiopen)
    challenges: Dict[int, int] = {}
    for phase in prover.phases:
        challenges = prover.commit_phase(phase, [fn(phase, challenges) for fn in witness_fns])
    prover.create_proof()
