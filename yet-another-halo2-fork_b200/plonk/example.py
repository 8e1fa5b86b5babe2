"""The circuit the reference exercises its frontend/backend split with (halo2_proofs/tests/frontend_backend_split.rs:60-468,
WIDTH_FACTOR = 1), laid out directly as a CompiledCircuit + witness columns -- a stand-in for the frontend, which is out of scope.

It has everything the prover's hot path sees in SURVEY.md 8d config 4: a degree-4 gate with a rotation, a second-phase column bound
to a challenge, copy constraints between advice, fixed and instance cells, a three-column dynamic lookup, a two-column shuffle.
Columns: advice a b c (phase 0) e (phase 1); fixed d s_lookup s_ltable s_shuffle s_stable + the selectors s_gate s_rlc s_instance
(selector compression off, :557); one instance column.  A 51-row unit is repeated down the usable rows (:437-456).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from .. import fields as F
from . import circuit as C

A_, B_, C_, E_ = 0, 1, 2, 3
D_, S_LOOKUP, S_LTABLE, S_SHUFFLE, S_STABLE, S_GATE, S_RLC, S_INSTANCE = range(8)
UNIT_ROWS = 51
R = F.R_MOD


def constraint_system() -> C.ConstraintSystemMid:
    a = lambda rot=0: C.query("advice", A_, rot)
    b, c, e = C.query("advice", B_), C.query("advice", C_), C.query("advice", E_)
    d = C.query("fixed", D_)
    fx = lambda i: C.query("fixed", i)
    one, ch = C.const(1), C.challenge(0)
    gate_a = C.mul(fx(S_GATE), C.sub(C.add(a(), C.mul(C.mul(b, c), d)), a(1)))                     # :165-174
    gate_rlc0 = C.mul(fx(S_RLC), C.sub(C.add(a(), C.mul(ch, b)), e))                                # :199-213
    gate_rlc1 = C.mul(fx(S_RLC), C.sub(C.add(c, C.mul(ch, d)), e))
    lookup = C.LookupArgument("lookup.0", [C.mul(x, fx(S_LOOKUP)) for x in (one, a(), b)],
                              [C.mul(x, fx(S_LTABLE)) for x in (one, d, c)])                          # :176-187
    shuffle = C.ShuffleArgument("shuffle.0", [C.mul(x, fx(S_SHUFFLE)) for x in (one, a())],
                                [C.mul(x, fx(S_STABLE)) for x in (one, b)])                           # :189-197
    return C.ConstraintSystemMid(
        num_fixed_columns=8, num_advice_columns=4, num_instance_columns=1, num_challenges=1,
        advice_column_phase=[0, 0, 0, 1], challenge_phase=[0],
        gates=[("gate_a.0", gate_a), ("gate_rlc.0", gate_rlc0), ("gate_rlc.0", gate_rlc1)],
        permutation_columns=[("advice", A_), ("advice", B_), ("fixed", D_), ("instance", 0)],
        lookups=[lookup], shuffles=[shuffle])


class MyCircuit:
    def __init__(self, k: int, input_value: int = 42):
        self.k, self.n, self.input = k, 1 << k, input_value
        unusable_rows = 2 + 3 + 1 + 1 + 1                                                           # :441
        max_rows = self.n - unusable_rows
        self.units = max_rows // UNIT_ROWS
        assert self.units >= 1, "k too small for one unit"
        self._unit = self._layout_unit()
        self._phase0 = None                                 # the first-phase columns do not depend on anything: tiled once
        self._buffers, self._pinned = {}, []

    # ---- one unit (synthesize_unit, :223-422) -----------------------------------------------------------------------------
    def _layout_unit(self):
        adv = {A_: [0] * UNIT_ROWS, B_: [0] * UNIT_ROWS, C_: [0] * UNIT_ROWS}
        fix = [[0] * UNIT_ROWS for _ in range(8)]
        rlc_rows, copies = [], []                          # copies: ((kind, col, row), (kind, col, row)) relative to the unit

        def gate(row, a, abcd):
            _, b, c, d = abcd
            fix[S_GATE][row] = 1
            if a is not None:
                adv[A_][row] = a
            adv[B_][row], adv[C_][row], fix[D_][row] = b, c, d
            adv[A_][row + 1] = (adv[A_][row] + b * c * d) % R
            return row + 1

        fix[S_INSTANCE][0] = 1
        adv[A_][0] = self.input
        copies.append((("instance", 0, 0), ("advice", A_, 0)))                                       # assign_advice_from_instance
        off = 0
        for abcd in ([0, 3, 4, 1], [0, 6, 7, 1], [0, 8, 9, 1], [0, 0xFFFFFFFF, 0xDEADBEEF, 1], [0, 0xABAD1D3A, 0x12345678, 0x42424242]):
            off = gate(off, None, abcd)
        self._instance_cells = [("advice", A_, 1), ("advice", A_, 2), ("advice", A_, 3)]
        off += 1
        rows = []
        for abcd in ([5, 2, 1, 1], [2, 3, 1, 1], [4, 2, 1, 1]):                                       # advice-advice copies
            rows.append(off)
            off = gate(off, abcd[0], abcd) + 1
        copies.append((("advice", B_, rows[0]), ("advice", A_, rows[1])))
        copies.append((("advice", A_, rows[1]), ("advice", B_, rows[2])))
        self._instance_cells += [("advice", B_, rows[0]), ("advice", A_, rows[1])]
        rows = []
        for abcd in ([5, 9, 1, 9], [2, 9, 1, 1], [9, 2, 1, 1]):                                       # advice-fixed copies
            rows.append(off)
            off = gate(off, abcd[0], abcd) + 1
        copies.append((("advice", B_, rows[0]), ("fixed", D_, rows[0])))
        copies.append((("advice", B_, rows[1]), ("fixed", D_, rows[0])))
        copies.append((("advice", A_, rows[2]), ("fixed", D_, rows[0])))
        table = [(exp, 2 ** exp) for exp in range(11)]                                               # dynamic lookup: powers of two
        lookups = [(2, 4), (2, 4), (10, 1024), (0, 1), (2, 4)] + [(0, 1)] * 6
        for (t0, t1), (l0, l1) in zip(table, lookups):
            fix[S_LOOKUP][off] = fix[S_LTABLE][off] = 1
            adv[A_][off], adv[B_][off] = l0, l1
            fix[D_][off], adv[C_][off] = t0, t1
            off += 1
        for abcd in ([3, 5, 3, 5], [8, 9, 8, 9], [111, 222, 111, 222]):                               # RLC gate
            fix[S_RLC][off] = 1
            rlc_rows.append((off, abcd[0], abcd[1]))
            off = gate(off, abcd[0], abcd) + 1
        shuffle = [0, 2, 4, 6, 8, 10, 12, 14, 1, 3, 5, 7, 9, 11, 13, 15]                              # dynamic shuffle
        for t, s in zip(range(16), shuffle):
            fix[S_SHUFFLE][off] = fix[S_STABLE][off] = 1
            adv[A_][off], adv[B_][off] = s, t
            off += 1
        assert off == UNIT_ROWS
        return adv, fix, rlc_rows, copies

    # ---- public pieces ------------------------------------------------------------------------------------------------------
    def instances(self) -> List[List[int]]:
        adv = self._unit[0]
        cell = lambda kind, col, row: adv[col][row]
        return [[self.input] + [cell(*c) for c in self._instance_cells]]                              # :117-135

    def _host_buffer(self, name: str) -> np.ndarray:
        """An (n, 4) uint64 host column, zeroed once and reused; page-locked when a CUDA device is present so that the prover's
        upload is a straight DMA (the witness calculator's output buffers; the rows past the last unit stay zero)."""
        if name not in self._buffers:
            buf = None
            try:
                import torch
                if torch.cuda.is_available():
                    t = torch.zeros((self.n, 4), dtype=torch.int64).pin_memory()
                    self._pinned.append(t)
                    buf = t.numpy().view(np.uint64)
            except ImportError:
                pass
            self._buffers[name] = buf if buf is not None else np.zeros((self.n, 4), dtype=np.uint64)
        return self._buffers[name]

    def _tile(self, unit_column: List[int], name: Optional[str] = None) -> np.ndarray:
        out = self._host_buffer(name) if name else np.zeros((self.n, 4), dtype=np.uint64)
        out[:self.units * UNIT_ROWS].reshape(self.units, UNIT_ROWS, 4)[:] = F.fr_array(unit_column)
        return out

    def compile(self) -> C.CompiledCircuit:
        _, fix, _, unit_copies = self._unit
        copies = []
        for u in range(self.units):
            base = u * UNIT_ROWS
            for (lk, lc, lr), (rk, rc, rr) in unit_copies:
                lrow = lr if lk == "instance" else base + lr
                copies.append((((lk, lc), lrow), ((rk, rc), base + rr)))
            if u == 0:                                                                               # constrain_instance, :449-453
                for i, (kind, col, row) in enumerate(self._instance_cells):
                    copies.append((((kind, col), row), (("instance", 0), 1 + i)))
        return C.CompiledCircuit(C.Preprocessing(copies, [self._tile(col) for col in fix]), constraint_system())

    def witness(self, phase: int, challenges: Dict[int, int]) -> List[Optional[np.ndarray]]:
        adv, _, rlc_rows, _ = self._unit
        if phase == 0:
            if self._phase0 is None:
                self._phase0 = [self._tile(adv[A_], "a"), self._tile(adv[B_], "b"), self._tile(adv[C_], "c")]
            return list(self._phase0) + [None]                  # read-only: the prover uploads them
        ch = challenges[0]
        e = [0] * UNIT_ROWS
        for row, a, b in rlc_rows:
            e[row] = (a + ch * b) % R
        return [None, None, None, self._tile(e, "e")]


class StandardPlonk:
    """The circuit of the reference's proving benchmark (halo2_proofs/benches/plonk.rs:28-266; BASELINE.json configs[0] at k = 12):
    three advice columns under one "combined add-mult" gate a*sa + b*sb + a*b*sm - c*sc, all three in the permutation, no lookups,
    no instance.  2^(k-1) - 3 times: a row a * a = a^2 and a row a + a^2 = c, tied by two copy constraints (:241-256).
    `minimum_degree` = 5 as in the benchmark (:201) gives one permutation set of three columns; None leaves the degree at 3 and
    chunk_len at 1, i.e. three chained permutation sets (permutation/prover.rs:96-190)."""

    def __init__(self, k: int, a: int = 3, minimum_degree: Optional[int] = 5):
        self.k, self.n, self.a, self.minimum_degree = k, 1 << k, a % R, minimum_degree
        self.reps = (1 << (k - 1)) - 3
        assert self.reps >= 1

    def constraint_system(self) -> C.ConstraintSystemMid:
        a, b, c = (C.query("advice", i) for i in range(3))
        sm, sa, sb, sc = (C.query("fixed", i) for i in range(4))
        gate = C.sub(C.add(C.add(C.mul(a, sa), C.mul(b, sb)), C.mul(C.mul(a, b), sm)), C.mul(c, sc))   # :217-229
        return C.ConstraintSystemMid(num_fixed_columns=4, num_advice_columns=3, num_instance_columns=0,
                                     advice_column_phase=[0, 0, 0], gates=[("Combined add-mult", gate)],
                                     permutation_columns=[("advice", 0), ("advice", 1), ("advice", 2)],
                                     minimum_degree=self.minimum_degree)

    def _columns(self, pair):
        out = np.zeros((self.n, 4), dtype=np.uint64)
        out[:2 * self.reps] = np.tile(F.fr_array(pair), (self.reps, 1))
        return out

    def compile(self) -> C.CompiledCircuit:
        fixed = [self._columns(p) for p in ([1, 0], [0, 1], [0, 1], [1, 1])]                         # sm, sa, sb, sc
        copies = []
        for i in range(self.reps):
            copies.append(((("advice", 0), 2 * i), (("advice", 0), 2 * i + 1)))
            copies.append(((("advice", 1), 2 * i + 1), (("advice", 2), 2 * i)))
        return C.CompiledCircuit(C.Preprocessing(copies, fixed), self.constraint_system())

    def instances(self) -> List[List[int]]:
        return []

    def witness(self, phase: int, challenges: Dict[int, int]) -> List[Optional[np.ndarray]]:
        a, sq = self.a, self.a * self.a % R
        return [self._columns([a, a]), self._columns([a, sq]), self._columns([sq, (sq + a) % R])]
