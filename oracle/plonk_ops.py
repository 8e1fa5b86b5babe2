"""RefOps: the CPU twin of yet-another-halo2-fork_b200/plonk/ops.py::DeviceOps on python ints -- TEST INFRASTRUCTURE ONLY.

The prover mirror (plonk/prover.py, keygen.py, shplonk.py) is written against the `ops` interface; running it once over DeviceOps
(the CUDA library) and once over RefOps (this file: the python-int restatements of oracle/pyref.py) with the same RNG must give
byte-identical proofs.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.

Polynomials are python lists of canonical ints.  Commitments use the SRS trapdoor: with the secret s known, commit(p) = [p(s)] G and
commit_lagrange(p) = [sum_i p_i L_i(s)] G are the same group elements the MSM against g / g_lagrange produces
(kzg/commitment.rs:64-131), at the cost of one scalar multiplication.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

from . import pyref as R

P = R.R_MOD
_M64 = (1 << 64) - 1


class RefOps:
    def __init__(self, k: int, s: int, domain: "R.EvaluationDomain"):
        self.k, self.n, self.s, self.domain = k, 1 << k, s % P, domain
        n = self.n
        root = R.ROOT_OF_UNITY
        for _ in range(k, R.FR_S):
            root = root * root % P
        mult = (pow(self.s, n, P) - 1) * R.fr_inv(n) % P
        self._lagrange_at_s, rp = [], 1
        for _ in range(n):                                       # commitment.rs:95-112
            self._lagrange_at_s.append(mult * rp % P * R.fr_inv((self.s - rp) % P) % P)
            rp = rp * root % P

    def with_domain(self, domain) -> "RefOps":
        o = RefOps.__new__(RefOps)
        o.__dict__.update(self.__dict__)
        o.domain = domain
        return o

    # ---- data movement ------------------------------------------------------------------------------------------------------
    def from_ints(self, values: Sequence[int]):
        return [v % P for v in values]

    def upload(self, mont: np.ndarray):
        r_inv = pow(1 << 256, -1, P)
        return [sum(int(x) << (64 * i) for i, x in enumerate(row)) * r_inv % P for row in np.asarray(mont).reshape(-1, 4)]

    # RefOps polynomials are canonical ints; "raw" bytes are the Montgomery limbs the device holds, so that serialised keys agree
    def download_bytes(self, p) -> bytes:
        return b"".join((v * (1 << 256) % P).to_bytes(32, "little") for v in p)

    def upload_raw(self, limbs: np.ndarray):
        return self.upload(limbs)

    def to_ints(self, p) -> List[int]:
        return list(p)

    def zeros(self, rows: int):
        return [0] * rows

    def clone(self, p):
        return list(p)

    def rows(self, p) -> int:
        return len(p)

    def write_ints(self, p, start: int, values: Sequence[int]) -> None:
        p[start:start + len(values)] = [v % P for v in values]

    def read_ints(self, p, start: int, count: int) -> List[int]:
        return list(p[start:start + count])

    def slice(self, p, start: int, count: int):
        return list(p[start:start + count])

    def resize(self, p, rows: int):
        return list(p[:rows]) + [0] * max(0, rows - len(p))

    def gather_rows(self, p, index):
        return [p[int(i)] for i in index]

    def concat(self, parts):
        return [v for part in parts for v in part]

    def full(self, rows: int, value: int):
        return [value % P] * rows

    def as_poly(self, x):
        if isinstance(x, np.ndarray) and x.ndim == 2:
            return self.upload(x)
        return self.from_ints(list(x))

    def to_canonical(self, p, rows: int) -> np.ndarray:
        return np.array([[(v >> (64 * i)) & _M64 for i in range(4)] for v in p[:rows]], dtype=np.uint64).reshape(-1, 4)

    def from_canonical(self, canon: np.ndarray, rows: int):
        vals = [sum(int(x) << (64 * i) for i, x in enumerate(row)) for row in canon]
        return vals + [0] * (rows - len(vals))

    def from_u512(self, limbs: np.ndarray):
        return [sum(int(row[i]) << (64 * i) for i in range(8)) % P for row in limbs]

    def permute_expression_pair(self, compressed_input, compressed_table, usable: int):
        from h2b200.plonk.arguments import permute_expression_pair
        pi, pt = permute_expression_pair(self.to_canonical(compressed_input, usable), self.to_canonical(compressed_table, usable))
        return self.from_canonical(pi, self.n), self.from_canonical(pt, self.n)

    def random_poly_chacha20(self, seeds, chunks):
        from h2b200.plonk.rng import ChaCha20Rng
        out = []
        for seed, (_, length) in zip(seeds, chunks):
            out += self.from_u512(ChaCha20Rng.from_seed(seed).random_fr_limbs(length))
        return out

    # ---- transforms ---------------------------------------------------------------------------------------------------------
    def lagrange_to_coeff(self, p):
        return self.domain.lagrange_to_coeff(p)

    def coeff_to_extended(self, p):
        return self.domain.coeff_to_extended(p)

    def coeff_to_extended_many(self, polys):
        return [self.coeff_to_extended(p) for p in polys]

    def extended_to_coeff(self, p, divide_by_vanishing: bool = False):
        if divide_by_vanishing:
            p = self.domain.divide_by_vanishing_poly(p)
        return self.domain.extended_to_coeff(p)

    # ---- commitments --------------------------------------------------------------------------------------------------------
    def commit(self, p):
        assert len(p) <= self.n
        return R.g1_mul(R.G1_GEN, R.eval_polynomial(p, self.s))

    def commit_lagrange(self, p):
        assert len(p) <= self.n
        return R.g1_mul(R.G1_GEN, sum(a * b for a, b in zip(p, self._lagrange_at_s)) % P)

    def commit_lagrange_many(self, polys):
        return [self.commit_lagrange(p) for p in polys]

    def commit_many(self, polys):
        return [self.commit(p) for p in polys]

    # ---- pointwise programs ---------------------------------------------------------------------------------------------------
    def graph(self, g, fixed, advice, instance, challenges, beta, gamma, theta, y, extended: bool, previous=None):
        d = self.domain
        size = d.extended_len() if extended else self.n
        rot_scale = 1 << (d.extended_k - d.k) if extended else 1
        return R.graph_evaluate(g.calculations, g.constants, g.rotations, fixed, advice, instance, challenges, beta, gamma, theta, y,
                                previous, size, rot_scale)

    def permutation_fold(self, values, sets, col_values, col_perms, chunk_len, l0, l_last, l_active, beta, gamma, y, delta,
                         blinding_factors):
        d = self.domain
        return R.permutation_fold(values, sets, col_values, col_perms, chunk_len, l0, l_last, l_active, beta, gamma, y, R.FR_ZETA,
                                  delta, d.extended_omega, blinding_factors, d.extended_len(), 1 << (d.extended_k - d.k))

    def lookup_fold(self, values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta, gamma, y):
        d = self.domain
        return R.lookup_fold(values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta, gamma, y,
                             d.extended_len(), 1 << (d.extended_k - d.k))

    def shuffle_fold(self, values, input_values, shuffle_values, product, l0, l_last, l_active, y):
        d = self.domain
        return R.shuffle_fold(values, input_values, shuffle_values, product, l0, l_last, l_active, y, d.extended_len(),
                              1 << (d.extended_k - d.k))

    # ---- scans ----------------------------------------------------------------------------------------------------------------
    def batch_invert(self, p):
        return R.batch_invert(p)

    def grand_product(self, v, init: int, out_len: int):
        return R.grand_product(v, init, out_len)

    def eval_polynomial(self, p, point: int) -> int:
        return R.eval_polynomial(p, point)

    def eval_polynomials(self, pairs) -> list:
        return [R.eval_polynomial(p, point) for p, point in pairs]

    def kate_division(self, p, b: int):
        return R.kate_division(p, b)
