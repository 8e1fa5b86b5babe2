"""DeviceOps: every vector operation the prover mirror needs, on device-resident polynomials through the C ABI.

A polynomial is a torch CUDA tensor of shape (rows, 4) int64 (Montgomery Fr limbs, the reference's `Vec<Fr>` layout).  Scalars that
cross this interface are canonical python ints; commitments come back as affine (x, y) int pairs (None = identity), i.e. after the
`to_affine` / `batch_normalize` the reference applies before `write_point` (prover.rs:443, lookup/prover.rs:146-155).

The prover (prover.py, keygen.py, shplonk.py) is written against this interface only, so the parity tests can run it a second time
over `oracle/plonk_ops.py` (python-int restatement, test infrastructure) and compare proof bytes.  There is no CPU fallback here:
constructing DeviceOps without the CUDA library or a device raises.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from .. import evaluation as EV
from .. import fields as F
from ..arithmetic import PolyOps
from ..domain import EvaluationDomain
from ..engine import B200Engine
from ..kzg import ParamsKZG


class DeviceOps:
    def __init__(self, params: ParamsKZG, engine: B200Engine, domain: EvaluationDomain):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("DeviceOps needs a CUDA device (there is no CPU fallback)")
        self.torch = torch
        self.params, self.engine, self.domain = params, engine, domain
        self.ctx = engine.ctx
        self.poly = PolyOps(self.ctx)
        self.k, self.n = domain.k, domain.n
        self.cosets = None          # a plonk.cosets.CosetProver: create_proof then evaluates the quotient coset by coset

    def with_domain(self, domain: EvaluationDomain) -> "DeviceOps":
        return DeviceOps(self.params, self.engine, domain)

    # ---- data movement ------------------------------------------------------------------------------------------------------
    # torch launches its copies / fills on its own stream while the library works on the context's streams: every torch-side
    # write is completed (`_done`) before the tensor can reach a library call.  Library calls are synchronous on return.
    def _done(self, t=None):
        self.torch.cuda.current_stream().synchronize()
        return t

    def from_ints(self, values: Sequence[int]):
        return self.upload(F.fr_array(list(values)))

    def upload(self, mont: np.ndarray):
        a = np.ascontiguousarray(mont, dtype=np.uint64).reshape(-1, 4)
        return self._done(self.torch.from_numpy(a.view(np.int64)).to("cuda"))

    def download(self, p) -> np.ndarray:
        return p.cpu().numpy().view(np.uint64)

    def download_bytes(self, p) -> bytes:
        """The device buffer as it is: rows x 32 bytes, little-endian limbs."""
        return self.download(p).tobytes()

    def upload_raw(self, limbs: np.ndarray):
        """(rows, 4) little-endian uint64 limbs taken as they are (no conversion)."""
        return self.upload(np.ascontiguousarray(limbs, dtype=np.uint64))

    def to_ints(self, p) -> List[int]:
        return [F.fr_int(row) for row in self.download(p)]

    def zeros(self, rows: int):
        return self._done(self.torch.zeros((rows, 4), dtype=self.torch.int64, device="cuda"))

    def clone(self, p):
        return self._done(p.clone())

    def rows(self, p) -> int:
        return p.shape[0]

    def write_ints(self, p, start: int, values: Sequence[int]) -> None:
        if len(values):
            p[start:start + len(values)] = self.upload(F.fr_array(list(values)))
            self._done()

    def read_ints(self, p, start: int, count: int) -> List[int]:
        return [F.fr_int(row) for row in self.download(p[start:start + count])]

    def slice(self, p, start: int, count: int):
        return self._done(p[start:start + count].clone())

    def resize(self, p, rows: int):
        if p.shape[0] >= rows:
            return self._done(p[:rows].clone())
        out = self.zeros(rows)
        out[:p.shape[0]] = p
        return self._done(out)

    def gather_rows(self, p, index: np.ndarray):
        idx = self.torch.from_numpy(np.ascontiguousarray(index, dtype=np.int64)).to("cuda")
        return self._done(p.index_select(0, idx))

    def concat(self, parts):
        return self._done(self.torch.cat(list(parts), dim=0))

    # ---- transforms (poly/domain.rs) ---------------------------------------------------------------------------------------
    def lagrange_to_coeff(self, p):
        return self.domain.lagrange_to_coeff(p)

    def coeff_to_extended(self, p):
        return self.domain.coeff_to_extended(p)

    def coeff_to_extended_many(self, polys):
        """The independent coset NTTs of evaluate_h (evaluation.rs:344-361, 497-505, 571), spread over the context's devices."""
        return self.domain.coeff_to_extended_many(polys)

    def extended_to_coeff(self, p, divide_by_vanishing: bool = False):
        return self.domain.extended_to_coeff(p, divide_by_vanishing=divide_by_vanishing)

    # ---- commitments (poly/kzg/commitment.rs:305-317, 354-366) + to_affine ---------------------------------------------
    def commit_lagrange(self, p):
        return F.jacobian_to_affine_ints(self.params.commit_lagrange(self.engine, p))

    def commit(self, p):
        return F.jacobian_to_affine_ints(self.params.commit(self.engine, p))

    # ---- pointwise programs (plonk/evaluation.rs GraphEvaluator) ---------------------------------------------------------
    def graph(self, g: EV.GraphEvaluator, fixed, advice, instance, challenges, beta: int, gamma: int, theta: int, y: int,
              extended: bool, previous=None):
        """GraphEvaluator::evaluate over every row of the extended domain (rot_scale = 2^(extended_k - k)) or, with extended=False,
        of the size-n domain with rot_scale 1 (`evaluate(expr, n, 1, ...)`, evaluation.rs:837-880)."""
        ext_k = self.domain.extended_k if extended else self.k
        return g.evaluate(self.ctx, list(fixed), list(advice), list(instance), list(challenges), beta, gamma, theta, y, self.k,
                          ext_k, previous=previous)

    def permutation_fold(self, values, sets, col_values, col_perms, chunk_len, l0, l_last, l_active, beta, gamma, y, delta,
                         blinding_factors):
        d = self.domain
        return EV.permutation_fold(self.ctx, values, sets, col_values, col_perms, chunk_len, l0, l_last, l_active, beta, gamma, y,
                                   F.FR_ZETA, delta, d.extended_omega, blinding_factors, d.k, d.extended_k)

    def lookup_fold(self, values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta, gamma, y):
        d = self.domain
        return EV.lookup_fold(self.ctx, values, table_values, product, permuted_input, permuted_table, l0, l_last, l_active, beta,
                              gamma, y, d.k, d.extended_k)

    def shuffle_fold(self, values, input_values, shuffle_values, product, l0, l_last, l_active, y):
        d = self.domain
        return EV.shuffle_fold(self.ctx, values, input_values, shuffle_values, product, l0, l_last, l_active, y, d.k, d.extended_k)

    # ---- scans (arithmetic.rs, the argument provers) ----------------------------------------------------------------------
    def batch_invert(self, p):
        return self.poly.batch_invert(p)

    def grand_product(self, v, init: int, out_len: int):
        return self.poly.grand_product(v, init, out_len)

    def eval_polynomial(self, p, point: int) -> int:
        return self.poly.eval_polynomial(p, point)

    def eval_polynomials(self, pairs) -> list:
        return self.poly.eval_polynomials(pairs)

    def kate_division(self, p, b: int):
        return self.poly.kate_division(p, b)

    # ---- conveniences built on the above ------------------------------------------------------------------------------------
    def full(self, rows: int, value: int):
        """`rows` copies of one field element, filled on the device (32 bytes cross PCIe, not rows * 32)."""
        one = self.torch.from_numpy(F.fr_mont(value).view(np.int64).reshape(1, 4)).to("cuda")
        return self._done(one.expand(rows, 4).contiguous())

    def as_poly(self, x):
        """A witness column as handed to commit_phase: a device tensor (adopted), an (n, 4) uint64 Montgomery array, or ints."""
        if hasattr(x, "is_cuda"):
            return x
        if isinstance(x, np.ndarray) and x.ndim == 2:
            return self.upload(x)
        return self.from_ints(list(x))

    def commit_lagrange_many(self, polys):
        """The per-column commits of one phase (prover.rs:436-440) through one batched MSM pipeline."""
        polys = list(polys)
        if not polys:
            return []
        out = self.engine.msm_batch(polys, self.params._gl_desc)
        return [F.jacobian_to_affine_ints(j) for j in out]

    def commit_many(self, polys):
        """The h-piece commits of the vanishing argument (vanishing/prover.rs:136-140), one batched MSM against `g`."""
        polys = list(polys)
        if not polys:
            return []
        return [F.jacobian_to_affine_ints(j) for j in self.engine.msm_batch(polys, self.params._g_desc)]

    def to_canonical(self, p, rows: int) -> np.ndarray:
        """First `rows` elements as canonical little-endian limbs (what `Fr: Ord` compares): a Montgomery multiplication by the
        element whose Montgomery form is 1 strips the factor R."""
        from . import circuit as C
        from . import polyexpr as PX
        r_inv = pow(F.MONT_R, -1, F.R_MOD)
        return self.download(PX.pointwise(self, C.mul(PX.col(0), C.const(r_inv)), [p])[:rows].contiguous())

    def from_canonical(self, canon: np.ndarray, rows: int):
        """(m, 4) canonical limbs -> a Montgomery polynomial of `rows` elements (zero padded)."""
        from . import circuit as C
        from . import polyexpr as PX
        raw = self.zeros(rows)
        raw[:canon.shape[0]] = self.upload(canon)
        self._done()
        return PX.pointwise(self, C.mul(PX.col(0), C.const(F.MONT_R % F.R_MOD)), [raw])

    def from_u512(self, limbs: np.ndarray):
        """`Fr::random` / `from_u512`: (m, 8) uint64 little-endian limbs of 512-bit integers, reduced mod r."""
        import ctypes
        limbs = np.ascontiguousarray(limbs, dtype=np.uint64).reshape(-1, 8)
        m = limbs.shape[0]
        src = self.torch.from_numpy(limbs.view(np.int64)).to("cuda")
        out = self.zeros(m)
        self.ctx.check(self.ctx.lib.h2b_fr_from_u512_device(self.ctx.handle, ctypes.c_void_p(src.data_ptr()), m,
                                                            ctypes.c_void_p(out.data_ptr())))
        return out

    def random_poly_chacha20(self, seeds, chunks):
        """vanishing/prover.rs:57-81: chunk i (offset, length) is filled by ChaCha20Rng::from_seed(seeds[i]) with consecutive
        Fr::random draws, on the device (h2b_fr_random_chacha20_device)."""
        import ctypes
        total = sum(length for _, length in chunks)
        offs = [0]
        for off, length in chunks:
            assert off == offs[-1], "chunks must tile the vector in order"
            offs.append(off + length)
        out = self.zeros(total)
        seed_buf = (ctypes.c_uint8 * (32 * len(seeds))).from_buffer_copy(b"".join(seeds))
        off_buf = (ctypes.c_uint64 * len(offs))(*offs)
        self.ctx.check(self.ctx.lib.h2b_fr_random_chacha20_device(self.ctx.handle, seed_buf, off_buf, len(seeds),
                                                                  ctypes.c_void_p(out.data_ptr())))
        return out

    def permute_expression_pair(self, compressed_input, compressed_table, usable: int):
        """lookup/prover.rs:398-494 (`permute_expression_pair`) on the usable rows, device resident (h2b_lookup_permute_device:
        256-bit radix sort of the canonical values, dense ranks, table histogram, one emit kernel).  Returns two n-row Montgomery
        polynomials whose last n - usable rows are zero (the caller blinds them).  Equals arguments.permute_expression_pair, which
        the CPU twin uses (tests/test_gpu_plonk.py compares the proof bytes)."""
        import ctypes
        from .. import _lib
        from .arguments import ConstraintSystemFailure
        out_in, out_tab = self.zeros(self.n), self.zeros(self.n)
        ptr = lambda t: ctypes.c_void_p(t.data_ptr())
        rc = self.ctx.lib.h2b_lookup_permute_device(self.ctx.handle, ptr(compressed_input), ptr(compressed_table), usable, self.n,
                                                    ptr(out_in), ptr(out_tab))
        if rc == _lib.H2B_E_CONSTRAINT:
            raise ConstraintSystemFailure("lookup input value missing from the table")
        self.ctx.check(rc)
        return out_in, out_tab
