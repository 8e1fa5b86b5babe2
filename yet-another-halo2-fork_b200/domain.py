"""EvaluationDomain: host mirror of halo2_backend/src/poly/domain.rs with the transforms on the GPU.

`EvaluationDomain.new(j, k)` restates domain.rs:38-144 (constants only, python ints); the three transforms

    lagrange_to_coeff(a)    domain.rs:216   ->  h2b_ntt(a, k, omega_inv, ifft_divisor)
    coeff_to_extended(a)    domain.rs:230   ->  h2b_coeff_to_extended
    extended_to_coeff(a)    domain.rs:271   ->  h2b_extended_to_coeff
    divide_by_vanishing_poly domain.rs:297  ->  fused into extended_to_coeff(..., divide_by_vanishing=True)

take `n x 4` uint64 arrays (host) or torch CUDA tensors (device-resident, result stays on the device) and
panic -- raise -- on a wrong length exactly where the reference `assert_eq!`s (domain.rs:217,234,272).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import fields as F
from .engine import Context, _is_torch_cuda


def _vp(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


class EvaluationDomain:
    def __init__(self, j: int, k: int, ctx: Optional[Context] = None):
        r = F.R_MOD
        self.quotient_poly_degree = j - 1                               # domain.rs:40
        self.k = k
        self.n = 1 << k
        extended_k = k
        while (1 << extended_k) < self.n * self.quotient_poly_degree:    # domain.rs:48-51
            extended_k += 1
        if extended_k > F.FR_S:                                          # domain.rs:54
            raise AssertionError("extended_k <= F::S")
        self.extended_k = extended_k
        ext_omega = F.ROOT_OF_UNITY
        for _ in range(extended_k, F.FR_S):                              # domain.rs:61-63
            ext_omega = ext_omega * ext_omega % r
        omega = ext_omega
        for _ in range(k, extended_k):                                   # domain.rs:72-74
            omega = omega * omega % r
        self.omega, self.extended_omega = omega, ext_omega
        self.omega_inv = pow(omega, -1, r)
        self.extended_omega_inv = pow(ext_omega, -1, r)
        self.g_coset = F.FR_ZETA                                         # domain.rs:83
        self.g_coset_inv = F.FR_ZETA * F.FR_ZETA % r                     # domain.rs:84
        orig, step = pow(F.FR_ZETA, self.n, r), pow(ext_omega, self.n, r)   # domain.rs:90-91
        t, cur = [], orig
        while True:
            t.append(cur)
            cur = cur * step % r
            if cur == orig:
                break
        assert len(t) == 1 << (extended_k - k)                           # domain.rs:100
        self.t_evaluations = [pow((c - 1) % r, -1, r) for c in t]
        self.ifft_divisor = pow(1 << k, -1, r)
        self.extended_ifft_divisor = pow(1 << extended_k, -1, r)
        self.barycentric_weight = pow(self.n, -1, r)
        # Montgomery encodings handed to the C ABI
        self._m = {name: F.fr_mont(getattr(self, name)) for name in (
            "omega", "omega_inv", "extended_omega", "extended_omega_inv", "g_coset", "g_coset_inv",
            "ifft_divisor", "extended_ifft_divisor")}
        self._t_evals_m = F.fr_array(self.t_evaluations)
        self.ctx = ctx if ctx is not None else Context()

    @staticmethod
    def new(j: int, k: int, ctx: Optional[Context] = None) -> "EvaluationDomain":
        return EvaluationDomain(j, k, ctx)

    def extended_len(self) -> int:
        return 1 << self.extended_k

    def get_omega(self) -> int:
        return self.omega

    def get_omega_inv(self) -> int:
        return self.omega_inv

    def get_extended_omega(self) -> int:
        return self.extended_omega

    def get_quotient_poly_degree(self) -> int:
        return self.quotient_poly_degree

    # ---------------------------------------------------------------------------------------------
    @staticmethod
    def _rows(a) -> int:
        if _is_torch_cuda(a):
            return a.numel() * a.element_size() // 32
        return np.asarray(a).shape[0]

    def lagrange_to_coeff(self, a):
        """domain.rs:216-226.  Consumes `a` (transforms in place), returns it as coefficients."""
        assert self._rows(a) == 1 << self.k, "lagrange_to_coeff: wrong length"       # domain.rs:217
        lib, c = self.ctx.lib, self.ctx
        if _is_torch_cuda(a):
            c.check(lib.h2b_ntt_device(c.handle, ctypes.c_void_p(a.data_ptr()), self.k, _vp(self._m["omega_inv"]),
                                       _vp(self._m["ifft_divisor"])))
            return a
        a = np.ascontiguousarray(a, dtype=np.uint64)
        c.check(lib.h2b_ntt(c.handle, _vp(a), self.k, _vp(self._m["omega_inv"]), _vp(self._m["ifft_divisor"])))
        return a

    def coeff_to_lagrange(self, a):
        """Forward transform on the size-n domain (best_fft(a, omega, k)); helper for tests / round trips."""
        assert self._rows(a) == 1 << self.k
        lib, c = self.ctx.lib, self.ctx
        if _is_torch_cuda(a):
            c.check(lib.h2b_ntt_device(c.handle, ctypes.c_void_p(a.data_ptr()), self.k, _vp(self._m["omega"]), None))
            return a
        a = np.ascontiguousarray(a, dtype=np.uint64)
        c.check(lib.h2b_ntt(c.handle, _vp(a), self.k, _vp(self._m["omega"]), None))
        return a

    def coeff_to_extended(self, a, out=None):
        """domain.rs:230-244."""
        assert self._rows(a) == 1 << self.k, "coeff_to_extended: wrong length"       # domain.rs:234
        lib, c = self.ctx.lib, self.ctx
        if _is_torch_cuda(a):
            import torch
            if out is None:
                out = torch.empty((self.extended_len(), 4), dtype=a.dtype, device=a.device)
            c.check(lib.h2b_coeff_to_extended_device(c.handle, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                                     self.k, self.extended_k, _vp(self._m["extended_omega"]),
                                                     _vp(self._m["g_coset"]), _vp(self._m["g_coset_inv"])))
            return out
        a = np.ascontiguousarray(a, dtype=np.uint64)
        if out is None:
            out = np.empty((self.extended_len(), 4), dtype=np.uint64)
        c.check(lib.h2b_coeff_to_extended(c.handle, _vp(a), _vp(out), self.k, self.extended_k, _vp(self._m["extended_omega"]),
                                          _vp(self._m["g_coset"]), _vp(self._m["g_coset_inv"])))
        return out

    def coeff_to_extended_batch(self, cols, outs=None):
        """The independent per-column coset NTTs of evaluate_h (plonk/evaluation.rs:344-361), one call.  `outs` = optional
        preallocated (n * e, 4) uint64 arrays (one per column) to receive the evaluations."""
        cols = [np.ascontiguousarray(x, dtype=np.uint64) for x in cols]
        for x in cols:
            assert x.shape[0] == 1 << self.k
        if outs is None:
            outs = [np.empty((self.extended_len(), 4), dtype=np.uint64) for _ in cols]
        assert len(outs) == len(cols) and all(o.shape == (self.extended_len(), 4) and o.dtype == np.uint64 and o.flags["C_CONTIGUOUS"] for o in outs)
        lib, c = self.ctx.lib, self.ctx
        inp = (ctypes.c_void_p * len(cols))(*[x.ctypes.data for x in cols])
        outp = (ctypes.c_void_p * len(cols))(*[x.ctypes.data for x in outs])
        c.check(lib.h2b_coeff_to_extended_batch(c.handle, inp, outp, len(cols), self.k, self.extended_k,
                                                _vp(self._m["extended_omega"]), _vp(self._m["g_coset"]),
                                                _vp(self._m["g_coset_inv"])))
        return outs

    def coeff_to_extended_many(self, cols, outs=None):
        """Device-resident columns, spread over the devices of the context (h2b_coeff_to_extended_batch_device)."""
        import torch
        cols = list(cols)
        if not cols:
            return []
        for x in cols:
            assert self._rows(x) == 1 << self.k, "coeff_to_extended: wrong length"
        if outs is None:
            outs = [torch.empty((self.extended_len(), 4), dtype=x.dtype, device=x.device) for x in cols]
        lib, c = self.ctx.lib, self.ctx
        inp = (ctypes.c_void_p * len(cols))(*[x.data_ptr() for x in cols])
        outp = (ctypes.c_void_p * len(cols))(*[x.data_ptr() for x in outs])
        c.check(lib.h2b_coeff_to_extended_batch_device(c.handle, inp, outp, len(cols), self.k, self.extended_k,
                                                       _vp(self._m["extended_omega"]), _vp(self._m["g_coset"]),
                                                       _vp(self._m["g_coset_inv"])))
        return outs

    def lagrange_to_coeff_batch(self, cols):
        cols = [np.ascontiguousarray(x, dtype=np.uint64) for x in cols]
        for x in cols:
            assert x.shape[0] == 1 << self.k
        lib, c = self.ctx.lib, self.ctx
        ptrs = (ctypes.c_void_p * len(cols))(*[x.ctypes.data for x in cols])
        c.check(lib.h2b_ntt_batch(c.handle, ptrs, len(cols), self.k, _vp(self._m["omega_inv"]), _vp(self._m["ifft_divisor"])))
        return cols

    def extended_to_coeff(self, a, divide_by_vanishing: bool = False):
        """domain.rs:271-293 (optionally preceded by divide_by_vanishing_poly, domain.rs:297-316, fused)."""
        assert self._rows(a) == self.extended_len(), "extended_to_coeff: wrong length"   # domain.rs:272
        out_len = self.n * self.quotient_poly_degree
        lib, c = self.ctx.lib, self.ctx
        t_ptr, n_t = (_vp(self._t_evals_m), len(self.t_evaluations)) if divide_by_vanishing else (None, 0)
        args = (self.extended_k, _vp(self._m["extended_omega_inv"]), _vp(self._m["extended_ifft_divisor"]),
                _vp(self._m["g_coset"]), _vp(self._m["g_coset_inv"]), out_len, t_ptr, n_t)
        if _is_torch_cuda(a):
            import torch
            out = torch.empty((out_len, 4), dtype=a.dtype, device=a.device)
            c.check(lib.h2b_extended_to_coeff_device(c.handle, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(out.data_ptr()), *args))
            return out
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.empty((out_len, 4), dtype=np.uint64)
        c.check(lib.h2b_extended_to_coeff(c.handle, _vp(a), _vp(out), *args))
        return out

    def rotate_omega(self, value: int, rotation: int) -> int:            # domain.rs:386-396
        if rotation >= 0:
            return value * pow(self.omega, rotation, F.R_MOD) % F.R_MOD
        return value * pow(self.omega_inv, -rotation, F.R_MOD) % F.R_MOD
