"""Host mirror of the reference's O(n) polynomial helpers that sit between MSM and NTT calls, on the GPU.

    eval_polynomial(poly, point)        halo2_backend/src/arithmetic.rs:56-82
    kate_division(a, b)                 halo2_backend/src/arithmetic.rs:99-120
    batch_invert(a)                     ff::BatchInvert, call sites plonk/permutation/prover.rs:124,
                                        plonk/lookup/prover.rs:225, plonk/shuffle/prover.rs:150
    grand_product(v, init, out_len)     the Z-polynomial running products: plonk/permutation/prover.rs:160-166,
                                        plonk/lookup/prover.rs:254-259, plonk/shuffle/prover.rs:161-166

Vectors are `n x 4` uint64 arrays (Montgomery Fr, host) or torch CUDA tensors (device-resident; results stay on the device);
scalars are python ints.  There is no CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import fields as F
from .engine import Context, _is_torch_cuda


def _vp(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _rows(a) -> int:
    if _is_torch_cuda(a):
        return a.numel() * a.element_size() // 32
    return np.asarray(a).shape[0]


class PolyOps:
    def __init__(self, ctx: Optional[Context] = None):
        self.ctx = ctx if ctx is not None else Context()

    def eval_polynomial(self, poly, point: int) -> int:
        """arithmetic.rs:56-82: sum_i poly[i] * point^i (a python int)."""
        c, lib = self.ctx, self.ctx.lib
        out = np.zeros(4, dtype=np.uint64)
        pt = F.fr_mont(point)
        n = _rows(poly)
        if _is_torch_cuda(poly):
            c.check(lib.h2b_eval_polynomial_device(c.handle, ctypes.c_void_p(poly.data_ptr()), n, _vp(pt), _vp(out)))
        else:
            a = np.ascontiguousarray(poly, dtype=np.uint64)
            c.check(lib.h2b_eval_polynomial(c.handle, _vp(a) if n else None, n, _vp(pt), _vp(out)))
        return F.fr_int(out)

    def eval_polynomials(self, pairs) -> list:
        """[eval_polynomial(poly, point) for poly, point in pairs] for device-resident polynomials of one length, in one library
        call (h2b_eval_polynomial_batch_device): the evaluation phase of create_proof (plonk/prover.rs:742-811)."""
        pairs = list(pairs)
        if not pairs:
            return []
        c, lib = self.ctx, self.ctx.lib
        n = _rows(pairs[0][0])
        if not all(_is_torch_cuda(p) and _rows(p) == n for p, _ in pairs):
            return [self.eval_polynomial(p, pt) for p, pt in pairs]
        count = len(pairs)
        ptrs = (ctypes.c_void_p * count)(*[p.data_ptr() for p, _ in pairs])
        pts = np.stack([F.fr_mont(pt) for _, pt in pairs])
        out = np.zeros((count, 4), dtype=np.uint64)
        c.check(lib.h2b_eval_polynomial_batch_device(c.handle, ctypes.cast(ptrs, ctypes.c_void_p), n, _vp(pts), count, _vp(out)))
        return [F.fr_int(out[i]) for i in range(count)]

    def kate_division(self, a, b: int):
        """arithmetic.rs:99-120: quotient of a(X) by (X - b), remainder dropped; len(a) - 1 coefficients."""
        c, lib = self.ctx, self.ctx.lib
        n = _rows(a)
        if n == 0:
            raise AssertionError("kate_division: empty polynomial")            # `a.len() - 1` underflows in the reference
        bm = F.fr_mont(b)
        if _is_torch_cuda(a):
            import torch
            q = torch.empty((max(n - 1, 0), 4), dtype=a.dtype, device=a.device)
            qp = ctypes.c_void_p(q.data_ptr()) if n > 1 else ctypes.c_void_p(a.data_ptr())
            c.check(lib.h2b_kate_division_device(c.handle, ctypes.c_void_p(a.data_ptr()), n, _vp(bm), qp))
            return q
        a = np.ascontiguousarray(a, dtype=np.uint64)
        q = np.empty((n - 1, 4), dtype=np.uint64)
        scratch = q if n > 1 else np.empty((1, 4), dtype=np.uint64)
        c.check(lib.h2b_kate_division(c.handle, _vp(a), n, _vp(bm), _vp(scratch)))
        return q

    def batch_invert(self, a):
        """ff::BatchInvert: every non-zero element replaced by its inverse, in place; zeros stay zero."""
        c, lib = self.ctx, self.ctx.lib
        n = _rows(a)
        if _is_torch_cuda(a):
            c.check(lib.h2b_batch_invert_device(c.handle, ctypes.c_void_p(a.data_ptr()), n))
            return a
        a = np.ascontiguousarray(a, dtype=np.uint64)
        c.check(lib.h2b_batch_invert(c.handle, _vp(a) if n else None, n))
        return a

    def mul(self, a, b, scale=None, out=None):
        """out[i] = a[i] * b[i] (* scale): device tensors of equal length (h2b_fr_mul_device); out defaults to a new tensor."""
        import torch
        c, n = self.ctx, _rows(a)
        assert _is_torch_cuda(a) and _is_torch_cuda(b) and _rows(b) == n
        if out is None:
            out = torch.empty_like(a)
        sc = F.fr_mont(scale) if scale is not None else None
        c.check(c.lib.h2b_fr_mul_device(c.handle, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                        _vp(sc) if sc is not None else None, ctypes.c_void_p(out.data_ptr()), n))
        return out

    def grand_product(self, v, init: int = 1, out_len: Optional[int] = None):
        """z[0] = init; z[i] = z[i-1] * v[i-1] for i < out_len (default len(v), as the Z polynomials take n values)."""
        c, lib = self.ctx, self.ctx.lib
        n = _rows(v)
        if out_len is None:
            out_len = n
        if out_len > n + 1:
            raise ValueError("grand_product: out_len > len(v) + 1")
        im = F.fr_mont(init)
        if _is_torch_cuda(v):
            import torch
            z = torch.empty((out_len, 4), dtype=v.dtype, device=v.device)
            c.check(lib.h2b_grand_product_device(c.handle, ctypes.c_void_p(v.data_ptr()), n, _vp(im),
                                                 ctypes.c_void_p(z.data_ptr()) if out_len else None, out_len))
            return z
        v = np.ascontiguousarray(v, dtype=np.uint64)
        z = np.empty((out_len, 4), dtype=np.uint64)
        c.check(lib.h2b_grand_product(c.handle, _vp(v) if n else None, n, _vp(im), _vp(z) if out_len else None, out_len))
        return z
