"""ParamsKZG: host mirror of halo2_backend/src/poly/kzg/commitment.rs with both MSMs on the GPU.

    commit_lagrange(engine, poly, blind)   commitment.rs:305-317  -> engine.msm(poly, g_lagrange[0..len])
    commit(engine, poly, blind)            commitment.rs:354-366  -> engine.msm(poly, g[0..len])

`g` and `g_lagrange` are registered once as base descriptors (resident SRS), exactly the caching half of
the ZAL API the reference defines (zal.rs:78-100).  The blind is ignored, as in the reference (`_: Blind`).
"""
from __future__ import annotations

import struct

import numpy as np

from .engine import B200Engine, BaseDescriptor


class ParamsKZG:
    def __init__(self, k: int, g: np.ndarray, g_lagrange: np.ndarray, engine: B200Engine):
        n = 1 << k
        if hasattr(g, "is_cuda"):                     # device-resident SRS (ParamsKZG.setup): adopted by the descriptors, not copied
            assert tuple(g.shape) == (n, 8) and tuple(g_lagrange.shape) == (n, 8)
        else:
            g = np.ascontiguousarray(g, dtype=np.uint64)
            g_lagrange = np.ascontiguousarray(g_lagrange, dtype=np.uint64)
            assert g.shape == (n, 8) and g_lagrange.shape == (n, 8)
        self.k, self.n = k, n
        self.g, self.g_lagrange = g, g_lagrange
        self._engine = engine
        self._g_desc = BaseDescriptor(engine.ctx, g)
        self._gl_desc = BaseDescriptor(engine.ctx, g_lagrange)

    @staticmethod
    def from_parts(k: int, g, g_lagrange, engine: B200Engine) -> "ParamsKZG":   # commitment.rs:135-154
        return ParamsKZG(k, g, g_lagrange, engine)

    @staticmethod
    def setup(k: int, s: int, engine: B200Engine) -> "ParamsKZG":
        """ParamsKZG::setup (commitment.rs:64-131) with the secret given instead of drawn (`let s = <E::Fr>::random(rng)`, :69):
        g[i] = [s^i] G (:74-83) and g_lagrange[i] = [(s^n - 1)/n * w^i / (s - w^i)] G (:95-112), all on the device: the scalar
        vectors through the scan / pointwise kernels, the 2n fixed-base multiplications by h2b_g1_fixed_base_mul_device."""
        import ctypes

        import torch

        from . import fields as F
        from .domain import EvaluationDomain
        from .plonk import circuit as C
        from .plonk import polyexpr as PX
        from .plonk.ops import DeviceOps
        ctx, n, r = engine.ctx, 1 << k, F.R_MOD
        dom = EvaluationDomain(3, k, ctx)
        ops = DeviceOps(None, engine, dom)
        s %= r
        powers = ops.grand_product(ops.full(n, s), 1, n)
        omegas = ops.grand_product(ops.full(n, dom.omega), 1, n)
        inv = ops.batch_invert(PX.pointwise(ops, C.sub(C.const(s), PX.col(0)), [omegas]))
        mult = (pow(s, n, r) - 1) * pow(n, -1, r) % r
        lagrange = PX.pointwise(ops, C.mul(C.mul(C.const(mult), PX.col(0)), PX.col(1)), [omegas, inv])
        out = []
        for scalars in (powers, lagrange):
            pts = torch.empty((n, 8), dtype=torch.int64, device="cuda")
            ctx.check(ctx.lib.h2b_g1_fixed_base_mul_device(ctx.handle, ctypes.c_void_p(scalars.data_ptr()), n,
                                                           ctypes.c_void_p(pts.data_ptr())))
            out.append(pts)
        return ParamsKZG(k, out[0], out[1], engine)

    @staticmethod
    def read_raw(data: bytes, engine: B200Engine) -> "ParamsKZG":
        """SerdeFormat::RawBytes prefix of ParamsKZG::write_custom (commitment.rs:167-181): k:u32 LE, then
        n x 64 B `g`, n x 64 B `g_lagrange` (Montgomery limbs: already the device layout); g2 / s_g2 ignored."""
        (k,) = struct.unpack_from("<I", data, 0)
        n = 1 << k
        g = np.frombuffer(data, dtype=np.uint64, count=n * 8, offset=4).reshape(n, 8)
        gl = np.frombuffer(data, dtype=np.uint64, count=n * 8, offset=4 + n * 64).reshape(n, 8)
        return ParamsKZG(k, g, gl, engine)

    def get_g(self) -> np.ndarray:                                               # commitment.rs:368
        return self.g

    def commit_lagrange(self, engine: B200Engine, poly, blind=None) -> np.ndarray:
        size = poly.shape[0] if hasattr(poly, "shape") else len(poly)
        assert self.n >= size, "bases.len() >= size"                              # commitment.rs:315
        return engine.msm_with_cached_base(poly, self._gl_desc)

    def commit(self, engine: B200Engine, poly, blind=None) -> np.ndarray:
        size = poly.shape[0] if hasattr(poly, "shape") else len(poly)
        assert self.n >= size, "bases.len() >= size"                              # commitment.rs:364
        return engine.msm_with_cached_base(poly, self._g_desc)
