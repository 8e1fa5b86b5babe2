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
