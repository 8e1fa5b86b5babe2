"""Synthetic benchmark inputs, generated without touching the CPU oracle.

bases   : [t0 + i]G computed on the device by libh2b200's `h2b_synth_bases` kernel.
scalars : Montgomery residues drawn uniformly below 2^252 (< r, so every value is a valid reduced Fr); the scalar a
          residue v stands for is v * R^-1 mod r, i.e. uniform over a 2^252-element subset of Fr.
"""
from __future__ import annotations

import ctypes

import numpy as np


def random_scalars(n: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64, endpoint=False)
    a[:, :3] ^= rng.integers(0, 1 << 63, size=(n, 3), dtype=np.uint64, endpoint=False) << np.uint64(1)
    a[:, 3] &= np.uint64((1 << 60) - 1)
    return a


def device_bases(ctx, n: int, t0: int = 1):
    """n x 64-byte affine points on the current CUDA device, as a torch int64 tensor of shape (n, 8)."""
    import torch
    out = torch.empty((n, 8), dtype=torch.int64, device="cuda")
    ctx.check(ctx.lib.h2b_synth_bases(ctx.handle, ctypes.c_void_p(out.data_ptr()), n, t0))
    return out
