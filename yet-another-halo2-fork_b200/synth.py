"""Synthetic benchmark inputs, generated without touching the CPU oracle.

The bench inputs come from the ChaCha20 seeds SURVEY.md 8(d) fixes, "so CPU oracle, GPU and any future Rust run agree":
    scalars : scalar i = the i-th `Fr::random` draw of ChaCha20Rng::from_seed([1u8; 32])   (`chacha_scalars_device`)
    bases   : [t0 + i]G with t0 = the first next_u64() of ChaCha20Rng::from_seed([2u8; 32])   (`bench_t0`, `device_bases`)
both produced on the device (h2b_fr_random_chacha20_device, h2b_synth_bases); oracle/cref.py generates the same values on the
CPU with its own ChaCha20 (tests/test_gpu_msm.py::test_seeded_inputs_match_the_oracle).
`random_scalars` (numpy Generator) remains for the quick columns of the proof-shaped sequences.
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


SEED_SCALARS, SEED_POINTS, SEED_SRS = bytes([1]) * 32, bytes([2]) * 32, bytes([3]) * 32


def bench_t0() -> int:
    """first next_u64() of rand_chacha::ChaCha20Rng::from_seed([2u8; 32])"""
    from .plonk import rng
    return rng.ChaCha20Rng(SEED_POINTS).next_u64()


def chacha_scalars_device(ctx, n: int, seed: bytes = SEED_SCALARS):
    """n consecutive `Fr::random` draws of ChaCha20Rng::from_seed(seed) as an (n, 4) int64 CUDA tensor (Montgomery)."""
    import torch
    assert len(seed) == 32
    out = torch.empty((n, 4), dtype=torch.int64, device="cuda")
    seed_buf = (ctypes.c_uint8 * 32).from_buffer_copy(seed)
    off_buf = (ctypes.c_uint64 * 2)(0, n)
    ctx.check(ctx.lib.h2b_fr_random_chacha20_device(ctx.handle, seed_buf, off_buf, 1, ctypes.c_void_p(out.data_ptr())))
    return out
