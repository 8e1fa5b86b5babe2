"""B200Engine: the GPU implementation of the reference's ZK-acceleration-layer MSM trait.

Mirrors `trait MsmAccel<C>` (halo2_middleware/src/zal.rs:57-103) method for method:

    msm(coeffs, base)                      zal.rs:58
    get_coeffs_descriptor(coeffs)          zal.rs:81
    get_base_descriptor(base)              zal.rs:82
    msm_with_cached_scalars(desc, base)    zal.rs:84
    msm_with_cached_base(coeffs, desc)     zal.rs:90
    msm_with_cached_inputs(desc, desc)     zal.rs:96

Scalars are `n x 4` uint64 arrays (Montgomery Fr), bases `n x 8` uint64 (affine, identity = (0,0)); results are
`uint64[12]` Jacobian, normalised.  Length mismatch raises (the reference `assert_eq!`s inside best_multiexp).
Descriptors hold device memory and release it on drop, as zal.rs:47-50 asks.
"""
from __future__ import annotations

import ctypes
import json
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import H2bError


def _is_torch_cuda(x) -> bool:
    return hasattr(x, "is_cuda") and bool(x.is_cuda)


def _host_ptr(a: np.ndarray, limbs: int):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if a.ndim != 2 or a.shape[1] != limbs:
        raise ValueError(f"expected an (n, {limbs}) uint64 array, got {a.shape}")
    return a, a.ctypes.data_as(ctypes.c_void_p), a.shape[0]


class Context:
    """Owns one `h2b_ctx` (devices, streams, resident bases, twiddle tables)."""

    def __init__(self, device_ids: Optional[Sequence[int]] = None):
        self.lib = _lib.load()
        self._ctx = ctypes.c_void_p()
        self.device_ids = [0] if device_ids is None else list(device_ids)
        if device_ids is None:
            rc = self.lib.h2b_init(ctypes.byref(self._ctx), None, 0)
        else:
            arr = (ctypes.c_int * len(device_ids))(*device_ids)
            rc = self.lib.h2b_init(ctypes.byref(self._ctx), arr, len(device_ids))
        if rc != 0:
            msg = self.lib.h2b_last_error(None)
            raise H2bError(rc, msg.decode() if msg else "h2b_init failed")

    def check(self, rc: int):
        if rc != 0:
            msg = self.lib.h2b_last_error(self._ctx)
            raise H2bError(rc, msg.decode() if msg else "?")

    @property
    def handle(self):
        return self._ctx

    def last_call_ms(self) -> float:
        return float(self.lib.h2b_last_call_ms(self._ctx))

    def launch_count(self) -> int:
        return int(self.lib.h2b_launch_count(self._ctx))

    def profile_enable(self, on: bool = True):
        self.check(self.lib.h2b_profile_enable(self._ctx, 1 if on else 0))

    def profile_read(self) -> dict:
        buf = ctypes.create_string_buffer(1 << 16)
        self.check(self.lib.h2b_profile_read(self._ctx, buf, len(buf)))
        return json.loads(buf.value.decode())

    def set_option(self, name: str, value=None):
        """Tuning knob of this context (`h2b_set_option`); None restores the default.  The H2B_* environment is only read once,
        when the context is created."""
        v = None if value is None else str(value).encode()
        self.check(self.lib.h2b_set_option(self._ctx, name.encode(), v))

    def last_notice(self) -> str:
        msg = self.lib.h2b_last_notice(self._ctx)
        return msg.decode() if msg else ""

    def imad_peak(self):
        a, b = ctypes.c_double(), ctypes.c_double()
        self.check(self.lib.h2b_imad_peak(self._ctx, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def close(self):
        if self._ctx:
            self.lib.h2b_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BaseDescriptor:
    """`MsmAccel::BaseDescriptor`: bases resident on the device(s)."""

    def __init__(self, ctx: Context, bases):
        self.ctx = ctx
        h = ctypes.c_uint64()
        if _is_torch_cuda(bases):
            self._keep = bases           # adopted, not copied
            self.n = bases.numel() * bases.element_size() // 64
            ctx.check(ctx.lib.h2b_bases_register_device(ctx.handle, ctypes.c_void_p(bases.data_ptr()), self.n, ctypes.byref(h)))
        else:
            a, p, self.n = _host_ptr(bases, 8)
            ctx.check(ctx.lib.h2b_bases_register(ctx.handle, p, self.n, ctypes.byref(h)))
        self.handle = h.value

    def info(self):
        """(window bits of the precomputed multiples, 0 = none; device bytes the descriptor owns)"""
        c, b = ctypes.c_uint32(), ctypes.c_uint64()
        self.ctx.check(self.ctx.lib.h2b_bases_info(self.ctx.handle, self.handle, ctypes.byref(c), ctypes.byref(b)))
        return c.value, b.value

    def release(self):
        if self.handle and self.ctx.handle:
            self.ctx.lib.h2b_bases_release(self.ctx.handle, self.handle)
        self.handle = 0

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class CoeffsDescriptor:
    """`MsmAccel::CoeffsDescriptor`: scalars staged once (a device tensor when torch is available)."""

    def __init__(self, ctx: Context, coeffs):
        self.ctx = ctx
        if _is_torch_cuda(coeffs):
            self.dev = coeffs
            self.n = coeffs.numel() * coeffs.element_size() // 32
        else:
            import torch
            a, _, self.n = _host_ptr(coeffs, 4)
            self.dev = torch.from_numpy(a.view(np.int64)).to("cuda")


class B200Engine:
    """GPU `MsmAccel<G1Affine>` (replaces `H2cEngine`, zal.rs:119,135-177)."""

    def __init__(self, ctx: Optional[Context] = None, device_ids: Optional[Sequence[int]] = None):
        self.ctx = ctx if ctx is not None else Context(device_ids)

    # --- zal.rs:58
    def msm(self, coeffs, base) -> np.ndarray:
        c, cp, n = _host_ptr(coeffs, 4)
        b, bp, nb = _host_ptr(base, 8)
        if n != nb:
            raise ValueError(f"msm: coeffs.len() = {n} != base.len() = {nb}")
        out = np.zeros(12, dtype=np.uint64)
        self.ctx.check(self.ctx.lib.h2b_msm_raw(self.ctx.handle, bp, cp, n, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    # --- zal.rs:81-82
    def get_coeffs_descriptor(self, coeffs) -> CoeffsDescriptor:
        return CoeffsDescriptor(self.ctx, coeffs)

    def get_base_descriptor(self, base) -> BaseDescriptor:
        return BaseDescriptor(self.ctx, base)

    # --- zal.rs:84
    def msm_with_cached_scalars(self, coeffs: CoeffsDescriptor, base) -> np.ndarray:
        desc = BaseDescriptor(self.ctx, base)
        try:
            return self.msm_with_cached_inputs(coeffs, desc)
        finally:
            desc.release()

    # --- zal.rs:90
    def msm_with_cached_base(self, coeffs, base: BaseDescriptor) -> np.ndarray:
        out = np.zeros(12, dtype=np.uint64)
        if _is_torch_cuda(coeffs):
            n = coeffs.numel() * coeffs.element_size() // 32
            self._check_len(n, base)
            self.ctx.check(self.ctx.lib.h2b_msm_device(self.ctx.handle, base.handle, ctypes.c_void_p(coeffs.data_ptr()), n,
                                                       out.ctypes.data_as(ctypes.c_void_p)))
        else:
            c, cp, n = _host_ptr(coeffs, 4)
            self._check_len(n, base)
            self.ctx.check(self.ctx.lib.h2b_msm(self.ctx.handle, base.handle, cp, n, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def msm_batch(self, columns, base: BaseDescriptor) -> np.ndarray:
        """Several scalar vectors of equal length against the same cached bases (the per-column commits of one prover phase,
        plonk/prover.rs:436-440); returns (len(columns), 12) normalised Jacobian points, equal to msm_with_cached_base per column."""
        columns = list(columns)
        out = np.zeros((len(columns), 12), dtype=np.uint64)
        if not columns:
            return out
        if all(_is_torch_cuda(c) for c in columns):
            n = columns[0].numel() * columns[0].element_size() // 32
            assert all(c.numel() * c.element_size() // 32 == n for c in columns), "msm_batch: columns differ in length"
            self._check_len(n, base)
            ptrs = (ctypes.c_void_p * len(columns))(*[c.data_ptr() for c in columns])
            self.ctx.check(self.ctx.lib.h2b_msm_batch_device(self.ctx.handle, base.handle, ptrs, len(columns), n,
                                                             out.ctypes.data_as(ctypes.c_void_p)))
            return out
        host = [_host_ptr(c, 4) for c in columns]
        n = host[0][2]
        assert all(h[2] == n for h in host), "msm_batch: columns differ in length"
        self._check_len(n, base)
        ptrs = (ctypes.c_void_p * len(columns))(*[h[0].ctypes.data for h in host])
        self.ctx.check(self.ctx.lib.h2b_msm_batch(self.ctx.handle, base.handle, ptrs, len(columns), n, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    # --- zal.rs:96
    def msm_with_cached_inputs(self, coeffs: CoeffsDescriptor, base: BaseDescriptor) -> np.ndarray:
        return self.msm_with_cached_base(coeffs.dev, base)

    @staticmethod
    def _check_len(n: int, base: BaseDescriptor):
        if n > base.n:
            raise ValueError(f"msm: {n} coefficients but only {base.n} cached bases")

    # --- multi-process sharding helpers (h2b_msm_windows / h2b_msm_combine)
    def msm_windows(self, coeffs, base: BaseDescriptor, n_total: int):
        out = np.zeros((_lib.MSM_MAX_WINDOWS, 16), dtype=np.uint64)
        c, w = ctypes.c_uint32(), ctypes.c_uint32()
        if _is_torch_cuda(coeffs):
            n = coeffs.numel() * coeffs.element_size() // 32
            ptr, on_dev = ctypes.c_void_p(coeffs.data_ptr()), 1
        else:
            a, ptr, n = _host_ptr(coeffs, 4)
            on_dev = 0
        self._check_len(n, base)
        self.ctx.check(self.ctx.lib.h2b_msm_windows(self.ctx.handle, base.handle, ptr, on_dev, n, n_total,
                                                    out.ctypes.data_as(ctypes.c_void_p), ctypes.byref(c), ctypes.byref(w)))
        return out[: w.value].copy(), c.value

    # --- the same sharding with the exchange inside the library, on the device (h2b_peer_* / h2b_msm_sharded)
    def peer_export(self) -> bytes:
        """Allocate this context's mailbox; returns its 64-byte CUDA IPC handle (self.peer_mailbox = the device pointer, which
        peers living in the same process use instead)."""
        handle = ctypes.create_string_buffer(_lib.PEER_HANDLE_BYTES)
        ptr = ctypes.c_void_p()
        self.ctx.check(self.ctx.lib.h2b_peer_export(self.ctx.handle, handle, ctypes.byref(ptr)))
        self.peer_mailbox = ptr.value
        return handle.raw

    def peer_attach(self, rank: int, world: int, all_gather_bytes=None, local_peers=None):
        """Join a group of `world` single-GPU contexts that shard MSMs by point range.  `all_gather_bytes(b)` must return the
        list of every rank's 64-byte blob in rank order (torch.distributed plumbing: see `dist_all_gather_bytes`);
        `local_peers` = {rank: B200Engine} for peers living in this process (each already `peer_export`ed).
        Attaching clears this rank's mailbox and call counter: barrier after it, before the first `msm_sharded` of any rank."""
        blob = self.peer_export()
        blobs = all_gather_bytes(blob) if all_gather_bytes is not None else [blob] * world
        assert len(blobs) == world and all(len(b) == _lib.PEER_HANDLE_BYTES for b in blobs)
        local = (ctypes.c_void_p * world)()
        for r, peer_engine in (local_peers or {}).items():
            local[r] = peer_engine.peer_mailbox
        self.ctx.check(self.ctx.lib.h2b_peer_attach(self.ctx.handle, rank, world, b"".join(blobs), local))
        self.peer_rank, self.peer_world = rank, world

    def msm_sharded(self, coeffs, base: BaseDescriptor, n_total: int) -> np.ndarray:
        """This rank's share of an n_total-point MSM; returns the WHOLE normalised result on every rank (h2b_msm_sharded)."""
        out = np.zeros(12, dtype=np.uint64)
        if _is_torch_cuda(coeffs):
            n = coeffs.numel() * coeffs.element_size() // 32
            ptr, on_dev = ctypes.c_void_p(coeffs.data_ptr()), 1
        else:
            a, ptr, n = _host_ptr(coeffs, 4)
            on_dev = 0
        self._check_len(n, base)
        self.ctx.check(self.ctx.lib.h2b_msm_sharded(self.ctx.handle, base.handle, ptr, on_dev, n, n_total,
                                                    out.ctypes.data_as(ctypes.c_void_p)))
        return out

    @staticmethod
    def msm_combine(parts: np.ndarray, c: int) -> np.ndarray:
        """parts: (n_parts, windows, 16) uint64 -> normalised Jacobian."""
        lib = _lib.load()
        parts = np.ascontiguousarray(parts, dtype=np.uint64)
        n_parts, windows = parts.shape[0], parts.shape[1]
        out = np.zeros(12, dtype=np.uint64)
        rc = lib.h2b_msm_combine(parts.ctypes.data_as(ctypes.c_void_p), n_parts, c, windows, out.ctypes.data_as(ctypes.c_void_p))
        if rc != 0:
            raise H2bError(rc, "h2b_msm_combine: invalid arguments")
        return out


def shard_range(n: int, rank: int, world: int):
    """Contiguous point range [lo, hi) of rank `rank` out of `world` (MSM sharded by point range)."""
    return n * rank // world, n * (rank + 1) // world


def dist_all_gather_bytes(blob: bytes, group=None):
    """`all_gather_bytes` for B200Engine.peer_attach over torch.distributed (any backend): every rank's blob, in rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


def window_bits(n_total: int, msm_c: int = 0) -> int:
    """The window size every shard of an n_total-point MSM uses (mirrors msm_choose_c in csrc/h2b200.cu; msm_c = the context's
    `msm_c` knob when it is set)."""
    if 4 <= msm_c <= 24:
        return msm_c
    lg = max(n_total, 1).bit_length() - 1
    return min(16, max(4, lg - 4))


class PlonkEngine:
    """zal.rs:182-185: `PlonkEngine { msm_backend }`."""

    def __init__(self, msm_backend: B200Engine):
        self.msm_backend = msm_backend


class PlonkEngineConfig:
    """zal.rs:188-243 builder: PlonkEngineConfig().set_msm(engine).build()."""

    def __init__(self):
        self._msm = None

    @staticmethod
    def new():
        return PlonkEngineConfig()

    def set_curve(self, _curve=None):
        return self

    def set_msm(self, engine: B200Engine):
        self._msm = engine
        return self

    def build(self) -> PlonkEngine:
        if self._msm is None:
            self._msm = B200Engine()
        return PlonkEngine(self._msm)
