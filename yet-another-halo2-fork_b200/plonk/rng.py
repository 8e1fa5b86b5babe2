"""The random number generators the reference's prover is driven with, restated so that a fixed seed gives a fixed proof.

    OneNg          halo2_proofs/tests/frontend_backend_split.rs:477-489  `BlockRng<OneNg>`: every u32 is 1
    ChaCha20Rng    rand_chacha 0.3 (halo2_backend/Cargo.toml:36; used by vanishing/prover.rs:70-73 for the per-thread streams and by
                   benches/commit_zk.rs:56): ChaCha20, 64-bit block counter in words 12-13, stream id 0, 64-word `BlockRng` buffer
    Fr::random     halo2curves 0.6 (absent from the tree): eight `next_u64()` as little-endian limbs of a 512-bit
                   integer, reduced mod r.  PARITY UNPINNED.

Every generator offers `next_u32`, `next_u64`, `fill_bytes(n)` (rand_core::RngCore) and `random_fr()`.
"""
from __future__ import annotations

import numpy as np

from ..fields import R_MOD


class _RngCore:
    def next_u32(self) -> int:
        raise NotImplementedError

    def next_u64(self) -> int:
        raise NotImplementedError

    def fill_bytes(self, n: int) -> bytes:
        raise NotImplementedError

    def random_fr(self) -> int:
        v = 0
        for i in range(8):
            v |= self.next_u64() << (64 * i)
        return v % R_MOD


class OneNg(_RngCore):
    def next_u32(self) -> int:
        return 1

    def next_u64(self) -> int:
        return (1 << 32) | 1

    def fill_bytes(self, n: int) -> bytes:
        return (b"\x01\x00\x00\x00" * ((n + 3) // 4))[:n]

    def random_fr(self) -> int:
        return _ONE_NG_FR


_ONE_NG_FR = sum(((1 << 32) | 1) << (64 * i) for i in range(8)) % R_MOD

_SIGMA = np.frombuffer(b"expand 32-byte k", dtype="<u4")


def _rotl(x, r):
    return (x << np.uint32(r)) | (x >> np.uint32(32 - r))


def chacha20_blocks(key: bytes, first_block: int, n_blocks: int) -> np.ndarray:
    """(n_blocks, 16) uint32 keystream words of blocks first_block .. first_block + n_blocks - 1, vectorised over blocks."""
    assert len(key) == 32
    kw = np.frombuffer(key, dtype="<u4")
    ctr = np.arange(first_block, first_block + n_blocks, dtype=np.uint64)
    init = [np.full(n_blocks, w, dtype=np.uint32) for w in list(_SIGMA) + list(kw)]
    init.append((ctr & np.uint64(0xFFFFFFFF)).astype(np.uint32))
    init.append((ctr >> np.uint64(32)).astype(np.uint32))
    init += [np.zeros(n_blocks, dtype=np.uint32), np.zeros(n_blocks, dtype=np.uint32)]
    x = [w.copy() for w in init]

    def qr(a, b, c, d):
        x[a] += x[b]; x[d] = _rotl(x[d] ^ x[a], 16)
        x[c] += x[d]; x[b] = _rotl(x[b] ^ x[c], 12)
        x[a] += x[b]; x[d] = _rotl(x[d] ^ x[a], 8)
        x[c] += x[d]; x[b] = _rotl(x[b] ^ x[c], 7)

    with np.errstate(over="ignore"):
        for _ in range(10):
            qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15)
            qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14)
        out = np.stack([x[i] + init[i] for i in range(16)], axis=1)
    return out


class ChaCha20Rng(_RngCore):
    """rand_chacha::ChaCha20Rng::from_seed(seed) behind rand_core's BlockRng (64-word buffer, four blocks per refill)."""

    BUF_WORDS = 64

    def __init__(self, seed: bytes):
        assert len(seed) == 32
        self.key = bytes(seed)
        self.block = 0                  # next block to generate
        self.results = np.zeros(self.BUF_WORDS, dtype=np.uint32)
        self.index = self.BUF_WORDS     # empty

    @staticmethod
    def from_seed(seed: bytes) -> "ChaCha20Rng":
        return ChaCha20Rng(seed)

    def _generate_and_set(self, index: int):
        self.results = chacha20_blocks(self.key, self.block, 4).reshape(-1)
        self.block += 4
        self.index = index

    def next_u32(self) -> int:
        if self.index >= self.BUF_WORDS:
            self._generate_and_set(0)
        v = int(self.results[self.index])
        self.index += 1
        return v

    def next_u64(self) -> int:
        n, i = self.BUF_WORDS, self.index
        if i < n - 1:
            self.index += 2
            return int(self.results[i]) | (int(self.results[i + 1]) << 32)
        if i >= n:
            self._generate_and_set(2)
            return int(self.results[0]) | (int(self.results[1]) << 32)
        x = int(self.results[n - 1])
        self._generate_and_set(1)
        return (int(self.results[0]) << 32) | x

    def fill_bytes(self, n: int) -> bytes:
        out = bytearray()
        while len(out) < n:
            if self.index >= self.BUF_WORDS:
                self._generate_and_set(0)
            avail = self.results[self.index:].astype("<u4").tobytes()
            take = min(len(avail), n - len(out))
            out += avail[:take]
            self.index += (take + 3) // 4
        return bytes(out)

    def random_fr_limbs(self, count: int) -> np.ndarray:
        """`count` consecutive `Fr::random` draws as (count, 8) uint64 little-endian limbs of the 512-bit values (not yet reduced).
        One draw is exactly one ChaCha block when the stream is block aligned, which holds for a fresh generator."""
        if self.index < self.BUF_WORDS and self.index % 16:
            return np.array([[self.next_u64() for _ in range(8)] for _ in range(count)], dtype=np.uint64).reshape(count, 8)
        have = []
        if self.index < self.BUF_WORDS:
            have.append(self.results[self.index:].reshape(-1, 16))
        got = sum(h.shape[0] for h in have)
        if count > got:
            need = count - got
            n_new = (need + 3) // 4 * 4
            have.append(chacha20_blocks(self.key, self.block, n_new))
            self.block += n_new
        words = np.concatenate(have, axis=0)
        used, total = count, words.shape[0]
        # what is left of the last four-block refill stays in the buffer
        left = words[used:total].reshape(-1)
        self.results = np.zeros(self.BUF_WORDS, dtype=np.uint32)
        self.index = self.BUF_WORDS - left.size
        if left.size:
            self.results[self.index:] = left
        w = np.ascontiguousarray(words[:used]).astype(np.uint64)
        return (w[:, 0::2] | (w[:, 1::2] << np.uint64(32))).astype(np.uint64)


def limbs512_to_ints(limbs: np.ndarray):
    """(count, 8) uint64 -> python ints reduced mod r (the `from_u512` map)."""
    out = []
    for row in limbs:
        v = 0
        for i in range(8):
            v |= int(row[i]) << (64 * i)
        out.append(v % R_MOD)
    return out
