"""Blake2b transcript and the canonical encodings it needs.

Mirrors halo2_backend/src/transcript.rs for the `Blake2bWrite` / `Blake2bRead` + `Challenge255` pair:

    init                      transcript.rs:308-322   BLAKE2b-512, personalisation "Halo2-Transcript"
    squeeze_challenge         transcript.rs:392-397   update [0]; clone; finalise; Fr::from_uniform_bytes (Challenge255::new :525-533)
    common_point              transcript.rs:399-411   update [1] || x.to_repr() || y.to_repr(); the identity is an error
    common_scalar             transcript.rs:413-418   update [2] || s.to_repr()
    write_point / write_scalar transcript.rs:358-368  common_* then 32 bytes to the proof
    read_point / read_scalar  transcript.rs:186-211

Field elements are python ints (canonical), points are affine (x, y) int pairs or None for the identity.  `to_repr` is the
32-byte little-endian canonical value.  The compressed point form (`G1Affine::to_bytes`, halo2curves 0.6 -- not in the reference
tree) is x little-endian with bit 6 of byte 31 = the low bit of y and bit 7 = identity: PARITY UNPINNED.
"""
from __future__ import annotations

import hashlib
from typing import Optional, Tuple

from ..fields import Q_MOD, R_MOD

Point = Optional[Tuple[int, int]]

BLAKE2B_PREFIX_CHALLENGE = b"\x00"   # transcript.rs:15
BLAKE2B_PREFIX_POINT = b"\x01"       # transcript.rs:18
BLAKE2B_PREFIX_SCALAR = b"\x02"      # transcript.rs:21


def fr_to_repr(s: int) -> bytes:
    return (s % R_MOD).to_bytes(32, "little")


def fr_from_repr(b: bytes) -> int:
    v = int.from_bytes(b, "little")
    if v >= R_MOD:
        raise ValueError("invalid field element encoding in proof")             # transcript.rs:203-208
    return v


def fr_from_uniform_bytes(b: bytes) -> int:
    """`Fr::from_uniform_bytes(&[u8; 64])`: the 512-bit little-endian integer reduced mod r."""
    assert len(b) == 64
    return int.from_bytes(b, "little") % R_MOD


def g1_to_bytes(p: Point) -> bytes:
    if p is None:
        out = bytearray(32)
        out[31] |= 0x80
        return bytes(out)
    x, y = p
    out = bytearray(x.to_bytes(32, "little"))
    out[31] |= (y & 1) << 6
    return bytes(out)


def g1_from_bytes(b: bytes) -> Point:
    assert len(b) == 32
    raw = bytearray(b)
    identity, sign = raw[31] >> 7, (raw[31] >> 6) & 1
    raw[31] &= 0x3F
    x = int.from_bytes(raw, "little")
    if identity:
        if x != 0 or sign:
            raise ValueError("invalid point encoding in proof")
        return None
    if x >= Q_MOD:
        raise ValueError("invalid point encoding in proof")
    y2 = (x * x * x + 3) % Q_MOD
    y = pow(y2, (Q_MOD + 1) // 4, Q_MOD)                                         # q = 3 mod 4
    if y * y % Q_MOD != y2:
        raise ValueError("invalid point encoding in proof")
    if (y & 1) != sign:
        y = Q_MOD - y
    return (x, y)


class _Blake2bTranscript:
    def __init__(self):
        self.state = hashlib.blake2b(digest_size=64, person=b"Halo2-Transcript")

    def squeeze_challenge_scalar(self) -> int:
        self.state.update(BLAKE2B_PREFIX_CHALLENGE)
        return fr_from_uniform_bytes(self.state.copy().digest())

    def common_point(self, p: Point) -> None:
        if p is None:
            raise IOError("cannot write points at infinity to the transcript")
        self.state.update(BLAKE2B_PREFIX_POINT)
        self.state.update(p[0].to_bytes(32, "little"))
        self.state.update(p[1].to_bytes(32, "little"))

    def common_scalar(self, s: int) -> None:
        self.state.update(BLAKE2B_PREFIX_SCALAR)
        self.state.update(fr_to_repr(s))


class Blake2bWrite(_Blake2bTranscript):
    """`Blake2bWrite<Vec<u8>, G1Affine, Challenge255<_>>`."""

    def __init__(self):
        super().__init__()
        self.writer = bytearray()

    @staticmethod
    def init() -> "Blake2bWrite":
        return Blake2bWrite()

    def write_point(self, p: Point) -> None:
        self.common_point(p)
        self.writer += g1_to_bytes(p)

    def write_scalar(self, s: int) -> None:
        self.common_scalar(s)
        self.writer += fr_to_repr(s)

    def finalize(self) -> bytes:
        return bytes(self.writer)


class Blake2bRead(_Blake2bTranscript):
    """`Blake2bRead<&[u8], G1Affine, Challenge255<_>>`."""

    def __init__(self, proof: bytes):
        super().__init__()
        self.reader = memoryview(bytes(proof))
        self.pos = 0

    @staticmethod
    def init(proof: bytes) -> "Blake2bRead":
        return Blake2bRead(proof)

    def _take(self, n: int) -> bytes:
        if self.pos + n > len(self.reader):
            raise IOError("failed to fill whole buffer")
        out = bytes(self.reader[self.pos:self.pos + n])
        self.pos += n
        return out

    def read_point(self) -> Point:
        p = g1_from_bytes(self._take(32))
        self.common_point(p)
        return p

    def read_scalar(self) -> int:
        s = fr_from_repr(self._take(32))
        self.common_scalar(s)
        return s

    def remaining(self) -> int:
        return len(self.reader) - self.pos
