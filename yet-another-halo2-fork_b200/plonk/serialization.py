"""Key serialisation: `VerifyingKey::write/read`, `ProvingKey::write/read` (halo2_backend/src/plonk.rs:57-118, 276-331), the
permutation keys (plonk/permutation.rs:22-87), `Polynomial::read/write` (poly.rs:170-198) and the `SerdeFormat` switch of
helpers.rs:9-95 -- the data formats on either side of the proving path (the reference's test: halo2_proofs/tests/serialization.rs).

    Processed          points compressed (32 B, `to_bytes`), scalars canonical little-endian (`to_repr`)
    RawBytes           points as (x, y) and scalars as their in-memory Montgomery limbs, little-endian (SerdeObject::write_raw);
                       reading checks that the values are reduced and, for points, on the curve
    RawBytesUnchecked  as RawBytes, read without checks

Lengths inside a key are big-endian u32 (`to_be_bytes`, poly.rs:192, helpers.rs:115), the fixed-commitment count little-endian
(plonk.rs:66).  The polynomials of a proving key are device resident: RawBytes is a straight copy of the device buffer.
"""
from __future__ import annotations

import struct
from typing import List, Tuple

import numpy as np

from .. import fields as F
from . import circuit as C
from . import keygen as K
from . import polyexpr as PX
from .transcript import fr_from_repr, fr_to_repr, g1_from_bytes, g1_to_bytes

VERSION = 0x04                                                      # plonk.rs:52
FORMATS = ("Processed", "RawBytes", "RawBytesUnchecked")


class _Reader:
    def __init__(self, data: bytes):
        self.data, self.pos = memoryview(bytes(data)), 0

    def take(self, n: int) -> bytes:
        if self.pos + n > len(self.data):
            raise IOError("failed to fill whole buffer")
        out = bytes(self.data[self.pos:self.pos + n])
        self.pos += n
        return out


# ---- points (helpers.rs:27-50) -------------------------------------------------------------------------------------------
def point_byte_length(fmt: str) -> int:
    return 32 if fmt == "Processed" else 64


def write_point(p, fmt: str) -> bytes:
    if fmt == "Processed":
        return g1_to_bytes(p)
    x, y = (0, 0) if p is None else p
    return F.fq_mont(x).tobytes() + F.fq_mont(y).tobytes()


def read_point(r: _Reader, fmt: str):
    if fmt == "Processed":
        return g1_from_bytes(r.take(32))
    raw = np.frombuffer(r.take(64), dtype="<u8")
    xm, ym = F.from_limbs(raw[:4]), F.from_limbs(raw[4:])
    if fmt == "RawBytes" and (xm >= F.Q_MOD or ym >= F.Q_MOD):
        raise IOError("invalid field element in a raw point")
    x, y = F.fq_int(raw[:4]), F.fq_int(raw[4:])
    if x == 0 and y == 0:
        return None
    if fmt == "RawBytes" and (y * y - x * x * x - 3) % F.Q_MOD != 0:
        raise IOError("raw point is not on the curve")
    return (x, y)


# ---- polynomials (poly.rs:170-198, helpers.rs:52-95) ------------------------------------------------------------------------
def _strip_mont(ops, p):
    """Montgomery -> canonical limbs on the device (a multiplication by the element whose Montgomery form is 1)."""
    rows = ops.rows(p)
    r_inv = C.const(pow(F.MONT_R, -1, F.R_MOD))
    assert rows in (ops.n, ops.domain.extended_len()), "key polynomials have n or n * e rows"
    return PX.pointwise(ops, C.mul(PX.col(0), r_inv), [p], extended=rows != ops.n)


def _add_mont(ops, p):
    rows = ops.rows(p)
    assert rows in (ops.n, ops.domain.extended_len())
    return PX.pointwise(ops, C.mul(PX.col(0), C.const(F.MONT_R % F.R_MOD)), [p], extended=rows != ops.n)


def write_polynomial(ops, p, fmt: str) -> bytes:
    body = ops.download_bytes(p if fmt != "Processed" else _strip_mont(ops, p))
    return struct.pack(">I", ops.rows(p)) + body


def read_polynomial(ops, r: _Reader, fmt: str):
    (length,) = struct.unpack(">I", r.take(4))
    raw = np.frombuffer(r.take(32 * length), dtype="<u8").reshape(length, 4).copy()
    if fmt != "RawBytesUnchecked":                                  # from_repr / read_raw reject unreduced values
        limbs = F.limbs(F.R_MOD)
        ge = np.zeros(length, dtype=bool)
        eq = np.ones(length, dtype=bool)
        for j in (3, 2, 1, 0):
            ge |= eq & (raw[:, j] > limbs[j])
            eq &= raw[:, j] == limbs[j]
        if np.any(ge | eq):
            raise IOError("Invalid prime field point encoding")
    p = ops.upload_raw(raw)
    return _add_mont(ops, p) if fmt == "Processed" else p


def write_polynomial_slice(ops, polys, fmt: str) -> bytes:
    return struct.pack(">I", len(polys)) + b"".join(write_polynomial(ops, p, fmt) for p in polys)


def read_polynomial_vec(ops, r: _Reader, fmt: str) -> list:
    (count,) = struct.unpack(">I", r.take(4))
    return [read_polynomial(ops, r, fmt) for _ in range(count)]


def polynomial_slice_byte_length(ops, polys) -> int:               # helpers.rs:131-136
    return 4 + len(polys) * (4 + 32 * (ops.rows(polys[0]) if polys else 0))


# ---- keys -------------------------------------------------------------------------------------------------------------------
def vk_write(vk: K.VerifyingKey, fmt: str) -> bytes:
    """plonk.rs:57-71 + permutation.rs:22-33."""
    k = vk.domain.k
    assert k <= F.FR_S
    out = bytes([VERSION, k]) + struct.pack("<I", len(vk.fixed_commitments))
    out += b"".join(write_point(c, fmt) for c in vk.fixed_commitments)
    out += b"".join(write_point(c, fmt) for c in vk.permutation_commitments)
    return out


def vk_bytes_length(vk: K.VerifyingKey, fmt: str) -> int:           # plonk.rs:131-139
    return 6 + (len(vk.fixed_commitments) + len(vk.permutation_commitments)) * point_byte_length(fmt)


def _vk_read(r: _Reader, fmt: str, cs: C.ConstraintSystemBack, make_domain) -> K.VerifyingKey:
    """plonk.rs:74-118: the constraint system comes from the circuit, the domain is rebuilt from it and k."""
    if r.take(1)[0] != VERSION:
        raise IOError("unexpected version byte")
    k = r.take(1)[0]
    if k > F.FR_S:
        raise IOError(f"circuit size value (k): {k} exceeds maxium: {F.FR_S}")
    domain = make_domain(cs.degree(), k)
    (num_fixed,) = struct.unpack("<I", r.take(4))
    fixed = [read_point(r, fmt) for _ in range(num_fixed)]
    perm = [read_point(r, fmt) for _ in cs.permutation_columns]
    vk = K.VerifyingKey(domain, fixed, perm, cs, cs.degree())
    vk.transcript_repr = K._transcript_repr(vk)                     # from_parts, plonk.rs:141-176
    return vk


def vk_read(data: bytes, fmt: str, cs: C.ConstraintSystemBack, make_domain, transcript_repr=None) -> K.VerifyingKey:
    """`transcript_repr`: as in keygen_vk -- a value taken from a Rust run, used instead of the restated Debug hash."""
    vk = _vk_read(_Reader(data), fmt, cs, make_domain)
    if transcript_repr is not None:
        vk.transcript_repr = transcript_repr % F.R_MOD
    return vk


def pk_write(ops, pk: K.ProvingKey, fmt: str) -> bytes:
    """plonk.rs:276-287 + permutation.rs:60-72."""
    out = vk_write(pk.vk, fmt)
    # (a coset-prover key, keygen_pk(extended=False), holds no extended-domain parts: they are rebuilt for the file)
    l0, l_last, l_active_row, fixed_cosets, perm_cosets = K.extended_parts(ops, pk)
    for p in (l0, l_last, l_active_row):
        out += write_polynomial(ops, p, fmt)
    for polys in (pk.fixed_values, pk.fixed_polys, fixed_cosets, pk.permutation.permutations, pk.permutation.polys, perm_cosets):
        out += write_polynomial_slice(ops, polys, fmt)
    return out


def pk_bytes_length(ops, pk: K.ProvingKey, fmt: str) -> int:        # plonk.rs:259-272
    rows = lambda p: ops.rows(p)
    if pk.l0 is None:          # coset-prover key: the extended parts are not held, their sizes follow from the domain
        ne = ops.n << (ops.domain.extended_k - ops.k)
        held = (pk.fixed_values, pk.fixed_polys, pk.permutation.permutations, pk.permutation.polys)
        return (vk_bytes_length(pk.vk, fmt) + 12 + 32 * 3 * ne + sum(polynomial_slice_byte_length(ops, s) for s in held)
                + 2 * 4 + (len(pk.fixed_polys) + len(pk.permutation.polys)) * (4 + 32 * ne))
    return (vk_bytes_length(pk.vk, fmt) + 12 + 32 * (rows(pk.l0) + rows(pk.l_last) + rows(pk.l_active_row))
            + sum(polynomial_slice_byte_length(ops, s) for s in (pk.fixed_values, pk.fixed_polys, pk.fixed_cosets,
                                                                   pk.permutation.permutations, pk.permutation.polys,
                                                                   pk.permutation.cosets)))


def pk_read(ops, data: bytes, fmt: str, cs: C.ConstraintSystemBack) -> K.ProvingKey:
    """plonk.rs:289-316.  `ops` must be bound to EvaluationDomain::new(cs.degree(), k) of the key being read."""
    r = _Reader(data)
    vk = _vk_read(r, fmt, cs, lambda degree, k: ops.domain)
    assert vk.domain.k == ops.k and vk.domain.quotient_poly_degree == cs.degree() - 1, "ops is bound to another domain"
    l0, l_last, l_active_row = (read_polynomial(ops, r, fmt) for _ in range(3))
    fixed_values, fixed_polys, fixed_cosets, permutations, polys, cosets = (read_polynomial_vec(ops, r, fmt) for _ in range(6))
    # the coefficient forms the coset prover reads (cosets.py) are not part of the file: rebuilt as keygen_pk builds them
    l0_poly, _, l_last_poly, l_active_poly = K.indicator_polys(ops, cs)
    return K.ProvingKey(vk, l0, l_last, l_active_row, fixed_values, fixed_polys, fixed_cosets,
                        K.PermutationProvingKey(permutations, polys, cosets), K.Evaluator(cs), l0_poly, l_last_poly, l_active_poly)
