"""Host-side BN254 constants and limb conversions (python ints <-> the reference's 4 x u64 Montgomery limbs).

Mirrors what the Rust host gets for free from halo2curves (`Fr::ROOT_OF_UNITY`, `Fr::ZETA`, `Fr::S`,
`to_repr`): SURVEY.md §A.1.  Only constants and O(1)-per-call scalar work live here; all vector work is
done by the CUDA library.
"""
from __future__ import annotations

import numpy as np

R_MOD = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001
Q_MOD = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47
FR_S = 28
ROOT_OF_UNITY = 0x03DDB9F5166D18B798865EA93DD31F743215CF6DD39329C8D34F1ED960C37C9C
FR_ZETA = 0x30644E72E131A029048B6E193FD84104CC37A73FEC2BC5E9B8CA0B2D36636F23
MONT_R = 1 << 256
_MASK64 = (1 << 64) - 1


def limbs(v: int, n: int = 4) -> np.ndarray:
    return np.array([(v >> (64 * i)) & _MASK64 for i in range(n)], dtype=np.uint64)


def from_limbs(a) -> int:
    a = np.asarray(a, dtype=np.uint64).reshape(-1)
    return sum(int(x) << (64 * i) for i, x in enumerate(a))


def fr_mont(x: int) -> np.ndarray:
    """python int -> 32-byte Montgomery Fr (np.uint64[4])."""
    return limbs((x % R_MOD) * MONT_R % R_MOD)


def fr_int(a) -> int:
    return from_limbs(a) * pow(MONT_R, -1, R_MOD) % R_MOD


def fq_mont(x: int) -> np.ndarray:
    return limbs((x % Q_MOD) * MONT_R % Q_MOD)


def fq_int(a) -> int:
    return from_limbs(a) * pow(MONT_R, -1, Q_MOD) % Q_MOD


def fr_array(xs) -> np.ndarray:
    out = np.empty((len(xs), 4), dtype=np.uint64)
    for i, x in enumerate(xs):
        out[i] = fr_mont(x)
    return out


def jacobian_to_affine_ints(j):
    """96-byte G1 (np.uint64[12]) -> (x, y) python ints, or None for the identity (z == 0)."""
    j = np.asarray(j, dtype=np.uint64).reshape(12)
    x, y, z = fq_int(j[0:4]), fq_int(j[4:8]), fq_int(j[8:12])
    if z == 0:
        return None
    zi = pow(z, -1, Q_MOD)
    return (x * zi * zi % Q_MOD, y * zi * zi * zi % Q_MOD)
