"""ctypes binding of oracle/libh2ref.so (the plain-C CPU restatement, oracle/h2ref.c).

TEST INFRASTRUCTURE ONLY -- see the header of oracle/h2ref.c.  Buffers are numpy uint64 arrays in the
boundary layout: Fr = 4 limbs (Montgomery), G1Affine = 8 limbs, G1 Jacobian = 12 limbs.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libh2ref.so")
_SRC = os.path.join(_HERE, "h2ref.c")


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(
            ["gcc", "-O3", "-fopenmp", "-shared", "-fPIC", "-o", _SO, _SRC, "-lm"], cwd=_HERE)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        try:
            build()                      # no-op when libh2ref.so is newer than h2ref.c
        except Exception:
            if not os.path.exists(_SO):
                raise
        _lib = ctypes.CDLL(_SO)
        vp, sz, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint64
        sigs = {
            "h2ref_best_multiexp": [vp, vp, sz, vp],
            "h2ref_msm_naive": [vp, vp, sz, vp],
            "h2ref_to_affine": [vp, vp],
            "h2ref_is_on_curve": [vp],
            "h2ref_affine_add": [vp, vp, vp],
            "h2ref_affine_mul": [vp, vp, vp],
            "h2ref_best_fft": [vp, vp, u32],
            "h2ref_ifft": [vp, vp, u32, vp],
            "h2ref_distribute_powers_zeta": [vp, sz, vp, vp],
            "h2ref_lagrange_to_coeff": [vp, u32, vp, vp],
            "h2ref_coeff_to_extended": [vp, vp, u32, u32, vp, vp, vp],
            "h2ref_extended_to_coeff": [vp, u32, vp, vp, vp, vp],
            "h2ref_divide_by_vanishing_poly": [vp, sz, vp, sz],
            "h2ref_eval_polynomial": [vp, sz, vp, vp],
            "h2ref_kate_division": [vp, sz, vp, vp],
            "h2ref_batch_invert": [vp, sz],
            "h2ref_grand_product": [vp, sz, vp, vp, sz],
            "h2ref_fr_mul": [vp, vp, vp], "h2ref_fq_mul": [vp, vp, vp],
            "h2ref_fr_add": [vp, vp, vp], "h2ref_fr_sub": [vp, vp, vp], "h2ref_fr_inv": [vp, vp],
            "h2ref_random_fr": [vp, sz, u64],
            "h2ref_random_fr_chacha20": [vp, sz, vp, u64],
            "h2ref_chacha20_block": [vp, u64, vp],
            "h2ref_gen_bases": [vp, sz, u64],
            "h2ref_kzg_setup": [u32, vp, vp, vp, vp],
            "h2ref_num_threads": [],
            "h2ref_set_num_threads": [ctypes.c_int],
        }
        for name, args in sigs.items():
            fn = getattr(_lib, name)
            fn.argtypes = args
            fn.restype = ctypes.c_int if name in ("h2ref_is_on_curve", "h2ref_num_threads") else None
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads() -> int:
    return lib().h2ref_num_threads()


def use_all_host_cores() -> int:
    """Give the OpenMP loops every core this process may run on (the launcher may have exported OMP_NUM_THREADS=1)."""
    import os
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().h2ref_set_num_threads(n)
    return num_threads()


def random_fr(n: int, seed: int) -> np.ndarray:
    out = np.empty((n, 4), dtype=np.uint64)
    lib().h2ref_random_fr(_p(out), n, seed)
    return out


def chacha20_block(seed: bytes, counter: int) -> np.ndarray:
    """16 little-endian keystream words of one rand_chacha ChaCha20 block (64-bit counter, stream 0)."""
    assert len(seed) == 32
    out = np.zeros(16, dtype=np.uint32)
    lib().h2ref_chacha20_block(ctypes.c_char_p(seed), counter, out.ctypes.data_as(ctypes.c_void_p))
    return out


def random_fr_chacha20(n: int, seed: bytes, first_block: int = 0) -> np.ndarray:
    """n consecutive `Fr::random` draws of ChaCha20Rng::from_seed(seed), starting at draw `first_block`."""
    assert len(seed) == 32
    out = np.empty((n, 4), dtype=np.uint64)
    lib().h2ref_random_fr_chacha20(_p(out), n, ctypes.c_char_p(seed), first_block)
    return out


# SURVEY.md 8(d) seeds of the synthetic benchmark inputs
SEED_SCALARS, SEED_POINTS, SEED_SRS = bytes([1]) * 32, bytes([2]) * 32, bytes([3]) * 32


def bench_t0() -> int:
    """bases are [t0 + i]G with t0 = the first next_u64() of ChaCha20Rng::from_seed([2u8; 32])"""
    w = chacha20_block(SEED_POINTS, 0)
    return int(w[0]) | (int(w[1]) << 32)


def gen_bases(n: int, t0: int = 1) -> np.ndarray:
    out = np.empty((n, 8), dtype=np.uint64)
    lib().h2ref_gen_bases(_p(out), n, t0)
    return out


def best_multiexp(coeffs: np.ndarray, bases: np.ndarray) -> np.ndarray:
    assert coeffs.shape[0] == bases.shape[0]
    out = np.zeros(12, dtype=np.uint64)
    lib().h2ref_best_multiexp(_p(coeffs), _p(bases), coeffs.shape[0], _p(out))
    return out


def msm_naive(coeffs: np.ndarray, bases: np.ndarray) -> np.ndarray:
    out = np.zeros(12, dtype=np.uint64)
    lib().h2ref_msm_naive(_p(coeffs), _p(bases), coeffs.shape[0], _p(out))
    return out


def to_affine(j: np.ndarray) -> np.ndarray:
    out = np.zeros(8, dtype=np.uint64)
    lib().h2ref_to_affine(_p(np.ascontiguousarray(j)), _p(out))
    return out


def is_on_curve(a: np.ndarray) -> bool:
    return bool(lib().h2ref_is_on_curve(_p(np.ascontiguousarray(a))))


def affine_add(a, b) -> np.ndarray:
    out = np.zeros(8, dtype=np.uint64)
    lib().h2ref_affine_add(_p(np.ascontiguousarray(a)), _p(np.ascontiguousarray(b)), _p(out))
    return out


def affine_mul(p, k) -> np.ndarray:
    out = np.zeros(8, dtype=np.uint64)
    lib().h2ref_affine_mul(_p(np.ascontiguousarray(p)), _p(np.ascontiguousarray(k)), _p(out))
    return out


def best_fft(a: np.ndarray, omega: np.ndarray, log_n: int) -> np.ndarray:
    a = np.ascontiguousarray(a).copy()
    assert a.shape[0] == 1 << log_n
    lib().h2ref_best_fft(_p(a), _p(omega), log_n)
    return a


def lagrange_to_coeff(a, k, omega_inv, ifft_divisor) -> np.ndarray:
    a = np.ascontiguousarray(a).copy()
    lib().h2ref_lagrange_to_coeff(_p(a), k, _p(omega_inv), _p(ifft_divisor))
    return a


def coeff_to_extended(a, k, ext_k, ext_omega, g_coset, g_coset_inv) -> np.ndarray:
    a = np.ascontiguousarray(a)
    out = np.empty((1 << ext_k, 4), dtype=np.uint64)
    lib().h2ref_coeff_to_extended(_p(a), _p(out), k, ext_k, _p(ext_omega), _p(g_coset), _p(g_coset_inv))
    return out


def extended_to_coeff(a, ext_k, ext_omega_inv, ext_ifft_divisor, g_coset, g_coset_inv, out_len) -> np.ndarray:
    a = np.ascontiguousarray(a).copy()
    lib().h2ref_extended_to_coeff(_p(a), ext_k, _p(ext_omega_inv), _p(ext_ifft_divisor), _p(g_coset), _p(g_coset_inv))
    return a[:out_len].copy()


def divide_by_vanishing_poly(a, t_evals) -> np.ndarray:
    a = np.ascontiguousarray(a).copy()
    t = np.ascontiguousarray(t_evals)
    lib().h2ref_divide_by_vanishing_poly(_p(a), a.shape[0], _p(t), t.shape[0])
    return a


def eval_polynomial(poly: np.ndarray, point: np.ndarray) -> np.ndarray:
    """arithmetic.rs:56-82; point and result are 4-limb Montgomery Fr."""
    poly = np.ascontiguousarray(poly, dtype=np.uint64)
    out = np.zeros(4, dtype=np.uint64)
    lib().h2ref_eval_polynomial(_p(poly), poly.shape[0], _p(np.ascontiguousarray(point, dtype=np.uint64)), _p(out))
    return out


def kate_division(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """arithmetic.rs:99-120; returns len(a) - 1 coefficients."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    assert a.shape[0] >= 1
    q = np.zeros((a.shape[0] - 1, 4), dtype=np.uint64)
    scratch = q if a.shape[0] > 1 else np.zeros((1, 4), dtype=np.uint64)
    lib().h2ref_kate_division(_p(a), a.shape[0], _p(np.ascontiguousarray(b, dtype=np.uint64)), _p(scratch))
    return q


def batch_invert(a: np.ndarray) -> np.ndarray:
    """ff::BatchInvert on a copy: non-zero elements inverted, zeros untouched."""
    a = np.array(a, dtype=np.uint64, copy=True, order="C")
    lib().h2ref_batch_invert(_p(a), a.shape[0])
    return a


def grand_product(v: np.ndarray, init: np.ndarray, out_len: int) -> np.ndarray:
    """z[0] = init, z[i] = z[i-1] * v[i-1] (permutation/prover.rs:160-166 and the lookup / shuffle analogues)."""
    v = np.ascontiguousarray(v, dtype=np.uint64)
    z = np.zeros((out_len, 4), dtype=np.uint64)
    lib().h2ref_grand_product(_p(v), v.shape[0], _p(np.ascontiguousarray(init, dtype=np.uint64)), _p(z), out_len)
    return z


def kzg_setup(k: int, s: np.ndarray, omega: np.ndarray):
    n = 1 << k
    g = np.empty((n, 8), dtype=np.uint64)
    gl = np.empty((n, 8), dtype=np.uint64)
    lib().h2ref_kzg_setup(k, _p(s), _p(omega), _p(g), _p(gl))
    return g, gl


def _binop(name, a, b):
    out = np.zeros(4, dtype=np.uint64)
    getattr(lib(), name)(_p(np.ascontiguousarray(a)), _p(np.ascontiguousarray(b)), _p(out))
    return out


def fr_mul(a, b): return _binop("h2ref_fr_mul", a, b)
def fq_mul(a, b): return _binop("h2ref_fq_mul", a, b)
def fr_add(a, b): return _binop("h2ref_fr_add", a, b)
def fr_sub(a, b): return _binop("h2ref_fr_sub", a, b)


def fr_inv(a):
    out = np.zeros(4, dtype=np.uint64)
    lib().h2ref_fr_inv(_p(np.ascontiguousarray(a)), _p(out))
    return out


# ---- conversions between python ints and limb arrays --------------------------------------------
def limbs_from_int(v: int, nlimbs: int = 4) -> np.ndarray:
    return np.array([(v >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(nlimbs)], dtype=np.uint64)


def int_from_limbs(a) -> int:
    a = np.asarray(a, dtype=np.uint64).reshape(-1)
    return sum(int(x) << (64 * i) for i, x in enumerate(a))


def mont_from_int(x: int, mod: int) -> np.ndarray:
    return limbs_from_int((x << 256) % mod)


def int_from_mont(a, mod: int) -> int:
    return int_from_limbs(a) * pow(1 << 256, -1, mod) % mod


def fr_array_from_ints(xs, mod: int) -> np.ndarray:
    return np.stack([mont_from_int(x, mod) for x in xs]) if len(xs) else np.empty((0, 4), dtype=np.uint64)


def ints_from_fr_array(a: np.ndarray, mod: int):
    rinv = pow(1 << 256, -1, mod)
    return [int_from_limbs(row) * rinv % mod for row in a]


def affine_from_ints(P, qmod: int) -> np.ndarray:
    if P is None:
        return np.zeros(8, dtype=np.uint64)
    return np.concatenate([mont_from_int(P[0], qmod), mont_from_int(P[1], qmod)])


def affine_to_ints(a, qmod: int):
    a = np.asarray(a, dtype=np.uint64).reshape(8)
    x, y = int_from_mont(a[:4], qmod), int_from_mont(a[4:], qmod)
    return None if (x == 0 and y == 0) else (x, y)
