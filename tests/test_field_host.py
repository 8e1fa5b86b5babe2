"""The limb-level algorithms the CUDA kernels run (field.cuh / curve.cuh / ntt.cuh / msm.cuh), compiled for the
host with the PTX carry-chain instructions emulated in C (csrc/host_check.cpp), against the oracle.  This is how the
kernels' arithmetic and index logic are validated in a container without a GPU."""
import ctypes
import os
import random
import subprocess
import sys

import numpy as np
import pytest

from oracle import cref as C
from oracle import pyref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "yet-another-halo2-fork_b200")


@pytest.fixture(scope="module")
def hc():
    so = os.path.join(PKG, "libh2b200_hostcheck.so")
    src = [os.path.join(PKG, "csrc", f) for f in os.listdir(os.path.join(PKG, "csrc"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call([sys.executable, "-c", "import sys; sys.path.insert(0, %r); import build; build.build_hostcheck(True)" % PKG])
    lib = ctypes.CDLL(so)
    lib.hc_ntt.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32] + [ctypes.c_void_p] * 4
    return lib


def p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def m(x):
    return C.mont_from_int(x, R.R_MOD)


def test_field_ops(hc):
    rnd = random.Random(1)
    for mod, nm in ((R.R_MOD, "fr"), (R.Q_MOD, "fq")):
        rinv = pow(1 << 256, -1, mod)
        edge = [0, 1, 2, mod - 1, mod - 2, (1 << 256) % mod, (1 << 255) % mod, (mod - 1) // 2, 0xFFFFFFFF, (1 << 128) - 1, mod - (1 << 32)]
        cases = [(a, b) for a in edge for b in edge] + [(rnd.randrange(mod), rnd.randrange(mod)) for _ in range(1500)]
        for a, b in cases:
            A, B = C.limbs_from_int(a), C.limbs_from_int(b)
            for op, exp in (("mul", a * b * rinv % mod), ("add", (a + b) % mod), ("sub", (a - b) % mod)):
                o = np.zeros(4, dtype=np.uint64)
                getattr(hc, f"hc_{nm}_{op}")(p(A), p(B), p(o))
                assert C.int_from_limbs(o) == exp, (nm, op, hex(a), hex(b))
    for a in [0, 1, R.R_MOD - 1] + [rnd.randrange(R.R_MOD) for _ in range(10)]:
        A = C.limbs_from_int(a)
        o = np.zeros(4, dtype=np.uint64)
        hc.hc_fr_neg(p(A), p(o)); assert C.int_from_limbs(o) == (-a) % R.R_MOD
        hc.hc_fr_from_mont(p(A), p(o)); assert C.int_from_limbs(o) == a * pow(1 << 256, -1, R.R_MOD) % R.R_MOD
        hc.hc_fr_to_mont(p(A), p(o)); assert C.int_from_limbs(o) == (a << 256) % R.R_MOD
        hc.hc_fr_inv(p(m(a)), p(o)); assert C.int_from_mont(o, R.R_MOD) == (pow(a, -1, R.R_MOD) if a else 0)
        # the windowed Fermat inversion the batch-inversion kernel uses (poly.cuh)
        hc.hc_fr_inv_window(p(m(a)), p(o)); assert C.int_from_mont(o, R.R_MOD) == (pow(a, -1, R.R_MOD) if a else 0)
        aq = a % R.Q_MOD
        hc.hc_fq_inv_window(p(C.mont_from_int(aq, R.Q_MOD)), p(o)); assert C.int_from_mont(o, R.Q_MOD) == (pow(aq, -1, R.Q_MOD) if aq else 0)


def test_curve_ops(hc):
    rnd = random.Random(2)
    bs = C.gen_bases(64, 7)
    pts = [C.affine_to_ints(x, R.Q_MOD) for x in bs]

    def chain(idx, negs):
        arr = np.ascontiguousarray(bs[idx])
        ng = np.array(negs, dtype=np.uint8)
        o = np.zeros(8, dtype=np.uint64)
        hc.hc_madd_chain(p(arr), ng.ctypes.data_as(ctypes.c_void_p), len(idx), p(o))
        exp = None
        for i, s in zip(idx, negs):
            exp = R.g1_add(exp, R.g1_neg(pts[i]) if s else pts[i])
        assert C.affine_to_ints(o, R.Q_MOD) == exp, (idx, negs)

    chain([0], [0]); chain([0, 0], [0, 0]); chain([0, 0], [0, 1]); chain([0, 0, 0], [0, 1, 0])
    chain([3, 4, 3, 4, 3], [0, 0, 0, 0, 1]); chain([5, 5, 5, 5], [1, 1, 1, 1]); chain([0, 1, 2], [0, 0, 0])
    for _ in range(20):
        k = rnd.randrange(1, 40)
        chain([rnd.randrange(64) for _ in range(k)], [rnd.randrange(2) for _ in range(k)])
    for n in (0, 1, 2, 3, 7, 16, 33):
        idx = [rnd.randrange(8) for _ in range(n)]
        arr = np.ascontiguousarray(bs[idx]) if n else np.zeros((0, 8), dtype=np.uint64)
        o = np.zeros(8, dtype=np.uint64)
        hc.hc_add_tree(p(arr), n, p(o))
        exp = None
        for i in idx:
            exp = R.g1_add(exp, pts[i])
        assert C.affine_to_ints(o, R.Q_MOD) == exp, n


def test_binary_gcd_inversion(hc):
    """fe_inv_bgcd (msm_ba.cuh): the ALU-pipe inversion the batched-affine accumulation uses once per block and round."""
    rnd = random.Random(5)
    for mod, nm in ((R.Q_MOD, "fq"), (R.R_MOD, "fr")):
        vals = [0, 1, 2, 3, mod - 1, mod - 2, (mod - 1) // 2, (mod + 1) // 2, 1 << 253, (1 << 253) - 1, (1 << 256) % mod,
                pow(1 << 256, -1, mod)] + [1 << b for b in range(1, 253, 17)] + [rnd.randrange(1, mod) for _ in range(400)]
        for a in vals:
            o = np.zeros(4, dtype=np.uint64)
            getattr(hc, f"hc_{nm}_inv_bgcd")(p(C.mont_from_int(a, mod)), p(o))
            assert C.int_from_mont(o, mod) == (pow(a, -1, mod) if a else 0), (nm, hex(a))


def test_batched_affine_bucket(hc):
    """One bucket through msm_ba.cuh's phases (pairs -> shared inversion -> affine sums -> mixed additions), 1 and 2 rounds:
    random buckets, every length around the pairing / KCAP boundaries, and the pairs the affine formula cannot take
    (equal points, P and -P, identity bases) in every position."""
    rnd = random.Random(6)
    bs = C.gen_bases(96, 5)
    pts = [C.affine_to_ints(x, R.Q_MOD) for x in bs]
    ident = np.zeros(8, dtype=np.uint64)

    def bucket(idx, negs, rounds):
        arr = np.ascontiguousarray(np.stack([ident if i is None else bs[i] for i in idx])) if idx else np.zeros((0, 8), dtype=np.uint64)
        ng = np.array(negs, dtype=np.uint8)
        o = np.zeros(8, dtype=np.uint64)
        hc.hc_msm_ba_bucket(p(arr), ng.ctypes.data_as(ctypes.c_void_p), len(idx), rounds, p(o))
        exp = None
        for i, s in zip(idx, negs):
            if i is not None:
                exp = R.g1_add(exp, R.g1_neg(pts[i]) if s else pts[i])
        assert C.affine_to_ints(o, R.Q_MOD) == exp, (idx, negs, rounds)

    for rounds in (1, 2):
        bucket([], [], rounds)
        bucket([0], [1], rounds)
        bucket([0, 1], [0, 0], rounds)
        bucket([0, 0], [0, 0], rounds)                       # doubling
        bucket([0, 0], [0, 1], rounds)                       # P + (-P)
        bucket([None, 3], [0, 0], rounds); bucket([3, None], [1, 0], rounds); bucket([None, None], [0, 0], rounds)
        bucket([0, 1, 0, 1], [0, 0, 0, 0], rounds)           # round 1 meets two equal sums
        bucket([0, 1, 0, 1], [0, 0, 1, 1], rounds)           # ... and two opposite ones
        bucket([2, 2, 2, 2, 2, 2, 2], [0] * 7, rounds)
        bucket([4, 5, 4, 5, 4, 5, None, 7, 7, 8], [0, 0, 1, 1, 0, 0, 0, 0, 1, 0], rounds)
        for n in list(range(1, 12)) + [31, 32, 33, 63, 64, 65, 127, 128, 129, 130, 131, 200]:
            bucket([rnd.randrange(96) for _ in range(n)], [rnd.randrange(2) for _ in range(n)], rounds)
        for _ in range(30):                                  # few distinct points: collisions everywhere
            n = rnd.randrange(2, 60)
            bucket([rnd.choice([None, 0, 1, 2, 3]) for _ in range(n)], [rnd.randrange(2) for _ in range(n)], rounds)


def test_stream_copy(hc):
    """The staging ring's copy (non-temporal stores above 4 KiB, csrc/streamcopy.hpp) for every alignment of source and
    destination and lengths around the 32 / 128-byte steps; the bytes next to the destination stay untouched."""
    hc.hc_stream_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
    rng = np.random.default_rng(11)
    src = rng.integers(0, 256, 1 << 17, dtype=np.uint8)
    for so in (0, 1, 7, 31, 33):
        for do in (0, 1, 16, 31, 32, 63):
            for ln in (0, 1, 31, 4095, 4096, 4097, 4096 + 127, 4096 + 128, 4096 + 129, 50000, 65536 + 5):
                dst = np.full(ln + 256, 0xA5, dtype=np.uint8)
                base = (-dst.ctypes.data) % 64                     # a 64-byte aligned origin inside dst
                hc.hc_stream_copy(dst.ctypes.data + base + do, src.ctypes.data + so, ln)
                got = dst[base + do: base + do + ln]
                assert np.array_equal(got, src[so: so + ln]), (so, do, ln)
                assert (dst[: base + do] == 0xA5).all() and (dst[base + do + ln:] == 0xA5).all(), (so, do, ln)


def hc_ntt(hc, a, log_n, omega, scale=None, in_len=None, out_len=None, in_pat=None, out_pat=None):
    N = 1 << log_n
    out = np.zeros((N, 4), dtype=np.uint64)
    a = np.ascontiguousarray(a)
    hc.hc_ntt(p(a), a.shape[0] if in_len is None else in_len, p(out), N if out_len is None else out_len, log_n,
              p(omega), p(scale), p(in_pat), p(out_pat))
    return out


@pytest.mark.parametrize("k,split", [(k, None) for k in range(1, 15)] +
                         [(12, "4,4,4"), (12, "3,3,3,3"), (13, "5,4,4"), (12, "9,3"), (12, "3,9"), (15, "7,8"), (11, "8,3"), (16, "8,8")])
def test_ntt_index_logic(hc, k, split, monkeypatch):
    if split:
        monkeypatch.setenv("H2B_NTT_SPLIT", split)
    else:
        monkeypatch.delenv("H2B_NTT_SPLIT", raising=False)
    a = C.random_fr(1 << k, 100 + k)
    w = m(pow(R.ROOT_OF_UNITY, 1 << (28 - k), R.R_MOD))
    assert np.array_equal(hc_ntt(hc, a, k, w), C.best_fft(a, w, k))
    sc = m(pow(1 << k, -1, R.R_MOD))
    wi = m(pow(R.ROOT_OF_UNITY, -(1 << (28 - k)), R.R_MOD))
    assert np.array_equal(hc_ntt(hc, a, k, wi, scale=sc), C.lagrange_to_coeff(a, k, wi, sc))


@pytest.mark.parametrize("j,k", [(5, 3), (3, 4), (5, 6), (9, 8), (5, 10), (6, 12)])
def test_ntt_fused_domain_ops(hc, j, k, monkeypatch):
    monkeypatch.delenv("H2B_NTT_SPLIT", raising=False)
    d = R.EvaluationDomain(j, k)
    ek, n = d.extended_k, 1 << k
    a = C.random_fr(n, k * 31 + j)
    pat_in = np.stack([m(d.g_coset), m(d.g_coset_inv)])
    pat_out = np.stack([m(d.g_coset_inv), m(d.g_coset)])
    ext = hc_ntt(hc, a, ek, m(d.extended_omega), in_len=n, in_pat=pat_in)
    exp = C.coeff_to_extended(a, k, ek, m(d.extended_omega), m(d.g_coset), m(d.g_coset_inv))
    assert np.array_equal(ext, exp)
    ol = n * (j - 1)
    back = hc_ntt(hc, exp, ek, m(d.extended_omega_inv), scale=m(d.extended_ifft_divisor), out_len=ol, out_pat=pat_out)
    exp2 = C.extended_to_coeff(exp, ek, m(d.extended_omega_inv), m(d.extended_ifft_divisor), m(d.g_coset), m(d.g_coset_inv), ol)
    assert np.array_equal(back[:ol], exp2) and not back[ol:].any()


@pytest.mark.parametrize("n,c,rt", [(1, 4, 0), (2, 4, 0), (33, 4, 4), (200, 5, 4), (1000, 8, 16), (1000, 13, 128), (300, 16, 128), (77, 20, 128)])
def test_msm_pipeline_logic(hc, n, c, rt):
    s = C.random_fr(n, n * 7 + c)
    b = C.gen_bases(n, 11)
    if n > 10:
        s[3] = 0; s[4] = m(1); s[5] = m(R.R_MOD - 1); b[6] = 0; b[8] = b[7]; s[8] = s[7]
    o = np.zeros(12, dtype=np.uint64)
    hc.hc_msm(p(s), p(b), n, c, rt, p(o))
    assert np.array_equal(o[:8], C.to_affine(C.best_multiexp(s, b)))
