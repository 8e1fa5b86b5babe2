"""The C-ABI boundary on a machine without a GPU: the library loads, exports exactly what include/h2b200.h declares,
and fails loudly (no CPU fallback) when asked to compute."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import h2b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "h2b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(h2b_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    names = header_functions()
    assert len(names) >= 25
    lib = h2b200.load()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/h2b200.h but not exported by libh2b200.so"
    assert sorted(h2b200.SYMBOLS) == names, "python binding table and header drifted apart"
    out = subprocess.run(["nm", "-D", "--defined-only", h2b200.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (h2b_[a-z0-9_]+)", out))
    assert set(names) <= exported


def test_rust_sys_crate_declares_every_header_symbol():
    # rust/h2b200-sys is the binding a maintainer adds on the reference side; no cargo here, so the check is textual
    src = open(os.path.join(ROOT, "rust", "h2b200-sys", "src", "lib.rs")).read()
    declared = set(re.findall(r"\bpub fn (h2b_[a-z0-9_]+)\s*\(", src))
    assert declared == set(header_functions())


@pytest.mark.skipif(not os.path.isdir("/root/reference/halo2_backend"), reason="the reference tree is not on this machine")
def test_ntt_binding_patch_applies_to_the_reference():
    """The EvaluationDomain side of the boundary has no trait: it ships as a patch of halo2_backend
    (domain.rs:216-293 + the coset loop of evaluation.rs:344-361).  `patch --dry-run` must accept it (nothing is written)."""
    patch = os.path.join(ROOT, "rust", "patches", "0001-halo2_backend-b200-ntt.patch")
    with open(patch) as fh:
        r = subprocess.run(["patch", "--dry-run", "-p1", "-d", "/root/reference"], stdin=fh, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    text = open(patch).read()
    for needle in ("h2b_ntt(", "h2b_coeff_to_extended(", "h2b_extended_to_coeff(", "h2b_coeff_to_extended_batch(", "is_bn256_fr"):
        assert needle in text
    # every library function the patch calls is declared by the sys crate with that name
    sys_src = open(os.path.join(ROOT, "rust", "h2b200-sys", "src", "lib.rs")).read()
    for fn in set(re.findall(r"sys::(h2b_[a-z0-9_]+)\(", text)):
        assert f"pub fn {fn}(" in sys_src, fn


def test_rust_engine_never_caches_by_address():
    """`msm` must not key resident bases on a (pointer, len) that can be freed and reused."""
    src = open(os.path.join(ROOT, "rust", "h2b200-engine", "src", "lib.rs")).read()
    assert "HashMap" not in src and "h2b_msm_raw" in src
    assert "impl Drop for B200BaseDesc" in src and "h2b_bases_release" in src


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", h2b200.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_product_path_through_the_oracle():
    """Nothing under the package may import or link the oracle (CPU code is test infrastructure only)."""
    pkg = os.path.join(ROOT, "yet-another-halo2-fork_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "h2ref" not in text and "import oracle" not in text and "from oracle" not in text, os.path.join(dirpath, f)
    ldd = subprocess.run(["ldd", h2b200.LIB_PATH], capture_output=True, text=True).stdout
    assert "h2ref" not in ldd


def test_oracle_prover_is_independent():
    """The converse: the prover oracle (and what it builds on) shares no code with the package -- its source never names the package,
    and importing it in a fresh interpreter loads neither the package nor its CUDA library."""
    for name in ("plonk_prover.py", "pyref.py"):
        text = open(os.path.join(ROOT, "oracle", name)).read()
        code = "\n".join(line for line in text.split("\n") if line.lstrip().startswith(("import ", "from ")))
        assert "h2b200" not in code and "yet-another-halo2-fork_b200" not in code and "yet_another" not in code, name
    probe = ("import sys; sys.path.insert(0, %r); from oracle import plonk_prover; "
             "bad = [m for m in sys.modules if m.startswith('h2b200') or 'halo2' in m or m == 'torch' or m == 'numpy']; print(bad)" % ROOT)
    out = subprocess.run([sys.executable, "-c", probe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "[]", out.stdout + out.stderr


@pytest.mark.skipif(h2b200.load().h2b_device_count() > 0, reason="GPU present")
def test_fails_loudly_without_gpu():
    with pytest.raises(h2b200.H2bError) as e:
        h2b200.Context()
    assert e.value.code == -4      # H2B_E_NODEVICE
    with pytest.raises(h2b200.H2bError):
        h2b200.B200Engine()


def test_combine_is_host_only_and_validates():
    lib = h2b200.load()
    out = np.zeros(12, dtype=np.uint64)
    assert lib.h2b_msm_combine(None, 1, 16, 16, out.ctypes.data_as(ctypes.c_void_p)) == -1
    # identity windows combine to the identity (0, 1, 0)
    parts = np.zeros((2, 16, 16), dtype=np.uint64)
    got = h2b200.B200Engine.msm_combine(parts, 16)
    assert not got[:4].any() and not got[8:].any() and h2b200.fields.fq_int(got[4:8]) == 1


def test_domain_constants_without_gpu():
    """EvaluationDomain::new (domain.rs:38-144) constants are pure host code; build them without a context."""
    from oracle import pyref as R
    class NoCtx:      # a stand-in so that no GPU is touched
        pass
    for j, k in [(1, 3), (5, 4), (4, 11), (9, 20), (5, 24)]:
        d = h2b200.EvaluationDomain(j, k, ctx=NoCtx())
        r = R.EvaluationDomain(j, k)
        for name in ("n", "k", "extended_k", "omega", "omega_inv", "extended_omega", "extended_omega_inv", "g_coset",
                     "g_coset_inv", "ifft_divisor", "extended_ifft_divisor", "t_evaluations", "barycentric_weight"):
            assert getattr(d, name) == getattr(r, name), name
    with pytest.raises(AssertionError):
        h2b200.EvaluationDomain(9, 26, ctx=NoCtx())     # extended_k > Fr::S (domain.rs:54)
