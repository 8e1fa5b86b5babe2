"""The C++ host mirror of the reference's operator interface (include/h2b200.hpp: B200Engine = MsmAccel, ParamsKZG::commit*,
EvaluationDomain) built with g++ against libh2b200.so and driven by tests/cpp/host_mirror_test.cpp, which compares every result
with the CPU oracle (oracle/libh2ref.so).  The reference's host is compiled code; this is the compiled host this image can build
(rust/h2b200-engine is the same binding as Rust source).

CPU part: the program builds and links, its host-side Fr arithmetic equals the oracle's, EvaluationDomain::new's constants equal
the python-int oracle's, and a Context without a device throws.  GPU part: MSM / commit / NTT parity through the mirror."""
import os
import shutil
import subprocess

import pytest

from oracle import cref, pyref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "yet-another-halo2-fork_b200")


@pytest.fixture(scope="module")
def binary(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    import h2b200
    h2b200.load()                                   # fails loudly if the CUDA library has not been built
    cref.lib()                                      # builds oracle/libh2ref.so if needed
    out = str(tmp_path_factory.mktemp("cpp") / "host_mirror_test")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp"), "-o", out,
           "-L", PKG, "-lh2b200", "-L", os.path.join(ROOT, "oracle"), "-lh2ref",
           "-Wl,-rpath," + PKG, "-Wl,-rpath," + os.path.join(ROOT, "oracle")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def _run(binary, *args):
    r = subprocess.run([binary, *args], capture_output=True, text=True, timeout=600)
    return r.returncode, r.stdout + r.stderr


def test_header_is_plain_c_too(tmp_path):
    """include/h2b200.h is the C ABI: a C compiler takes it as it is."""
    src = tmp_path / "t.c"
    src.write_text('#include "h2b200.h"\nint main(void) { return h2b_device_count() < 0; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o",
                        str(tmp_path / "t.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_host_arithmetic_and_domain_identities(binary):
    rc, out = _run(binary, "host")
    assert rc == 0 and "PASSED" in out and "FAIL" not in out, out


@pytest.mark.parametrize("j,k", [(2, 4), (5, 8), (6, 20), (5, 24), (9, 3), (1, 7)])
def test_domain_constants_match_the_oracle(binary, j, k):
    """EvaluationDomain::new in C++ == oracle/pyref.py's restatement of domain.rs:38-144, field by field."""
    rc, out = _run(binary, "consts", str(j), str(k))
    assert rc == 0, out
    got = {}
    for line in out.strip().split("\n"):
        name, value = line.split()
        got.setdefault(name, []).append(int(value, 0))
    d = R.EvaluationDomain(j, k)
    assert got["extended_k"] == [d.extended_k] and got["quotient_poly_degree"] == [d.quotient_poly_degree]
    for name in ("omega", "omega_inv", "extended_omega", "extended_omega_inv", "g_coset", "g_coset_inv", "ifft_divisor",
                 "extended_ifft_divisor", "barycentric_weight"):
        assert got[name] == [getattr(d, name)], name
    assert got["t_evaluation"] == list(d.t_evaluations)


def test_context_without_a_device_throws(binary):
    import h2b200
    if h2b200.load().h2b_device_count() > 0:
        pytest.skip("a device is present")
    rc, out = _run(binary, "nodevice")
    assert rc == 0 and "PASSED" in out, out


@pytest.mark.gpu
def test_mirror_matches_the_oracle_on_the_gpu(binary):
    rc, out = _run(binary, "gpu")
    assert rc == 0 and "PASSED" in out and "FAIL" not in out, out
    assert "launches" in out
