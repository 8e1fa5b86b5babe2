"""In-tree build of libh2b200.so (sm_100a only) and the host unit-test shim."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libh2b200.so")
HOSTCHECK = os.path.join(HERE, "libh2b200_hostcheck.so")

def _newer(target: str, sources) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources():
    out = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    out.append(os.path.join(HERE, "..", "include", "h2b200.h"))
    return out


def build_lib(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a (in parallel) and link libh2b200.so in-tree."""
    if not force and _newer(LIB, _sources()):
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    cus = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    cflags = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC"]
    if verbose:
        cflags += ["-Xptxas", "-v"]

    def deps(path, seen=None):
        """the file plus every quoted #include reachable from it (so touching msm.cuh does not rebuild the NTT objects)"""
        seen = set() if seen is None else seen
        path = os.path.normpath(path)
        if path in seen or not os.path.exists(path):
            return seen
        seen.add(path)
        with open(path) as fh:
            for line in fh:
                line = line.strip()
                if line.startswith("#include \""):
                    deps(os.path.join(os.path.dirname(path), line.split('"')[1]), seen)
        return seen

    def compile_one(f):
        obj = os.path.join(objdir, f[:-3] + ".o")
        if not force and not verbose and _newer(obj, deps(os.path.join(CSRC, f))):
            return obj
        subprocess.check_call([nvcc] + cflags + ["-c", "-o", obj, os.path.join(CSRC, f)], cwd=HERE)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(cus))) as ex:
        objs = list(ex.map(compile_one, cus))
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs, cwd=HERE)
    return LIB


def build_hostcheck(force: bool = False) -> str:
    if not force and _newer(HOSTCHECK, _sources()):
        return HOSTCHECK
    cmd = ["g++", "-O2", "-x", "c++", "-std=c++17", "-shared", "-fPIC", "-Wno-unknown-pragmas",
           "-DH2B_HOST_EMULATE_PTX", "-o", HOSTCHECK, os.path.join(CSRC, "host_check.cpp")]
    subprocess.check_call(cmd, cwd=HERE)
    return HOSTCHECK


if __name__ == "__main__":
    import sys
    build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_hostcheck(force="--force" in sys.argv)
    print(LIB)
