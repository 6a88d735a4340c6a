"""Build ``libgadmm_b200.so`` from ``csrc/*.cu`` with nvcc for sm_100a (cross-compiles without a GPU).

The shared object stays in-tree (git-ignored) so that it travels with the repository snapshot to
the GPU box; nothing is installed into site-packages and there is no JIT cache.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
_CSRC = os.path.join(_ROOT, "csrc")
_INC = os.path.join(_ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def lib_path() -> str:
    return os.path.join(_HERE, "libgadmm_b200.so")


def _sources():
    return sorted(glob.glob(os.path.join(_CSRC, "*.cu")))


def _deps():
    return _sources() + sorted(glob.glob(os.path.join(_CSRC, "*.cuh"))) + \
        sorted(glob.glob(os.path.join(_INC, "*.h")))


def needs_build() -> bool:
    lib = lib_path()
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(f) > t for f in _deps())


def nvcc_path() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; the native library must be built before use")
    return cand


def build(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> str:
    """Compile every ``csrc/*.cu`` (one object per file, in parallel) and link the shared object."""
    if not force and not needs_build():
        return lib_path()
    nvcc = nvcc_path()
    objdir = os.path.join(_HERE, "obj")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in _sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_info else []) + \
            ["-I", _INC, "-I", _CSRC, "-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    log = []
    for src, p in procs:
        out, _ = p.communicate()
        text = out.decode(errors="replace")
        log.append(text)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{text}")
    tmp = lib_path() + ".tmp"
    # the arch flag keeps nvcc's link step from adding an (empty) cubin for its default sm_52
    cmd = [nvcc, "-shared", "--cudart", "shared", "-gencode", "arch=compute_100a,code=sm_100a",
           "-o", tmp] + objs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout.decode(errors="replace"))
    os.replace(tmp, lib_path())
    if ptxas_info or verbose:
        print("\n".join(log), file=sys.stderr)
    return lib_path()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, ptxas_info="--ptxas" in sys.argv))
