"""In-tree native library: build (nvcc, sm_100a) and load (ctypes, C ABI of include/gadmm_b200.h)."""
from .build import build, lib_path, needs_build

__all__ = ["build", "lib_path", "needs_build"]
