"""ctypes binding of ``libgadmm_b200.so`` (C ABI declared in ``include/gadmm_b200.h``).

The library is loaded lazily and *loudly*: on a machine with a CUDA device a missing or stale
shared object is an error, never a silent PyTorch fallback.  Tensors are passed as raw device
pointers; the current torch CUDA stream is used unless a stream is given.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .._native import build as _build
from .._native import lib_path, needs_build

_lib = None

STORE_F64 = 0
STORE_F32 = 1
TILE = 64

_vp = C.c_void_p


class GadmmChain(C.Structure):
    """Mirror of ``struct GadmmChain`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("gid", _vp), ("left", _vp), ("right", _vp), ("color", _vp), ("cmul", _vp), ("Mptr", _vp),
        ("push0", _vp), ("push1", _vp), ("owner", _vp), ("b", _vp), ("hyy", _vp), ("nu", _vp),
        ("theta", _vp), ("flags", _vp), ("theta_peer", _vp), ("flags_peer", _vp),
        ("objpart", _vp), ("objflag", _vp), ("objpart_peer", _vp), ("objflag_peer", _vp),
        ("rowpart", _vp), ("colpart", _vp), ("counters", _vp), ("objw", _vp), ("ctrl", _vp),
        ("obj_hist", _vp), ("loss_hist", _vp), ("t_hist", _vp),
        ("rho", C.c_double), ("obj0", C.c_double), ("acc", C.c_double),
        ("d", C.c_int), ("dp", C.c_int), ("Nglob", C.c_int), ("Nloc", C.c_int), ("rank", C.c_int),
        ("nranks", C.c_int), ("store", C.c_int), ("max_iter", C.c_int), ("pdl", C.c_int),
        ("split_finish", C.c_int), ("wready", _vp), ("sblk", _vp), ("fin_counters", _vp),
        ("nsplit", C.c_int), ("reserved", C.c_int), ("order", _vp),
        ("sync", _vp), ("nuflag", _vp),
    ]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if os.environ.get("GADMM_B200_LIB"):          # A/B experiments with an alternative build
        path = os.environ["GADMM_B200_LIB"]
    elif needs_build():
        if os.environ.get("GADMM_B200_NO_BUILD") == "1" and os.path.exists(path):
            pass
        else:
            _build()
    _lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    _lib.gadmm_last_error.restype = C.c_char_p
    _lib.gadmm_packed_elems.restype = C.c_size_t
    _lib.gadmm_packed_elems.argtypes = [C.c_int]
    _lib.gadmm_symv_scratch_elems.restype = C.c_size_t
    _lib.gadmm_symv_scratch_elems.argtypes = [C.c_int]
    _lib.gadmm_logistic_xtwx_scratch_elems.restype = C.c_size_t
    _lib.gadmm_logistic_xtwx_scratch_elems.argtypes = [C.c_int, C.c_int, C.c_int]
    _lib.gadmm_logi_graph_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    _lib.gadmm_logi_graph_launch.argtypes = [C.c_void_p, C.c_void_p]
    _lib.gadmm_logi_graph_destroy.argtypes = [C.c_void_p]
    return _lib


class NativeError(RuntimeError):
    pass


def check(code: int):
    if code != 0:
        msg = lib().gadmm_last_error()
        raise NativeError(f"gadmm_b200 native call failed ({code}): "
                          f"{msg.decode(errors='replace') if msg else '?'}")


def ptr(t):
    if t is None:
        return _vp(0)
    if isinstance(t, int):
        return _vp(t)
    assert t.is_cuda and t.is_contiguous(), "native ops need contiguous CUDA tensors"
    return _vp(t.data_ptr())


def stream_ptr(stream=None):
    s = torch.cuda.current_stream() if stream is None else stream
    return _vp(s.cuda_stream)


def padded(d: int) -> int:
    return (d + TILE - 1) // TILE * TILE


def packed_elems(dp: int) -> int:
    return int(lib().gadmm_packed_elems(dp))


def symv_scratch_elems(dp: int) -> int:
    return int(lib().gadmm_symv_scratch_elems(dp))


class GadmmShared(C.Structure):
    """Mirror of ``struct GadmmShared`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("theta", _vp), ("Y", _vp), ("fw", _vp), ("ghat", _vp), ("theta_hat", _vp), ("dnorm", _vp),
        ("Hmax", _vp), ("uni", _vp), ("cumprob", _vp), ("obj_hist", _vp), ("loss_hist", _vp),
        ("comm_hist", _vp), ("t_hist", _vp), ("ctrl", _vp), ("counters", _vp),
        ("xbuf", _vp), ("xflag", _vp), ("xbuf_peer", _vp), ("xflag_peer", _vp), ("dnpart", _vp),
        ("counters2", _vp),
        ("step", C.c_double), ("thr", C.c_double), ("obj0", C.c_double), ("acc", C.c_double),
        ("N", C.c_int), ("d", C.c_int), ("dp", C.c_int), ("kind", C.c_int), ("max_iter", C.c_int),
        ("logistic", C.c_int), ("Nglob", C.c_int), ("gid0", C.c_int), ("rank", C.c_int),
        ("nranks", C.c_int), ("split_server", C.c_int), ("reserved", C.c_int),
    ]


class GadmmOwn(C.Structure):
    """Mirror of ``struct GadmmOwn`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("Theta", _vp), ("Y", _vp), ("fw", _vp), ("Z", _vp), ("obj_hist", _vp), ("loss_hist", _vp),
        ("t_hist", _vp), ("ctrl", _vp), ("counters", _vp),
        ("Ytab", _vp), ("Ztab", _vp), ("rowflag", _vp), ("Ytab_peer", _vp), ("Ztab_peer", _vp),
        ("rowflag_peer", _vp), ("fx", _vp), ("fxflag", _vp), ("fx_peer", _vp), ("fxflag_peer", _vp),
        ("ftot", _vp),
        ("step", C.c_double), ("obj0", C.c_double), ("acc", C.c_double),
        ("N", C.c_int), ("d", C.c_int), ("dp", C.c_int), ("max_iter", C.c_int), ("jacobi", C.c_int),
        ("Nglob", C.c_int), ("gid0", C.c_int), ("rank", C.c_int), ("nranks", C.c_int),
        ("reserved", C.c_int),
    ]


FO_GD, FO_LAG_PS, FO_LAG_WK, FO_IAG_CYCLIC, FO_IAG_RANDOM = range(5)


class GadmmStar(C.Structure):
    """Mirror of ``struct GadmmStar`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("Mptr", _vp), ("Ms", _vp), ("b", _vp), ("hyy", _vp), ("bs", _vp), ("lam", _vp),
        ("theta", _vp), ("z", _vp), ("U", _vp), ("fw", _vp), ("ubuf", _vp), ("uflag", _vp),
        ("ubuf_peer", _vp), ("uflag_peer", _vp), ("rowpart", _vp), ("colpart", _vp),
        ("counters", _vp), ("ctrl", _vp), ("obj_hist", _vp), ("loss_hist", _vp), ("t_hist", _vp),
        ("ubuf_mc", _vp), ("z_mc", _vp), ("zobj", _vp), ("zobj_mc", _vp), ("zflag", _vp),
        ("zflag_peer", _vp),
        ("rho", C.c_double), ("obj0", C.c_double), ("acc", C.c_double), ("hyy_s", C.c_double),
        ("d", C.c_int), ("dp", C.c_int), ("N", C.c_int), ("Nw", C.c_int), ("rank", C.c_int),
        ("nranks", C.c_int), ("store", C.c_int), ("max_iter", C.c_int), ("nvls", C.c_int),
        ("server_rank", C.c_int), ("nsplit", C.c_int), ("reserved", C.c_int), ("zblk", _vp),
    ]


class GadmmLogistic(C.Structure):
    """Mirror of ``struct GadmmLogistic`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("X", _vp), ("y", _vp), ("theta", _vp), ("lam", _vp), ("fw", _vp), ("inner_steps", _vp),
        ("ctrl", _vp), ("counters", _vp), ("obj_hist", _vp), ("loss_hist", _vp), ("t_hist", _vp),
        ("rho", C.c_double), ("lam_log", C.c_double), ("gd_step", C.c_double),
        ("inner_tol", C.c_double), ("obj0", C.c_double), ("acc", C.c_double),
        ("N", C.c_int), ("s", C.c_int), ("d", C.c_int), ("dp", C.c_int), ("max_iter", C.c_int),
        ("max_inner", C.c_int), ("n0", C.c_int), ("nloc", C.c_int),
    ]


class GadmmLogi(C.Structure):
    """Mirror of ``struct GadmmLogi`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("X", _vp), ("y", _vp), ("Dw", _vp), ("table", _vp), ("flags", _vp), ("table_peer", _vp),
        ("flags_peer", _vp), ("objpart", _vp), ("objflag", _vp), ("objpart_peer", _vp),
        ("objflag_peer", _vp), ("lam", _vp),
        ("x", _vp), ("g", _vp), ("p", _vp), ("r", _vp), ("q", _vp), ("hq", _vp), ("xtry", _vp),
        ("lin", _vp), ("vlast", _vp),
        ("mu", _vp), ("f", _vp), ("gn", _vp), ("eta", _vp), ("rs", _vp), ("t", _vp), ("slope", _vp),
        ("losstry", _vp), ("losslast", _vp), ("live", _vp), ("objw", _vp), ("scratch", _vp),
        ("ctrl", _vp), ("cond", _vp), ("counters", _vp), ("obj_hist", _vp), ("loss_hist", _vp),
        ("t_hist", _vp),
        ("rho", C.c_double), ("lam_log", C.c_double), ("obj0", C.c_double), ("acc", C.c_double),
        ("tol", C.c_double),
        ("N", C.c_int), ("Nloc", C.c_int), ("g0", C.c_int), ("s", C.c_int), ("d", C.c_int),
        ("rank", C.c_int), ("nranks", C.c_int), ("store", C.c_int), ("hv_mode", C.c_int),
        ("max_iter", C.c_int), ("max_newton", C.c_int), ("max_cg", C.c_int), ("max_ls", C.c_int),
        ("G", C.c_int * 2),
    ]


class GadmmRechain(C.Structure):
    """Mirror of ``struct GadmmRechain`` (include/gadmm_b200.h)."""
    _fields_ = [
        ("paths", _vp), ("pos", _vp), ("tab", _vp), ("Mptr", _vp), ("owner", _vp), ("theta", _vp),
        ("flags", _vp), ("theta_peer", _vp), ("ctrl", _vp),
        ("base_c1", C.c_longlong), ("base_c2", C.c_longlong), ("stride_bytes", C.c_longlong),
        ("dyn", _vp), ("tau", C.c_longlong),
        ("N", C.c_int), ("Nloc", C.c_int), ("dp", C.c_int), ("rank", C.c_int), ("nranks", C.c_int),
        ("K", C.c_int), ("order", _vp),
    ]
