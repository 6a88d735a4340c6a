"""MAT-file (Level 5) and text-dataset IO with the reference's layouts.

* ``inputData.mat``  = {``X_fede``: (N*s) x d double, ``y_fede``: (N*s) x 1} -- MATLAB compacts the
  integer-valued labels to int16, ``scipy.io.loadmat`` returns int16, we cast to float64
  (SURVEY.md section 2.1).
* ``optimalSol.mat`` = {``obj0``: 1x1 double, ``opt_obj``: 1 x num_iter double = obj0*ones}
  (LinearRegression_Synthetic.m:60; consumer GD_DGD_LAG_logistic.m:79).
* datasets ``dataNN/{data,y}.txt`` (LinearRegression_Synthetic.m:6-11) are plain whitespace text.

No HDF5/v7.3: ``h5py`` is absent and the fixtures are v5.
"""
from __future__ import annotations

import csv
import json
import os

import numpy as np
import scipy.io as sio


def load_input_data(path: str):
    """Return ``(X_fede float64 (Ns,d), y_fede float64 (Ns,))``."""
    m = sio.loadmat(path)
    X = np.ascontiguousarray(m["X_fede"], dtype=np.float64)
    y = np.ascontiguousarray(m["y_fede"], dtype=np.float64).reshape(-1)
    return X, y


def save_input_data(path: str, X_fede, y_fede, compact_labels: bool = True):
    """Write the ``inputData.mat`` layout.  ``compact_labels`` stores integer-valued labels as
    int16 like MATLAB's integer compaction does, so a round trip is class-identical."""
    X = np.asarray(X_fede, dtype=np.float64)
    y = np.asarray(y_fede).reshape(-1, 1)
    if compact_labels and np.all(np.equal(np.mod(y, 1), 0)) and np.all(np.abs(y) < 2 ** 15):
        y = y.astype(np.int16)
    else:
        y = y.astype(np.float64)
    sio.savemat(path, {"X_fede": X, "y_fede": y}, format="5", do_compression=False)


def load_optimal_sol(path: str):
    """Return ``(obj0 float, opt_obj float64 (num_iter,))``."""
    m = sio.loadmat(path)
    return float(np.asarray(m["obj0"]).reshape(-1)[0]), \
        np.asarray(m["opt_obj"], dtype=np.float64).reshape(-1)


def save_optimal_sol(path: str, obj0: float, num_iter: int):
    sio.savemat(path, {"obj0": np.array([[float(obj0)]]),
                       "opt_obj": float(obj0) * np.ones((1, int(num_iter)))},
                format="5", do_compression=False)


def load_text_dataset(directory: str):
    """``load('dataNN/data.txt')`` / ``load('dataNN/y.txt')`` (user-supplied; not in the repo)."""
    X = np.loadtxt(os.path.join(directory, "data.txt"), dtype=np.float64, ndmin=2)
    y = np.loadtxt(os.path.join(directory, "y.txt"), dtype=np.float64).reshape(-1)
    return X, y


def save_curves(path: str, curves: dict):
    """Emit curves as ``.mat`` (v5), ``.csv`` or ``.json`` by extension -- the stand-in for the
    reference's ``semilogy`` figures (no graphics stack on this box)."""
    ext = os.path.splitext(path)[1].lower()
    arrs = {k: np.asarray(v, dtype=np.float64) for k, v in curves.items()}
    if ext == ".mat":
        sio.savemat(path, {k: v.reshape(1, -1) if v.ndim <= 1 else v for k, v in arrs.items()},
                    format="5")
    elif ext == ".json":
        with open(path, "w") as f:
            json.dump({k: v.tolist() for k, v in arrs.items()}, f)
    elif ext == ".csv":
        keys = list(arrs)
        n = max((a.size for a in arrs.values()), default=0)
        with open(path, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["iter"] + keys)
            for i in range(n):
                w.writerow([i + 1] + [repr(float(arrs[k].ravel()[i])) if i < arrs[k].size else ""
                                      for k in keys])
    else:
        raise ValueError(f"unsupported curve format {ext!r}")


_STATE_ARRAYS = ("theta", "dual", "path", "obj", "loss", "time")


def save_state(path: str, state: dict):
    """Write a solver checkpoint (``return_state=True`` / ``ChainEngine.state()``: ``theta`` d x N,
    ``dual`` d x N, ``iteration``, optional ``path`` and curves so far) as ``.mat`` (v5), ``.npz``
    or ``.pt`` by extension.  The reference keeps this state in function locals and never saves
    it (``%save`` is commented out, GD_DGD_LAG.m:29); SURVEY.md section 5.4."""
    ext = os.path.splitext(path)[1].lower()
    rec = {}
    for k, v in state.items():
        if v is None:
            continue
        rec[k] = np.asarray(v, dtype=np.int64 if k == "path" else np.float64) \
            if k in _STATE_ARRAYS or isinstance(v, np.ndarray) else v
    if ext == ".mat":
        sio.savemat(path, {k: (np.atleast_2d(v) if isinstance(v, np.ndarray) else v)
                           for k, v in rec.items()}, format="5")
    elif ext == ".npz":
        np.savez(path, **rec)
    elif ext == ".pt":
        import torch
        torch.save({k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else v)
                    for k, v in rec.items()}, path)
    else:
        raise ValueError(f"unsupported checkpoint format {ext!r}")
    return path


def load_state(path: str) -> dict:
    """Inverse of :func:`save_state`: arrays come back as float64 (``path`` as int64, 1-based as
    the solvers return it), scalars as Python numbers; pass the result as ``state=``."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".mat":
        raw = {k: v for k, v in sio.loadmat(path).items() if not k.startswith("__")}
    elif ext == ".npz":
        with np.load(path, allow_pickle=False) as z:
            raw = {k: z[k] for k in z.files}
    elif ext == ".pt":
        import torch
        raw = {k: (v.numpy() if hasattr(v, "numpy") else v)
               for k, v in torch.load(path, weights_only=True).items()}
    else:
        raise ValueError(f"unsupported checkpoint format {ext!r}")
    out = {}
    for k, v in raw.items():
        a = np.asarray(v)
        if k in ("theta", "dual"):
            out[k] = np.ascontiguousarray(a, dtype=np.float64)   # MAT files come back column-major
        elif k == "path":
            out[k] = np.asarray(a, dtype=np.int64).reshape(-1)
        elif a.ndim == 0 or a.size == 1:
            x = a.reshape(-1)[0]
            out[k] = int(x) if float(x).is_integer() and k in ("iteration", "inner_steps", "launches", "g0") \
                else float(x)
        else:
            out[k] = np.asarray(a, dtype=np.float64).reshape(-1) if a.ndim <= 2 and 1 in a.shape else a
    return out
