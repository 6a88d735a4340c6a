"""Shared pieces of the experiment drivers (the reference's seven scripts).

Every driver is ``run(cfg) -> dict[str, ndarray]`` plus a CLI (``python -m
gadmm_b200.experiments.<name> --backend b200 --out curves.mat``).  Hyper-parameters are dataclass
fields whose defaults are the literals of the corresponding ``.m`` script; ``--quick`` divides the
baseline iteration caps so that a smoke run finishes in seconds.  Figures are not drawn (no
graphics stack): the arrays the reference plots are returned / written instead.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
from typing import Dict

import numpy as np

from ..data import surrogate_labels, surrogate_targets
from ..io import load_text_dataset, save_curves
from ..results import cumulative_com, error_percent


def add_common_args(ap: argparse.ArgumentParser):
    ap.add_argument("--backend", default="auto", choices=["auto", "oracle", "b200", "nccl", "gloo"])
    ap.add_argument("--out", default=None, help="write curves to .mat / .csv / .json")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--quick", action="store_true", help="shrink iteration caps (smoke run)")
    ap.add_argument("--data-dir", default=None, help="directory holding dataNN/{data,y}.txt")
    ap.add_argument("--fig", default=None, help="write the reference's semilogy figure as an .svg file")


def optimum_backend(backend: str) -> str:
    """Backend for the set-up optimum ``obj0``: the GPU path when the run itself is on the GPU
    engine, the CPU oracle otherwise (nccl / gloo baselines have no optimum solver of their own)."""
    if backend in (None, "auto"):
        try:
            import torch
            return "b200" if torch.cuda.is_available() else "oracle"
        except Exception:  # pragma: no cover
            return "oracle"
    return "b200" if backend == "b200" else "oracle"


def apply_overrides(cfg, args):
    for f in dataclasses.fields(cfg):
        if hasattr(args, f.name) and getattr(args, f.name) is not None:
            setattr(cfg, f.name, getattr(args, f.name))
    return cfg


def dataset_or_surrogate(data_dir, name, n_rows, n_cols, rng, labels=False):
    """Load ``<data_dir>/<name>/{data,y}.txt`` or synthesise a stand-in of the same shape
    (the reference's data files are not shipped, README.md:21-22)."""
    if data_dir and os.path.exists(os.path.join(data_dir, name, "data.txt")):
        X, y = load_text_dataset(os.path.join(data_dir, name))
        return X, y, False
    X = rng.standard_normal((n_rows, n_cols))
    y = surrogate_labels(n_rows, rng) if labels else surrogate_targets(n_rows, rng)
    return X, y, True


def finish(curves: Dict[str, np.ndarray], out, summary: dict, fig=None):
    if out:
        save_curves(out, curves)
    if fig:
        from ..plot import figure_from_curves
        figure_from_curves(curves, fig, title=str(summary.get("driver", "")))
    print(json.dumps(summary))
    return curves


def gadmm_comm(prefix, curves, obj, loss, it, num_workers, opt_obj=None):
    curves[f"obj_{prefix}"] = obj
    curves[f"loss_{prefix}"] = loss
    curves[f"Iter_{prefix}"] = np.array([it], dtype=np.float64)
    curves[f"cumulative_com_{prefix}"] = cumulative_com("GADMM", it, num_workers)
    if opt_obj is not None:
        curves[f"errorPer_{prefix}"] = error_percent(loss, opt_obj)


def baseline_curves(curves, out19, num_workers, obj0):
    names = ["obj_GD", "loss_GD", "time_GD", "obj_DGD", "loss_DGD", "time_DGD", "obj_LAG_PS",
             "loss_LAG_PS", "time_lagPS", "comm_iter_final_LAG_PS", "obj_LAG_WK", "loss_LAG_WK",
             "time_lagWK", "comm_iter_final_LAG_WK", "obj_cyclic_IAG", "loss_cyclic_IAG",
             "obj_R_IAG", "loss_R_IAG"]
    for k, nm in enumerate(names):
        curves[nm] = np.asarray(out19[k])
    curves["stepsize2"] = np.array([out19[18]])
    curves["cumulative_com_GD"] = cumulative_com("GD", len(out19[0]), num_workers)
    curves["cumulative_com_DGD"] = cumulative_com("DGD", len(out19[3]), num_workers)
    curves["cumulative_com_LAG_PS"] = cumulative_com("LAG_PS", len(out19[6]), num_workers, out19[9])
    curves["cumulative_com_LAG_WK"] = cumulative_com("LAG_WK", len(out19[10]), num_workers, out19[13])
    if obj0:
        curves["errorPer_GD"] = error_percent(out19[1], obj0)
        curves["errorPer_LAG_PS"] = error_percent(out19[7], obj0)
        curves["errorPer_LAG_WK"] = error_percent(out19[11], obj0)
