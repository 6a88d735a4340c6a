"""Public API: the reference's entry points with identical argument order and return tuples.

Every function mirrors one ``.m`` file and takes an extra keyword
``backend``:

  ``"oracle"``  CPU float64 restatement of the reference semantics (:mod:`gadmm_b200.oracle`);
  ``"b200"``    the product: hand-written sm_100a kernels, one process per GPU, NVLink P2P halo
                exchange from inside the kernels (:mod:`gadmm_b200.engine`);
  ``"nccl"``    plain ``torch.distributed`` send/recv + all_reduce implementation -- the timed
                baseline the product has to beat (:mod:`gadmm_b200.parallel.nccl_baseline`);
  ``"gloo"``    the same message pattern on CPU/gloo (plumbing, BASELINE.json config 1);
  ``"auto"``    ``"b200"`` when a CUDA device is present, else ``"oracle"``.

Worker ids are 1-based at this boundary (``path``, ``center``), curves are 1-D float64 NumPy
arrays of length ``Iter`` (or ``num_iter`` for the fixed-length baselines).
"""
from __future__ import annotations

from . import oracle as _oracle
from .topology import calc_cost, findPath, findPath2  # noqa: F401  (re-exported)

_GPU_BACKENDS = ("b200", "nccl", "gloo")


def _resolve(backend: str) -> str:
    if backend in (None, "auto"):
        try:
            import torch
            return "b200" if torch.cuda.is_available() else "oracle"
        except Exception:  # pragma: no cover
            return "oracle"
    if backend not in ("oracle",) + _GPU_BACKENDS:
        raise ValueError(f"unknown backend {backend!r}")
    return backend


def _engine():
    from . import engine
    return engine


def _dist(backend):
    from .parallel import dist_backends
    return dist_backends.get(backend)


# ------------------------------------------------------------------------------- GADMM (linear)

def group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                          backend: str = "auto", **kw):
    """``[obj_GADMM, loss_GADMM, Iter, gadmm_time]`` -- group_ADMM_closedForm.m:1."""
    b = _resolve(backend)
    if b == "oracle":
        return _oracle.group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples,
                                             num_iter, obj0, acc, **kw)
    if b == "b200":
        return _engine().gadmm_closed_form(XX, YY, rho, no_workers, num_feature, noSamples,
                                           num_iter, obj0, acc, **kw)
    return _dist(b).gadmm_closed_form(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                      obj0, acc, **kw)


def group_ADMM(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
               backend: str = "auto", **kw):
    """``[obj_GADMM, loss_GADMM, Iter]`` -- group_ADMM.m:1 (exact local solves; 3 outputs)."""
    obj, loss, it = group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples,
                                          num_iter, obj0, acc, backend=backend, **kw)[:3]
    return obj, loss, it


def static_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                 acc, coherence_Time, pathCost_static_matrix,
                                 backend: str = "auto", **kw):
    """5 outputs incl. ``com_cost`` -- static_group_ADMM_closedForm.m:1-2."""
    b = _resolve(backend)
    if b == "oracle":
        return _oracle.static_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature,
                                                    noSamples, num_iter, obj0, acc,
                                                    coherence_Time, pathCost_static_matrix, **kw)
    if b == "b200":
        return _engine().static_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                      obj0, acc, coherence_Time, pathCost_static_matrix, **kw)
    return _dist(b).static_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                 acc, coherence_Time, pathCost_static_matrix, **kw)


def dynamic_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                  acc, path, pathCost, coherence_Time, backend: str = "auto",
                                  **kw):
    """D-GADMM, live re-chaining -- dynamic_group_ADMM_closedForm.m:1-2."""
    b = _resolve(backend)
    if b == "oracle":
        return _oracle.dynamic_group_ADMM_closedForm(XX, YY, rho, no_workers, num_feature,
                                                     noSamples, num_iter, obj0, acc, path,
                                                     pathCost, coherence_Time, **kw)
    if b == "b200":
        return _engine().dynamic_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                       obj0, acc, path, pathCost, coherence_Time, **kw)
    return _dist(b).dynamic_gadmm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                                  acc, path, pathCost, coherence_Time, **kw)


def dynamic_group_ADMM_closedForm_v0(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                     obj0, acc, path_matrix, pathCost_matrix, coherence_Time,
                                     backend: str = "auto", **kw):
    """D-GADMM, pre-generated chains -- dynamic_group_ADMM_closedForm_v0.m:1-2."""
    b = _resolve(backend)
    if b == "oracle":
        return _oracle.dynamic_group_ADMM_closedForm_v0(XX, YY, rho, no_workers, num_feature,
                                                        noSamples, num_iter, obj0, acc,
                                                        path_matrix, pathCost_matrix,
                                                        coherence_Time, **kw)
    if b == "b200":
        return _engine().dynamic_gadmm_v0(XX, YY, rho, no_workers, num_feature, noSamples,
                                          num_iter, obj0, acc, path_matrix, pathCost_matrix,
                                          coherence_Time, **kw)
    return _dist(b).dynamic_gadmm_v0(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                     obj0, acc, path_matrix, pathCost_matrix, coherence_Time, **kw)


# ----------------------------------------------------------------------------- GADMM (logistic)

def group_ADMM_logistic(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                        lambda_logistic, acc, backend: str = "auto", **kw):
    """Exact local solves (CVX in the reference; Newton here) -- group_ADMM_logistic.m:1."""
    b = _resolve(backend)
    if b == "b200":
        # solver="dense": explicit Hessians (tcgen05 Gram + Cholesky), right for small square shards;
        # solver="cg": Hessian-free Newton-CG, for wide shards and for multi-GPU runs (dist=...)
        solver = kw.pop("solver", "auto")
        if solver == "auto":
            many = kw.get("dist") is not None and getattr(kw["dist"], "world", 1) > 1
            solver = "cg" if many or int(num_feature) >= 2048 else "dense"
        if solver == "cg":
            return _engine().gadmm_logistic_cg(XX, YY, rho, no_workers, num_feature, per_samples,
                                               num_iter, obj0, lambda_logistic, acc, **kw)
        kw.pop("dist", None)
        return _engine().gadmm_logistic_newton(XX, YY, rho, no_workers, num_feature, per_samples,
                                               num_iter, obj0, lambda_logistic, acc, **kw)
    kw.pop("newton_tol", None)
    kw.pop("solver", None)
    kw.pop("dist", None)
    if b in ("nccl", "gloo"):
        return _dist(b).gadmm_logistic(XX, YY, rho, no_workers, num_feature, per_samples, num_iter,
                                       obj0, lambda_logistic, acc, **kw)
    return _oracle.group_ADMM_logistic(XX, YY, rho, no_workers, num_feature, per_samples,
                                       num_iter, obj0, lambda_logistic, acc, **kw)


def group_ADMM_logistic_GD(XX, YY, rho, no_workers, num_feature, per_samples, num_iter, obj0,
                           lambda_logistic, acc, gdStep, backend: str = "auto", **kw):
    """Inexact local solves (<=100 frozen-penalty GD steps) -- group_ADMM_logistic_GD.m:1-2."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().gadmm_logistic_gd(XX, YY, rho, no_workers, num_feature, per_samples,
                                           num_iter, obj0, lambda_logistic, acc, gdStep, **kw)
    kw.pop("dist", None)
    if b in ("nccl", "gloo"):
        return _dist(b).gadmm_logistic_gd(XX, YY, rho, no_workers, num_feature, per_samples,
                                          num_iter, obj0, lambda_logistic, acc, gdStep, **kw)
    return _oracle.group_ADMM_logistic_GD(XX, YY, rho, no_workers, num_feature, per_samples,
                                          num_iter, obj0, lambda_logistic, acc, gdStep, **kw)


def logReg_GD(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, backend: str = "auto",
              **kw):
    """``x = logReg_GD(...)`` -- logReg_GD.m:1."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().logreg_gd_single(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep,
                                          **kw)
    return _oracle.logReg_GD(X, Y, x, C1, C2, term1, term2, lambda_logistic, gdStep, **kw)


# ------------------------------------------------------------------------------------ star ADMM

def standared_ADMM(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                   backend: str = "auto", **kw):
    """Standard (star) ADMM -- standared_ADMM.m:1-2 (spelling as in the reference)."""
    b = _resolve(backend)
    if b == "oracle":
        return _oracle.standared_ADMM(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                      obj0, acc, **kw)
    if b == "b200":
        return _engine().star_admm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter,
                                   obj0, acc, **kw)
    return _dist(b).star_admm(XX, YY, rho, no_workers, num_feature, noSamples, num_iter, obj0,
                              acc, **kw)


standard_ADMM = standared_ADMM


# ------------------------------------------------------------------------------------ baselines

def GD_DGD_LAG(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede, y_fede, num_sample,
               backend: str = "auto", **kw):
    """19 outputs -- GD_DGD_LAG.m:1-4."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().gd_dgd_lag(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede,
                                    y_fede, num_sample, **kw)
    if b in ("nccl", "gloo"):
        return _dist(b).gd_dgd_lag(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede,
                                   y_fede, num_sample, **kw)
    return _oracle.GD_DGD_LAG(X, y, num_workers, num_feature, Hmax, num_iter, obj0, X_fede,
                              y_fede, num_sample, **kw)


def GD_DGD_LAG_logistic(X, y, num_workers, num_feature, Hmax, num_iter, X_fede, y_fede,
                        num_sample, lam, accuracy, backend: str = "auto", **kw):
    """20 outputs (last = ``obj1``, the GD-defined optimum) -- GD_DGD_LAG_logistic.m:1-5."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().gd_dgd_lag_logistic(X, y, num_workers, num_feature, Hmax, num_iter,
                                             X_fede, y_fede, num_sample, lam, accuracy, **kw)
    if b in ("nccl", "gloo"):
        return _dist(b).gd_dgd_lag_logistic(X, y, num_workers, num_feature, Hmax, num_iter,
                                            X_fede, y_fede, num_sample, lam, accuracy, **kw)
    return _oracle.GD_DGD_LAG_logistic(X, y, num_workers, num_feature, Hmax, num_iter, X_fede,
                                       y_fede, num_sample, lam, accuracy, **kw)


def dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc, alpha,
                   backend: str = "auto", **kw):
    """dual_averaging.m:1."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter,
                                        obj0, acc, alpha, **kw)
    if b in ("nccl", "gloo"):
        return _dist(b).dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter,
                                       obj0, acc, alpha, **kw)
    return _oracle.dual_averaging(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                                  alpha, **kw)


def dual_averaging_logisticReg(XX, YY, no_workers, num_feature, noSamples, num_iter, obj0, acc,
                               alpha, eta, backend: str = "auto", **kw):
    """dual_averaging_logisticReg.m:1."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().dual_averaging_logistic(XX, YY, no_workers, num_feature, noSamples,
                                                 num_iter, obj0, acc, alpha, eta, **kw)
    if b in ("nccl", "gloo"):
        return _dist(b).dual_averaging_logistic(XX, YY, no_workers, num_feature, noSamples,
                                                num_iter, obj0, acc, alpha, eta, **kw)
    return _oracle.dual_averaging_logisticReg(XX, YY, no_workers, num_feature, noSamples,
                                              num_iter, obj0, acc, alpha, eta, **kw)


# -------------------------------------------------------------------------------------- optimum

def opt_sol_closedForm(XX, YY, backend: str = "auto", **kw):
    """opt_sol_closedForm.m:1 -- returns the optimal objective *value*."""
    b = _resolve(backend)
    if b == "b200":
        return _engine().opt_sol_closed_form(XX, YY, **kw)
    return _oracle.opt_sol_closedForm(XX, YY, **kw)


def opt_sol(XX, YY, size=None, backend: str = "auto", **kw):
    """opt_sol.m:1 (CVX least squares; same minimiser as the normal equations)."""
    if _resolve(backend) == "b200":
        import numpy as _np
        X = _np.asarray(XX, dtype=_np.float64)
        return _engine().opt_sol_closed_form(X[:, : int(size)] if size is not None else X, YY, **kw)
    return _oracle.opt_sol(XX, YY, size, **kw)


def opt_sol_logistic(XX, YY, dim, lam, num_workers, backend: str = "auto", **kw):
    """opt_sol_logistic.m:1 (CVX in the reference; Newton to 1e-12 here; Newton-CG on the GPU)."""
    if _resolve(backend) == "b200":
        return _engine().opt_sol_logistic_gpu(XX, YY, dim, lam, num_workers, **kw)
    return _oracle.opt_sol_logistic(XX, YY, dim, lam, num_workers, **kw)
