"""GPU engine entry points behind ``backend="b200"`` (see :mod:`gadmm_b200.api`)."""
from .gadmm_api import (dynamic_gadmm, dynamic_gadmm_v0, gadmm_closed_form, static_gadmm,
                        opt_sol_closed_form)
from .chain import ChainEngine

__all__ = ["ChainEngine", "gadmm_closed_form", "static_gadmm", "dynamic_gadmm",
           "dynamic_gadmm_v0", "opt_sol_closed_form"]


# entry points that live in their own modules, imported on first use (importing the chain engine
# does not load every family)
_LAZY = {
    "star_admm": "star", "StarEngine": "star",
    "gadmm_logistic_gd": "logistic", "gadmm_logistic_newton": "logistic",
    "logreg_gd_single": "logistic", "opt_sol_logistic_gpu": "logistic",
    "gadmm_logistic_cg": "logistic_scaled", "gadmm_logistic_scaled": "logistic_scaled",
    "ScaledLogisticGadmm": "logistic_scaled",
}


def __getattr__(name):
    import importlib
    if name.startswith("__"):
        raise AttributeError(name)
    if name in _LAZY:
        return getattr(importlib.import_module(f"{__name__}.{_LAZY[name]}"), name)
    families = importlib.import_module(__name__ + ".families")
    if hasattr(families, name):
        return getattr(families, name)
    raise AttributeError(name)
