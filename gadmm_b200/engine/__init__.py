"""GPU engine entry points behind ``backend="b200"`` (see :mod:`gadmm_b200.api`)."""
from .gadmm_api import (dynamic_gadmm, dynamic_gadmm_v0, gadmm_closed_form, static_gadmm,
                        opt_sol_closed_form)
from .chain import ChainEngine

__all__ = ["ChainEngine", "gadmm_closed_form", "static_gadmm", "dynamic_gadmm",
           "dynamic_gadmm_v0", "opt_sol_closed_form"]


def __getattr__(name):
    # late-bound so importing the chain engine does not require every family to be loaded
    import importlib
    if name.startswith("__"):
        raise AttributeError(name)
    families = importlib.import_module(__name__ + ".families")
    if hasattr(families, name):
        return getattr(families, name)
    raise AttributeError(name)
