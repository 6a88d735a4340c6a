"""gadmm_b200 -- Blackwell-native decentralised convex optimisation (GADMM family).

Public entry points keep the names, argument order and return tuples of the MATLAB reference
(``dayoladejo/GADMM``); see :mod:`gadmm_b200.api`.  The compute path is hand-written sm_100a CUDA
(``csrc/``, loaded through a C ABI by :mod:`gadmm_b200.ops`) driven by a PyTorch runtime with one
process per GPU.
"""
from .api import (GD_DGD_LAG, GD_DGD_LAG_logistic, calc_cost, dual_averaging,
                  dual_averaging_logisticReg, dynamic_group_ADMM_closedForm,
                  dynamic_group_ADMM_closedForm_v0, findPath, findPath2, group_ADMM,
                  group_ADMM_closedForm, group_ADMM_logistic, group_ADMM_logistic_GD, logReg_GD,
                  opt_sol, opt_sol_closedForm, opt_sol_logistic, standard_ADMM, standared_ADMM,
                  static_group_ADMM_closedForm)
from .results import Result, cumulative_com, error_percent

__version__ = "0.1.0"

__all__ = ["GD_DGD_LAG", "GD_DGD_LAG_logistic", "calc_cost", "dual_averaging",
           "dual_averaging_logisticReg", "dynamic_group_ADMM_closedForm",
           "dynamic_group_ADMM_closedForm_v0", "findPath", "findPath2", "group_ADMM",
           "group_ADMM_closedForm", "group_ADMM_logistic", "group_ADMM_logistic_GD", "logReg_GD",
           "opt_sol", "opt_sol_closedForm", "opt_sol_logistic", "standard_ADMM", "standared_ADMM",
           "static_group_ADMM_closedForm", "Result", "cumulative_com", "error_percent"]
