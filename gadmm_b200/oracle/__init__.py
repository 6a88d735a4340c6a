"""CPU float64 oracle: reference semantics of every algorithm."""
from .gadmm import (group_ADMM, group_ADMM_closedForm, group_ADMM_logistic, group_ADMM_logistic_GD,
                    logReg_GD, static_group_ADMM_closedForm, dynamic_group_ADMM_closedForm,
                    dynamic_group_ADMM_closedForm_v0)
from .star import standared_ADMM
from .baselines import GD_DGD_LAG, GD_DGD_LAG_logistic
from .dual_avg import dual_averaging, dual_averaging_logisticReg
from .optimum import opt_sol, opt_sol_closedForm, opt_sol_logistic
