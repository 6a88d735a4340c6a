"""Problem families ("models") the framework optimises: per-worker least squares and
l2-regularised logistic regression.  See :mod:`gadmm_b200.models.problems`."""
from .problems import (LinearProblem, LogisticProblem, shard_rows, stack_shards,
                       linear_objective, logistic_objective, logistic_gradient)

__all__ = ["LinearProblem", "LogisticProblem", "shard_rows", "stack_shards",
           "linear_objective", "logistic_objective", "logistic_gradient"]
