from .synth import (linreg_synthetic, logreg_synthetic, split_dataset, surrogate_targets,
                    surrogate_labels, hmax_linear, hmax_logistic)

__all__ = ["linreg_synthetic", "logreg_synthetic", "split_dataset", "surrogate_targets",
           "surrogate_labels", "hmax_linear", "hmax_logistic"]
