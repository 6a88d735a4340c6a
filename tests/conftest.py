import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")
    config.addinivalue_line("markers", "slow: long-running CPU test")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        n = 0
    for it in items:
        if "gpu" in it.keywords and n == 0:
            it.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in it.keywords and n < 2:
            it.add_marker(pytest.mark.skip(reason="needs >= 2 CUDA devices"))


@pytest.fixture
def small_linear():
    """N=6 workers, s=d=12: the reference's synthetic construction at toy size."""
    from gadmm_b200.data import linreg_synthetic
    rng = np.random.default_rng(7)
    X, y, Hmax, XX, YY = linreg_synthetic(N=6, s=12, growth=1.3, rng=rng,
                                          y=rng.standard_normal(12))
    return dict(X=X, y=y, Hmax=Hmax, XX=XX, YY=YY, N=6, s=12, d=12)


@pytest.fixture
def ref_linear():
    """The reference default shape N=24, s=d=50 (LinearRegression_Synthetic.m:17,21-22)."""
    from gadmm_b200.data import linreg_synthetic
    rng = np.random.default_rng(0)
    X, y, Hmax, XX, YY = linreg_synthetic(N=24, s=50, growth=1.3, rng=rng,
                                          y=rng.standard_normal(50))
    return dict(X=X, y=y, Hmax=Hmax, XX=XX, YY=YY, N=24, s=50, d=50)


@pytest.fixture
def small_logistic():
    from gadmm_b200.data import logreg_synthetic
    rng = np.random.default_rng(3)
    lam = 1e-3
    X, y, Hmax, XX, YY = logreg_synthetic(N=6, s=12, lam=lam, rng=rng)
    return dict(X=X, y=y, Hmax=Hmax, XX=XX, YY=YY, N=6, s=12, d=12, lam=lam)
