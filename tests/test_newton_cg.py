"""Batched Newton-CG (the local solver of the scaled logistic GADMM) on CPU tensors."""
import numpy as np
import torch

from gadmm_b200.engine.logistic_scaled import _pick_ctas_per_shard, newton_cg
from gadmm_b200.oracle.gadmm import _logistic_local_newton


def test_newton_cg_matches_dense_newton():
    rng = np.random.default_rng(0)
    B, s, d = 5, 30, 12
    X = rng.standard_normal((B, s, d))
    y = np.where(rng.standard_normal((B, s)) > 0, 1.0, -1.0)
    lin = rng.standard_normal((B, d))
    mu = np.array([0.3, 0.5, 1.0, 0.05, 2.0])
    Xt, yt, lint = torch.from_numpy(X), torch.from_numpy(y), torch.from_numpy(lin)
    mut = torch.from_numpy(mu).unsqueeze(1)
    state = {}

    def evaluate(x):
        z = torch.einsum("bsd,bd->bs", Xt, x) * yt
        sig = torch.sigmoid(-z)
        state["D"] = sig * (1 - sig)
        f = torch.nn.functional.softplus(-z).sum(1) + 0.5 * mut.squeeze(1) * (x * x).sum(1) \
            + (lint * x).sum(1)
        g = -torch.einsum("bsd,bs->bd", Xt, yt * sig) + mut * x + lint
        return f, g

    def hessvec(p):
        u = torch.einsum("bsd,bd->bs", Xt, p) * state["D"]
        return torch.einsum("bsd,bs->bd", Xt, u) + mut * p

    stats = {}
    x, f, g = newton_cg(evaluate, hessvec, torch.zeros((B, d), dtype=torch.float64), tol=1e-11,
                        stats=stats)
    assert float(g.norm(dim=1).max()) < 1e-9
    assert stats["hessvecs"] > 0 and stats["evals"] > 1
    for b in range(B):
        ref = _logistic_local_newton(X[b], y[b], np.zeros(d), lin[b], mu[b], np.zeros(d), 0.0, False,
                                     tol=1e-12)
        np.testing.assert_allclose(x[b].numpy(), ref, rtol=1e-7, atol=1e-9)


def test_newton_cg_already_converged_rows_stay_put():
    A = torch.diag(torch.tensor([1.0, 2.0, 3.0], dtype=torch.float64))
    b = torch.tensor([[1.0, 1.0, 1.0], [0.0, 0.0, 0.0]], dtype=torch.float64)

    def evaluate(x):
        return 0.5 * ((x @ A) * x).sum(1) - (b * x).sum(1), x @ A - b

    x, f, g = newton_cg(evaluate, lambda p: p @ A, torch.zeros((2, 3), dtype=torch.float64), tol=1e-12)
    np.testing.assert_allclose(x[0].numpy(), [1.0, 0.5, 1.0 / 3.0], rtol=1e-12)
    assert float(x[1].abs().max()) == 0.0


def test_ctas_per_shard_fills_waves():
    assert _pick_ctas_per_shard(148, 2048) == 1
    g = _pick_ctas_per_shard(256, 2048)
    assert 1 <= g <= 16 and -(-256 * g // 148) / g <= 1.76
    assert _pick_ctas_per_shard(3, 8) == 1
