"""tcgen05 split-tf32 Gram vs a plain fp64 PyTorch reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,s,d,dtype", [(2, 64, 128, torch.float32), (3, 50, 50, torch.float64),
                                         (1, 700, 300, torch.float32), (2, 4500, 256, torch.float32),
                                         # dT >= 512: CTA-pair kernel (cta_group::2, 256x256 tiles);
                                         # 600 -> 640 rows = 2.5 tiles (TMA zero-fills the rest)
                                         (2, 300, 512, torch.float32), (1, 5000, 600, torch.float32),
                                         (1, 96, 1100, torch.float64)])
def test_gram_tc_matches_fp64(N, s, d, dtype):
    from gadmm_b200.ops import linalg as LA
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(N * 1000 + s)
    X = torch.randn(N, s, d, dtype=dtype, generator=g).to(dev)
    G = LA.gram_tc(X)
    dp = G.shape[1]
    Xd = X.double()
    ref = Xd.transpose(1, 2) @ Xd
    scale = ref.abs().max()
    err = (G[:, :d, :d] - ref).abs().max() / scale
    assert err < 3e-5, float(err)     # fp32-grade: tf32 split + tensor-core fp32 accumulation
    assert torch.count_nonzero(G[:, d:, :]) == 0 and torch.count_nonzero(G[:, :, d:]) == 0
    torch.testing.assert_close(G, G.transpose(1, 2), rtol=0, atol=0)   # exactly symmetric
    # K-segments are summed in fixed point (integer atomics): bit-identical run to run
    for _ in range(2):
        assert torch.equal(G, LA.gram_tc(X))


def test_gram_tc_weighted_and_accumulate():
    from gadmm_b200.ops import linalg as LA
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(5)
    X = torch.randn(2, 333, 200, dtype=torch.float64, generator=g).to(dev)
    w = torch.rand(2, 333, dtype=torch.float64, generator=g).to(dev)
    G = LA.gram_tc(X, w=w)
    ref = (X * w.unsqueeze(2)).transpose(1, 2) @ X
    assert ((G[:, :200, :200] - ref).abs().max() / ref.abs().max()) < 3e-5
    G2 = LA.gram_tc(X, w=w, out=G.clone(), accumulate=True)
    assert ((G2[:, :200, :200] - 2 * ref).abs().max() / ref.abs().max()) < 6e-5


def test_scaled_builder_tc_matches_fp64_path():
    """The tensor-core setup path (Philox shards + tcgen05 Gram, sample streaming) produces a
    solvable problem: GADMM on it converges to its own optimum like the fp64 path does."""
    import numpy as np
    from gadmm_b200.engine.chain import ChainEngine
    from gadmm_b200.engine.scaled import build_scaled_linreg
    res = {}
    for gram in ("f64", "tc"):
        eng = ChainEngine(6, 200, 1.0, 0.0, 1e-6, 3000, stop=True)
        obj0 = build_scaled_linreg(eng, 700, seed=1, gram=gram, s_chunk=256)
        eng.set_path(np.arange(6))
        it = eng.run(3000)
        obj, loss, it, tv = eng.results(it)
        assert it < 3000 and loss[-1] < 1e-6 and obj0 > 0
        res[gram] = (it, obj0)
    # different generators (torch vs Philox) -> different data, same statistics
    assert abs(res["f64"][1] - res["tc"][1]) / res["f64"][1] < 0.5
