import numpy as np
import pytest
import torch

import gadmm_b200 as g

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N", [1, 2, 24, 50, 777])
def test_find_path_device_matches_host(N):
    from gadmm_b200.ops.topology import find_path_device
    rng = np.random.default_rng(N)
    for model, area, fn in ((0, 50.0, g.findPath), (1, 250.0, g.findPath2)):
        pts = rng.random((N, 2)) * area
        ref = fn(N, points=(pts[:, 0], pts[:, 1]))
        out = find_path_device(torch.from_numpy(pts).cuda(), model)
        np.testing.assert_array_equal(out[0], ref[0])
        np.testing.assert_allclose(out[1], ref[1], rtol=1e-15)
        np.testing.assert_allclose(out[2], ref[2], rtol=1e-15)
        if model == 1:
            np.testing.assert_allclose(out[3], ref[3], rtol=1e-15)
            assert out[4] == ref[4]


def test_philox_fills():
    from gadmm_b200.ops.topology import philox_normal, philox_uniform
    u = philox_uniform(200001, seed=7, scale=250.0)
    assert 0.0 <= float(u.min()) and float(u.max()) < 250.0
    assert abs(float(u.mean()) - 125.0) < 1.0
    assert torch.equal(u, philox_uniform(200001, seed=7, scale=250.0))
    assert not torch.equal(u, philox_uniform(200001, seed=8, scale=250.0))
    for dt in (torch.float64, torch.float32):
        z = philox_normal((1000, 501), seed=3, dtype=dt)
        assert abs(float(z.double().mean())) < 0.01 and abs(float(z.double().std()) - 1.0) < 0.01
        k = float(((z.double()) ** 4).mean())
        assert abs(k - 3.0) < 0.1


def test_find_path_batch_matches_single():
    from gadmm_b200.ops.topology import find_path_batch_device, find_path_device, philox_uniform
    K, N = 5, 97
    paths, cost_sums = find_path_batch_device(K, N, seed=21, stream_id=3, model=1)
    xy = philox_uniform(K * N * 2, seed=21, stream_id=3, scale=250.0).view(K, N, 2)
    for k in range(K):
        ref = g.findPath2(N, points=(xy[k, :, 0].cpu().numpy(), xy[k, :, 1].cpu().numpy()))
        np.testing.assert_array_equal(paths[k].cpu().numpy() + 1, ref[0])
        assert float(cost_sums[k]) == pytest.approx(ref[1].sum(), rel=1e-14)
        one = find_path_device(xy[k].contiguous(), 1, want_grid=False)
        np.testing.assert_array_equal(one[0], ref[0])
