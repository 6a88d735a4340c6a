"""gloo world_size=2 worker for tests/test_gloo.py (BASELINE.json config 1: plumbing, no GPU)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(rank, world, port, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    import gadmm_b200 as g
    from gadmm_b200 import oracle as O
    from gadmm_b200.data import linreg_synthetic
    rng = np.random.default_rng(0)
    X, y, Hmax, XX, YY = linreg_synthetic(N=24, s=50, growth=1.3, rng=rng, y=rng.standard_normal(50))
    obj0 = O.opt_sol_closedForm(XX, YY)
    base = (XX, YY, 5.0, 24, 50, 50, 1000, obj0, 1e-4)
    a = O.group_ADMM_closedForm(*base)
    b = g.group_ADMM_closedForm(*base, backend="gloo")
    assert a[2] == b[2], (a[2], b[2])
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    prng = np.random.default_rng(1)
    paths, costs = [], []
    for _ in range(300):
        pa, co, _ = g.findPath(24, rng=prng)
        paths.append(pa)
        costs.append(co)
    a = O.dynamic_group_ADMM_closedForm_v0(*base, np.array(paths), np.array(costs), 10)
    b = g.dynamic_group_ADMM_closedForm_v0(*base, np.array(paths), np.array(costs), 10, backend="gloo")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    np.testing.assert_allclose(a[4], b[4], rtol=1e-12)
    a = O.standared_ADMM(XX, YY, 1.0, 24, 50, 50, 2000, obj0, 1e-4)
    b = g.standared_ADMM(XX, YY, 1.0, 24, 50, 50, 2000, obj0, 1e-4, backend="gloo")
    assert a[2] == b[2]
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9)
    # first-order baselines with the workers split over the two ranks (all_gather / all_reduce)
    rb = np.random.default_rng(3)
    Nb, sb, db = 7, 12, 9
    Xb = rb.standard_normal((Nb * sb, db))
    Yb = Xb @ rb.standard_normal(db) + 0.1 * rb.standard_normal(Nb * sb)
    Xl = [Xb[i * sb:(i + 1) * sb] for i in range(Nb)]
    yl = [Yb[i * sb:(i + 1) * sb] for i in range(Nb)]
    Hm = np.array([np.linalg.eigvalsh(x.T @ x)[-1] for x in Xl])
    ob = O.opt_sol_closedForm(Xb, Yb)
    argsb = (Xl, yl, Nb, db, Hm, 120, ob, Xb, Yb, sb)
    ra = O.GD_DGD_LAG(*argsb, rng=5)
    rb2 = g.GD_DGD_LAG(*argsb, rng=5, backend="gloo")
    for k in (0, 3, 6, 10, 14, 16):
        np.testing.assert_allclose(ra[k], rb2[k], rtol=1e-9, atol=1e-12, err_msg=f"baseline {k}")
    np.testing.assert_array_equal(ra[9], rb2[9])
    np.testing.assert_array_equal(ra[13], rb2[13])
    da = O.dual_averaging(Xb, Yb, Nb, db, sb, 80, ob, 1e-6, ra[18])
    db2 = g.dual_averaging(Xb, Yb, Nb, db, sb, 80, ob, 1e-6, ra[18], backend="gloo")
    assert da[2] == db2[2]
    np.testing.assert_allclose(da[0], db2[0], rtol=1e-10)
    Yc = np.where(Yb > 0, 1.0, -1.0)
    ylc = [Yc[i * sb:(i + 1) * sb] for i in range(Nb)]
    Hl = 0.25 * Hm + 1e-2
    argsl = (Xl, ylc, Nb, db, Hl, 60, Xb, Yc, sb, 1e-2, 1e-9)
    la = O.GD_DGD_LAG_logistic(*argsl, rng=2)
    lb = g.GD_DGD_LAG_logistic(*argsl, rng=2, backend="gloo")
    for k in (0, 3, 6, 10, 14, 16):
        np.testing.assert_allclose(la[k], lb[k], rtol=1e-9, atol=1e-12, err_msg=f"logistic baseline {k}")
    # logistic GADMM (exact and inexact local solves), chain split over the two ranks
    star = O.opt_sol_logistic(Xb, Yc, db, 1e-3, Nb)
    ea = O.group_ADMM_logistic(Xb, Yc, 5.0, Nb, db, sb, 25, star, 1e-3, 1e-3)
    eb = g.group_ADMM_logistic(Xb, Yc, 5.0, Nb, db, sb, 25, star, 1e-3, 1e-3, backend="gloo")
    assert ea[2] == eb[2]
    np.testing.assert_allclose(ea[0], eb[0], rtol=1e-11)
    ia = O.group_ADMM_logistic_GD(Xb, Yc, 0.01, Nb, db, sb, 20, star, 1e-3, 0.0, 0.05, return_state=True)
    ib = g.group_ADMM_logistic_GD(Xb, Yc, 0.01, Nb, db, sb, 20, star, 1e-3, 0.0, 0.05, backend="gloo",
                                  return_state=True)
    assert ia[4]["inner_steps"] == ib[4]["inner_steps"]
    np.testing.assert_allclose(ia[0], ib[0], rtol=1e-11)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, "ok", int(b[2])))
