"""Multi-rank worker launched by tests/test_multigpu.py through torchrun (one rank per GPU)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import gadmm_b200 as g
    from gadmm_b200 import oracle as O
    from gadmm_b200.parallel.runtime import Dist, init_process_group

    init_process_group("nccl")
    D = Dist()
    rng = np.random.default_rng(5)
    N, s, d = 13, 30, 70
    XX = rng.standard_normal((N * s, d))
    YY = rng.standard_normal(N * s)
    obj0 = O.opt_sol_closedForm(XX, YY)
    base = (XX, YY, 1.5, N, d, s, 400, obj0, 1e-6)
    a = O.group_ADMM_closedForm(*base)
    b = g.group_ADMM_closedForm(*base, backend="b200", dist=D)
    assert a[2] == b[2], (a[2], b[2])
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)
    # tiny problem without an explicit dist: every rank solves it with the persistent kernel
    rs_ = np.random.default_rng(3)
    Xs_, Ys_ = rs_.standard_normal((6 * 15, 20)), rs_.standard_normal(6 * 15)
    os_ = O.opt_sol_closedForm(Xs_, Ys_)
    a = O.group_ADMM_closedForm(Xs_, Ys_, 2.0, 6, 20, 15, 500, os_, 1e-6)
    b = g.group_ADMM_closedForm(Xs_, Ys_, 2.0, 6, 20, 15, 500, os_, 1e-6, backend="b200")
    assert a[2] == b[2], ("solo", a[2], b[2])
    np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)
    # dynamic chains: neighbours land on arbitrary GPUs, re-chained every 3 iterations
    prng = np.random.default_rng(0)
    paths, costs = [], []
    for _ in range(420):
        pa, co, _ = g.findPath(N, rng=prng)
        paths.append(pa)
        costs.append(co)
    paths, costs = np.array(paths), np.array(costs)
    for tau in (1, 3):
        a = O.dynamic_group_ADMM_closedForm_v0(*base, paths, costs, tau)
        b = g.dynamic_group_ADMM_closedForm_v0(*base, paths, costs, tau, backend="b200", dist=D)
        assert a[2] == b[2], (tau, a[2], b[2])
        np.testing.assert_allclose(a[0], b[0], rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(a[4], b[4], rtol=1e-12)
    # long run on the static chain: flag protocol stress (no stop), windows replayed as graphs
    c = g.group_ADMM_closedForm(XX, YY, 1.5, N, d, s, 3000, obj0, 0.0, backend="b200", dist=D)
    r = O.group_ADMM_closedForm(XX, YY, 1.5, N, d, s, 3000, obj0, 0.0)
    np.testing.assert_allclose(c[0], r[0], rtol=1e-9, atol=1e-9)
    # star ADMM: in-kernel reduction by P2P stores, and by NVLS multimem.ld_reduce / multimem.st
    from gadmm_b200.engine.star import star_admm
    rs = O.standared_ADMM(XX, YY, 1.0, N, d, s, 500, obj0, 1e-6)
    used = []
    for mode in (False, "auto"):
        out = star_admm(XX, YY, 1.0, N, d, s, 500, obj0, 1e-6, dist=D, nvls=mode, return_engine=True)
        assert out[2] == rs[2], (mode, out[2], rs[2])
        np.testing.assert_allclose(out[0], rs[0], rtol=1e-9, atol=1e-9)
        used.append(out[4].nvls)
    # parameter-server baselines with the workers split over the GPUs (fused P2P reduce + step)
    rng2 = np.random.default_rng(8)
    Nb, sb, db = 11, 20, 40
    Xb = rng2.standard_normal((Nb * sb, db))
    Yb = Xb @ rng2.standard_normal(db) + 0.1 * rng2.standard_normal(Nb * sb)
    ob = O.opt_sol_closedForm(Xb, Yb)
    Xl = [Xb[i * sb:(i + 1) * sb] for i in range(Nb)]
    yl = [Yb[i * sb:(i + 1) * sb] for i in range(Nb)]
    Hm = np.array([np.linalg.eigvalsh(x.T @ x)[-1] for x in Xl])
    argsb = (Xl, yl, Nb, db, Hm, 400, ob, Xb, Yb, sb)
    ra = O.GD_DGD_LAG(*argsb, rng=3)
    rb = g.GD_DGD_LAG(*argsb, rng=3, backend="b200", dist=D)
    for kk in (0, 3, 6, 10, 14, 16):
        np.testing.assert_allclose(ra[kk], rb[kk], rtol=1e-7, atol=1e-9, err_msg=f"baseline curve {kk}")
    live = int(np.nonzero(ra[7] < 1e-7)[0][0]) if np.any(ra[7] < 1e-7) else 400
    np.testing.assert_array_equal(ra[9][:live], rb[9][:live])
    # dual averaging along the chain with Z halos over NVLink (Gauss-Seidel order and Jacobi)
    for mode in ("faithful", "jacobi"):
        da = O.dual_averaging(Xb, Yb, Nb, db, sb, 300, ob, 1e-4, ra[18], mode=mode, return_state=True)
        dbb = g.dual_averaging(Xb, Yb, Nb, db, sb, 300, ob, 1e-4, ra[18], mode=mode, backend="b200",
                               dist=D, return_state=True)
        assert da[2] == dbb[2], (mode, da[2], dbb[2])
        np.testing.assert_allclose(da[0], dbb[0], rtol=1e-9, err_msg=f"dual averaging {mode}")
        np.testing.assert_allclose(da[4]["theta"], dbb[4]["theta"], rtol=1e-8, atol=1e-12)
    # logistic GADMM, exact local solves by Newton-CG, chain split over the GPUs
    rl = np.random.default_rng(4)
    Nl, sl, dl, laml, rhol = 9, 24, 18, 1e-3, 5.0
    Xl_ = rl.standard_normal((Nl * sl, dl))          # double data, fp64 shards
    Yl_ = np.where(rl.standard_normal(Nl * sl) > 0, 1.0, -1.0)
    ol = O.opt_sol_logistic(Xl_, Yl_, dl, laml, Nl)
    argl = (Xl_, Yl_, rhol, Nl, dl, sl, 300, ol, laml, 1e-3)
    la = O.group_ADMM_logistic(*argl, naive=False, return_state=True)
    lb = g.group_ADMM_logistic(*argl, backend="b200", dist=D, return_state=True)
    assert la[2] == lb[2] and la[2] < 300, ("logistic", la[2], lb[2])
    np.testing.assert_allclose(la[0], lb[0], rtol=1e-8)
    g0l = lb[4]["g0"]
    np.testing.assert_allclose(la[4]["theta"][:, g0l:g0l + lb[4]["theta"].shape[1]], lb[4]["theta"],
                               rtol=1e-5, atol=1e-7)
    # the reference's inexact local solve (<= 100 frozen-penalty GD steps), chain split over the GPUs
    argg = (Xl_, Yl_, 0.01, Nl, dl, sl, 60, ol, laml, 0.0, 0.05)
    ga = O.group_ADMM_logistic_GD(*argg, return_state=True)
    gb = g.group_ADMM_logistic_GD(*argg, backend="b200", dist=D, return_state=True)
    assert ga[2] == gb[2], ("logistic GD", ga[2], gb[2])
    np.testing.assert_allclose(ga[0], gb[0], rtol=1e-9)
    assert ga[4]["inner_steps"] == gb[4]["inner_steps"], (ga[4]["inner_steps"], gb[4]["inner_steps"])
    # synthetic shards are drawn per global worker id: same problem for any number of ranks
    from gadmm_b200.engine.logistic_scaled import gadmm_logistic_scaled
    from gadmm_b200.parallel.runtime import SoloDist
    syn_a = gadmm_logistic_scaled(8, 256, 64, 0.5, 1e-3, 12, 0.0, dist=D, return_state=True)
    syn_b = gadmm_logistic_scaled(8, 256, 64, 0.5, 1e-3, 12, 0.0, dist=SoloDist(), return_state=True)
    assert abs(syn_a[4]["obj0"] - syn_b[4]["obj0"]) <= 1e-9 * abs(syn_b[4]["obj0"])
    np.testing.assert_allclose(syn_a[0], syn_b[0], rtol=1e-8)
    # plain torch.distributed (NCCL) baselines of the first-order and logistic families
    nb = g.GD_DGD_LAG(*argsb, rng=3, backend="nccl")
    for kk in (0, 3, 6, 10, 14, 16):
        np.testing.assert_allclose(ra[kk], nb[kk], rtol=1e-8, atol=1e-10, err_msg=f"nccl baseline {kk}")
    np.testing.assert_array_equal(ra[9][:live], nb[9][:live])
    dn = g.dual_averaging(Xb, Yb, Nb, db, sb, 300, ob, 1e-4, ra[18], backend="nccl")
    do = O.dual_averaging(Xb, Yb, Nb, db, sb, 300, ob, 1e-4, ra[18])
    assert dn[2] == do[2]
    np.testing.assert_allclose(do[0], dn[0], rtol=1e-9)
    ln = g.group_ADMM_logistic_GD(*argg, backend="nccl", return_state=True)
    assert ln[4]["inner_steps"] == ga[4]["inner_steps"]
    np.testing.assert_allclose(ga[0], ln[0], rtol=1e-10)
    # optional protocol stress: GADMM_STRESS_ITERS iterations, a new random
    # chain every iteration (drawn and switched on the device), split run == single-GPU run
    stress = int(os.environ.get("GADMM_STRESS_ITERS", "0"))
    if stress:
        Ns = int(os.environ.get("GADMM_STRESS_WORKERS", str(N)))      # e.g. 64: >= 2 workers per GPU at 8 GPUs
        if Ns != N:
            rs2 = np.random.default_rng(17)
            XXs, YYs = rs2.standard_normal((Ns * s, d)), rs2.standard_normal(Ns * s)
            objs = O.opt_sol_closedForm(XXs, YYs)
        else:
            XXs, YYs, objs = XX, YY, obj0
        p0 = np.arange(1, Ns + 1)
        sargs = (XXs, YYs, 1.5, Ns, d, s, stress, objs, 0.0, p0, np.ones(Ns - 1), 1)
        sm = g.dynamic_group_ADMM_closedForm(*sargs, backend="b200", dist=D, chain_gen="device", rng=5)
        ss = g.dynamic_group_ADMM_closedForm(*sargs, backend="b200", dist=SoloDist(),
                                             chain_gen="device", rng=5)
        assert sm[2] == ss[2] == stress
        np.testing.assert_allclose(sm[0], ss[0], rtol=1e-9, atol=1e-9)
        if D.rank == 0:
            print(f"STRESS_OK workers={Ns} iters={stress} final_gap={sm[1][-1]:.3e} "
                  f"us_per_iter={1e6 * sm[3][-1] / stress:.1f}")
    D.barrier()
    if D.rank == 0:
        print(f"STAR_NVLS_USED={used}")
        print(f"MULTIGPU_OK world={D.world}")
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
