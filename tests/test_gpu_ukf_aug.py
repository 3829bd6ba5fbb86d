"""GPU parity of the augmented-state UKF (csrc/ukf_aug.cuh through lslam_ukf_aug_step) against the
literal filterpy restatement (oracle/ukf.py step_literal_augmented) up to BASELINE.json configs[3]'s
n = 403 (L = 200).  Bar: 1e-5 relative; where the filter itself is ill-conditioned (alpha = 1e-4, large
n) the bar is 8 x the oracle's own sensitivity to a one-ulp change of its input (tests/ukf_aug_util.py)."""
import numpy as np
import pytest

from tests import ukf_aug_util as au

pytestmark = pytest.mark.gpu


def run_gpu(eng, x, P, u, z, flags=None, params=None, workspace_filters=None):
    import torch
    dev = eng.device
    F, n = x.shape
    L = (n - 3) // 2
    st = eng.alloc_ukf_aug(F, L, workspace_filters=workspace_filters)
    st["x"] = torch.from_numpy(x).to(dev)
    st["P"] = torch.from_numpy(P).to(dev)
    if params is not None:
        st["params"] = list(params)
    eng.ukf_aug_step(st, torch.from_numpy(z).to(dev), u=torch.from_numpy(u).to(dev), flags=flags)
    torch.cuda.synchronize()
    return st["x"].cpu().numpy(), st["P"].cpu().numpy(), st["status"].cpu().numpy()


@pytest.mark.parametrize("L,F", [(1, 40), (5, 64), (16, 16), (32, 5), (40, 6), (64, 3), (200, 2), (256, 1)])   # dz = 64, 128, 512: tile-aligned edges
def test_aug_step_matches_literal(gpu_engine, L, F):
    x, P, u, z = au.aug_case(F, L, seed=L)
    want = au.oracle_aug(x, P, u, z)
    cond = au.oracle_conditioning(x, P, u, z, want)
    xg, Pg, st = run_gpu(gpu_engine, x, P, u, z)
    assert not st.any()
    ex, eP = au.assert_aug_close(want, (xg, Pg), "L=%d" % L, cond=cond)
    print("L=%d err x %.2e P %.2e (oracle one-ulp sensitivity %.2e %.2e)" % (L, ex, eP, cond[0], cond[1]))


@pytest.mark.parametrize("L,F", [(40, 4), (200, 1)])
def test_aug_literal_and_paired_cross_covariance(gpu_engine, L, F):
    """ukf_aug.pzx = 0 (the literal sum over all 2n + 1 sigma points, zero k-tiles skipped) and 1 (default: + / - points
    paired, half the depth) both hold the parity bar, and differ from each other by far less than the bar."""
    x, P, u, z = au.aug_case(F, L, seed=100 + L)
    want = au.oracle_aug(x, P, u, z)
    cond = au.oracle_conditioning(x, P, u, z, want)
    got = {}
    try:
        for mode in (0, 1):
            gpu_engine.set_option("ukf_aug.pzx", mode)
            xg, Pg, st = run_gpu(gpu_engine, x, P, u, z)
            assert not st.any()
            ex, eP = au.assert_aug_close(want, (xg, Pg), "pzx mode %d L=%d" % (mode, L), cond=cond)
            print("L=%d pzx=%d err x %.2e P %.2e (oracle one-ulp sensitivity %.2e %.2e)" % (L, mode, ex, eP, cond[0], cond[1]))
            got[mode] = (xg, Pg)
    finally:
        gpu_engine.set_option("ukf_aug.pzx", 1)
    dx, dP = au._errors(got[0], got[1])
    assert dx < 1e-6 and dP < 1e-6, (dx, dP)



@pytest.mark.parametrize("flags", [1, 2, 3])
def test_aug_halves(gpu_engine, flags):
    x, P, u, z = au.aug_case(8, 20, seed=flags)
    want = au.oracle_aug(x, P, u, z, flags)
    cond = au.oracle_conditioning(x, P, u, z, want, flags)
    xg, Pg, st = run_gpu(gpu_engine, x, P, u, z, flags=flags)
    assert not st.any()
    au.assert_aug_close(want, (xg, Pg), "flags %d" % flags, cond=cond)


def test_aug_strict_tolerance_with_wider_sigma_spread(gpu_engine):
    """alpha = 0.05: well conditioned, plain 1e-5 (in fact ~1e-9) without the conditioning term, n = 403."""
    from lidar_slam_b200 import batch
    x, P, u, z = au.aug_case(1, 200, seed=9)
    want = au.oracle_aug(x, P, u, z, 7, alpha=0.05)
    xg, Pg, st = run_gpu(gpu_engine, x, P, u, z, params=batch.Engine.ukf_params(alpha=0.05))
    assert not st.any()
    ex, eP = au.assert_aug_close(want, (xg, Pg), "alpha 0.05 L=200")
    assert ex < 1e-7 and eP < 1e-7


@pytest.mark.parametrize("L,flags", [(40, 7), (40, 3), (40, 2), (200, 7)])
def test_aug_scratch_guards_and_nan_workspace(gpu_engine, L, flags):
    """Guard bands around x, P and the workspace stay untouched, and a workspace full of NaNs gives the same bits as a
    zeroed one (no kernel reads scratch the step did not write first; TMA over-fetch columns are never used)."""
    import torch
    eng = gpu_engine
    dev = eng.device
    F = 3
    n = 3 + 2 * L
    x, P, u, z = au.aug_case(F, L, seed=300 + L)
    nbytes = eng.handle.ukf_aug_workspace_bytes(L, F)
    G = 4096
    outs = []
    for fill in (0, 0xFF):
        big = torch.full((nbytes + 2 * G,), fill, dtype=torch.uint8, device=dev)
        big[:G] = 0xA5; big[G + nbytes:] = 0xA5
        xb = torch.full((F * n + 64,), 7.25, dtype=torch.float64, device=dev)
        Pb = torch.full((F * n * n + 64,), 7.25, dtype=torch.float64, device=dev)
        st = eng.alloc_ukf_aug(F, L, workspace_filters=1)
        st["workspace"] = big[G:G + nbytes]
        st["x"] = xb[32:32 + F * n].view(F, n); st["x"].copy_(torch.from_numpy(x))
        st["P"] = Pb[32:32 + F * n * n].view(F, n, n); st["P"].copy_(torch.from_numpy(P))
        eng.ukf_aug_step(st, torch.from_numpy(z).to(dev), u=torch.from_numpy(u).to(dev), flags=flags)
        torch.cuda.synchronize()
        assert not st["status"].any().item()
        assert (big[:G] == 0xA5).all().item() and (big[G + nbytes:] == 0xA5).all().item(), "workspace guard"
        for buf, m in ((xb, F * n), (Pb, F * n * n)):
            assert (buf[:32] == 7.25).all().item() and (buf[32 + m:] == 7.25).all().item(), "state guard"
        outs.append((st["x"].cpu().numpy().copy(), st["P"].cpu().numpy().copy()))
    assert np.isfinite(outs[1][0]).all() and np.isfinite(outs[1][1]).all()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])


def test_aug_batching_failures_and_repeat_steps(gpu_engine):
    import torch
    L, F = 12, 37
    x, P, u, z = au.aug_case(F, L, seed=6)
    P[5, 7, 7] = -2.0
    a = run_gpu(gpu_engine, x, P, u, z)
    b = run_gpu(gpu_engine, x, P, u, z, workspace_filters=4)        # 10 batches, the last one short
    for p, q in zip(a, b):
        assert np.array_equal(p, q)
    assert a[2][5] == 1 and a[2].sum() == 1
    assert np.array_equal(a[0][5], x[5]) and np.array_equal(a[1][5], P[5])
    # posterior stays symmetric positive definite over repeated steps
    eng = gpu_engine
    st = eng.alloc_ukf_aug(4, 30)
    xs, Ps, us, zs = au.aug_case(4, 30, seed=8)
    st["x"] = torch.from_numpy(xs).to(eng.device)
    st["P"] = torch.from_numpy(Ps).to(eng.device)
    zt = torch.from_numpy(zs).to(eng.device)
    for _ in range(5):
        eng.ukf_aug_step(st, zt, u=torch.from_numpy(us).to(eng.device))
    torch.cuda.synchronize()
    assert not st["status"].any().item()
    Pn = st["P"].cpu().numpy()
    assert np.abs(Pn - Pn.transpose(0, 2, 1)).max() <= 1e-6 * np.abs(Pn).max()
    assert np.all(np.linalg.eigvalsh((Pn + Pn.transpose(0, 2, 1)) / 2) > 0)


def test_aug_against_extended_precision_truth(gpu_engine):
    """BASELINE.json configs[3] at the reference's alpha = 1e-4, L = 200 (n = 403), 16 distinct filters, against a TRUTH
    value instead of a self-chosen bar: tests/golden/ukf_aug_kat.npz holds the same step evaluated in x87 extended precision
    (eps 1.1e-19; tests/golden/make_ukf_aug_kat.py).  The kernel must be as close to the truth as the fp64 oracle is:
    |GPU - truth| <= max(1e-5, 1.5 |oracle - truth|) for x and P, in both cross-covariance modes, taken over the 16 filters
    as the worst case and as the root mean square (per filter both errors are one realisation of the same rounding noise,
    so single-filter ratios scatter around 1: printed).  If the kernel were further from the truth than the oracle, that
    would be a kernel accuracy bug, not conditioning."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ukf_aug_kat.npz"))
    F, L = int(g["F"]), int(g["L"])
    x, P, u, z = au.aug_case(F, L, seed=int(g["seed"]))
    xo, Po = au.oracle_aug(x, P, u, z)
    ox, oP = au.truth_errors(g, xo, Po)
    print("oracle vs truth:   x %.2e  P %.2e (max over %d filters)" % (ox.max(), oP.max(), F))
    try:
        for mode in (0, 1):
            gpu_engine.set_option("ukf_aug.pzx", mode)
            xg, Pg, st = run_gpu(gpu_engine, x, P, u, z)
            assert not st.any()
            gx, gP = au.truth_errors(g, xg, Pg)
            print("pzx=%d GPU vs truth: x %.2e  P %.2e; worst ratio to the oracle's error: x %.2f  P %.2f"
                  % (mode, gx.max(), gP.max(), (gx / np.maximum(ox, 1e-300)).max(), (gP / np.maximum(oP, 1e-300)).max()))
            rms = lambda v: float(np.sqrt(np.mean(v * v)))
            print("       rms over filters: GPU x %.2e P %.2e, oracle x %.2e P %.2e" % (rms(gx), rms(gP), rms(ox), rms(oP)))
            for name, ge, oe in (("x", gx, ox), ("P", gP, oP)):
                assert ge.max() <= max(1e-5, 1.5 * oe.max()), "pzx=%d: worst %s error %.3g against the oracle's %.3g" % (mode, name, ge.max(), oe.max())
                assert rms(ge) <= max(1e-5, 1.5 * rms(oe)), "pzx=%d: rms %s error %.3g against the oracle's %.3g" % (mode, name, rms(ge), rms(oe))
    finally:
        gpu_engine.set_option("ukf_aug.pzx", 1)


def test_failed_redraw_or_S_leaves_the_prior_in_both_filters(gpu_engine):
    """As tests/test_emu_ukf_aug.py: a failed re-draw (status 2) or S factorisation (status 4) after a predict leaves
    the PRIOR - bit for bit what a predict-only call returns - in the augmented path and in the n = 3 kernel."""
    import torch
    from oracle import cport
    eng = gpu_engine
    bad_q = cport.ukf_params(Q=-1e3 * np.eye(3))
    bad_r = cport.ukf_params(var_dist=-1e12)
    x, P, u, z = au.aug_case(5, 40, seed=9)
    for params, bit in ((bad_q, 2), (bad_r, 4)):
        prior = run_gpu(eng, x, P, u, z, flags=1, params=params)
        got = run_gpu(eng, x, P, u, z, flags=7, params=params, workspace_filters=2)
        assert got[2].tolist() == [bit] * 5 and not prior[2].any()
        assert np.array_equal(got[0], prior[0]) and np.array_equal(got[1], prior[1])
        assert not np.array_equal(got[0], x)
    rng = np.random.default_rng(4)
    F, L = 70, 8
    lm = rng.normal(0, 1500, (F, L, 2))
    z3 = torch.from_numpy(np.stack([np.hypot(lm[..., 0], lm[..., 1]), np.arctan2(lm[..., 1], lm[..., 0])], -1)).to(eng.device)
    lmt = torch.from_numpy(lm).to(eng.device)
    for params, bit in ((bad_q, 2), (bad_r, 4)):
        outs = []
        for flags in (1, 7):
            st = eng.alloc_ukf(F, L)
            st["params"] = list(params)
            st["u"].fill_(1.0)
            eng.ukf_step(st, z3, lmt, flags=flags)
            torch.cuda.synchronize()
            outs.append((st["x"].cpu().numpy(), st["P"].cpu().numpy(), st["status"].cpu().numpy()))
        assert outs[1][2].tolist() == [bit] * F and not outs[0][2].any()
        assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
