"""Sub-steps of the UKF restatement pinned to the REAL libraries filterpy itself calls (both importable here):

  * oracle/ukf.py::cholesky_upper  ==  scipy.linalg.cholesky  (what MerweScaledSigmaPoints uses as ``sqrt_method``)
  * the gain K = Pxz inv(S) with ``inv`` = np.linalg.inv  (what UnscentedKalmanFilter.update uses) against a solve,
    and the W-form update the augmented CUDA path uses (x = xbar + W^T v, P = Pbar - W^T W, W = Us^-T [Pzx | r])
    against filterpy's literal x + K y, P - K S K^T
  * Python's float ``%`` in normalize_angle == np.mod / math.fmod-with-floor (floor-mod)

and, when someone with filterpy has made tests/golden/filterpy_golden.npz (tests/golden/make_filterpy_golden.py), the
whole engine against upstream's own numbers."""
import math
import os

import numpy as np
import pytest
import scipy.linalg

from oracle import ukf
from tests import ukf_aug_util as au

GOLD = os.path.join(os.path.dirname(__file__), "golden", "filterpy_golden.npz")


def spd(rng, n, cond=1e4):
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    return (Q * np.geomspace(1.0, cond, n)) @ Q.T


@pytest.mark.parametrize("n", [1, 2, 3, 16, 27, 203])
def test_cholesky_upper_is_scipy_cholesky(n):
    rng = np.random.default_rng(n)
    for cond in (1e2, 1e8):
        A = spd(rng, n, cond)
        A = 0.5 * (A + A.T)
        U, Us = ukf.cholesky_upper(A), scipy.linalg.cholesky(A)          # upper: U^T U = A (scipy's default lower=False)
        np.testing.assert_allclose(U, Us, rtol=1e-10 * max(1.0, cond / 1e4), atol=1e-13 * np.sqrt(cond))
        assert np.array_equal(U, np.triu(U)) and np.array_equal(Us, np.triu(Us))
    # both read only the upper triangle, both raise LinAlgError on a matrix that is not positive definite
    B = A.copy()
    B[np.tril_indices(n, -1)] = 1e9
    np.testing.assert_array_equal(ukf.cholesky_upper(B), ukf.cholesky_upper(A))
    np.testing.assert_array_equal(scipy.linalg.cholesky(B), scipy.linalg.cholesky(A))
    bad = A - 2 * np.linalg.eigvalsh(A).max() * np.eye(n)
    with pytest.raises(np.linalg.LinAlgError):
        ukf.cholesky_upper(bad)
    with pytest.raises(np.linalg.LinAlgError):
        scipy.linalg.cholesky(bad)


def test_sigma_points_use_the_rows_of_the_scipy_factor():
    rng = np.random.default_rng(1)
    P = spd(rng, 5, 100.0)
    x = rng.normal(size=5)
    pts = ukf.MerweScaledSigmaPoints(5, 0.3, 2.0, 1.0)
    sg = pts.sigma_points(x, P)
    lam = 0.3 ** 2 * (5 + 1.0) - 5
    U = scipy.linalg.cholesky((5 + lam) * P)
    np.testing.assert_allclose(sg[1:6] - x, U, rtol=1e-11, atol=1e-13)       # x + U[k] (rows)
    np.testing.assert_allclose(x - sg[6:11], U, rtol=1e-11, atol=1e-13)


@pytest.mark.parametrize("dz,n", [(2, 3), (16, 3), (24, 27), (400, 403)])
def test_gain_forms_agree_with_numpy_inv(dz, n):
    """filterpy: SI = inv(S); K = Pxz SI; x += K y; P -= K S K^T.  The CUDA paths solve instead of inverting (n = 3: two
    triangular solves per row of K; augmented: W form).  All three are the same update to rounding."""
    rng = np.random.default_rng(dz)
    S = spd(rng, dz, 1e3)
    S = 0.5 * (S + S.T)
    Pxz = rng.normal(size=(n, dz))
    y = rng.normal(size=dz)
    Pbar = spd(rng, n, 1e3) * 50 + Pxz @ np.linalg.solve(S, Pxz.T)              # keeps the posterior positive definite
    xbar = rng.normal(size=n)
    K = Pxz @ np.linalg.inv(S)                                                   # upstream, literally
    x_lit, P_lit = xbar + K @ y, Pbar - K @ (S @ K.T)
    # solve form
    K2 = np.linalg.solve(S, Pxz.T).T
    np.testing.assert_allclose(K2, K, rtol=1e-9, atol=1e-12)
    # W form: S = Us^T Us, W = Us^-T [Pzx | y]
    Us = scipy.linalg.cholesky(S)
    W = scipy.linalg.solve_triangular(Us, np.column_stack([Pxz.T, y]), trans="T")
    x_w, P_w = xbar + W[:, :n].T @ W[:, n], Pbar - W[:, :n].T @ W[:, :n]
    np.testing.assert_allclose(x_w, x_lit, rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(P_w, P_lit, rtol=1e-8, atol=1e-9 * np.abs(P_lit).max())


def test_normalize_angle_is_floor_mod():
    for a in (0.0, math.pi, -math.pi, 2 * math.pi, 3 * math.pi + 1e-9, 3 * math.pi - 1e-9, -7.5, 1e6, -1e6):
        m = a % (2 * math.pi)
        assert m == float(np.mod(a, 2 * math.pi)) and 0.0 <= m < 2 * math.pi
        w = ukf.normalize_angle(a)
        assert -math.pi < w <= math.pi and abs(math.remainder(w - a, 2 * math.pi)) < 1e-9 * max(1.0, abs(a))


# ---- upstream's own numbers, when someone with filterpy has made the golden file --------------------------------
needs_gold = pytest.mark.skipif(not os.path.exists(GOLD), reason="tests/golden/filterpy_golden.npz absent: run "
                                "tests/golden/make_filterpy_golden.py where filterpy is installed")


@needs_gold
def test_restatement_matches_filterpy_golden():
    g = np.load(GOLD)
    sysm = ukf.System([ukf._Pos(*p) for p in g["lm"]])
    for k in range(5):
        sysm.ukf.predict(u=g["u"])
        np.testing.assert_allclose(sysm.ukf.x, g["x_prior"][k], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(sysm.ukf.P, g["P_prior"][k], rtol=1e-9, atol=1e-14)
        sysm.ukf.update(g["z"][k], landmarks=sysm.landmarks)
        np.testing.assert_allclose(sysm.ukf.x, g["x"][k], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(sysm.ukf.P, g["P"][k], rtol=1e-9, atol=1e-14)
    want = au.oracle_aug(g["aug_x0"], g["aug_P0"], g["aug_u"], g["aug_z"])
    au.assert_aug_close((g["aug_x"], g["aug_P"]), want, "augmented restatement vs filterpy", rtol=1e-8)


@needs_gold
@pytest.mark.gpu
def test_gpu_matches_filterpy_golden(gpu_engine):
    import torch
    from tests import ukf_util as uu
    g = np.load(GOLD)
    eng = gpu_engine
    dev = eng.device
    st = eng.alloc_ukf(1)
    st["x"] = torch.zeros((1, 3), dtype=torch.float64, device=dev)
    st["P"] = torch.diag(torch.tensor([.1, .1, .05], dtype=torch.float64, device=dev))[None].contiguous()
    lm = torch.from_numpy(g["lm"][None].copy()).to(dev)
    for k in range(5):
        z = torch.from_numpy(g["z"][k].reshape(1, 8, 2).copy()).to(dev)
        eng.ukf_step(st, z, lm, u=torch.from_numpy(g["u"][None].copy()).to(dev))
        torch.cuda.synchronize()
        status = st["status"].cpu().numpy()
        uu.assert_ukf_close((g["x"][k][None], g["P"][k][None], status), (st["x"].cpu().numpy(), st["P"].cpu().numpy(), status),
                            "filterpy golden step %d" % k)
    F, n = g["aug_x0"].shape
    L = (n - 3) // 2
    sa = eng.alloc_ukf_aug(F, L)
    sa["x"] = torch.from_numpy(g["aug_x0"].copy()).to(dev)
    sa["P"] = torch.from_numpy(g["aug_P0"].copy()).to(dev)
    eng.ukf_aug_step(sa, torch.from_numpy(g["aug_z"].copy()).to(dev), u=torch.from_numpy(g["aug_u"].copy()).to(dev))
    torch.cuda.synchronize()
    want = (g["aug_x"], g["aug_P"])
    cond = au.oracle_conditioning(g["aug_x0"], g["aug_P0"], g["aug_u"], g["aug_z"], want)
    au.assert_aug_close(want, (sa["x"].cpu().numpy(), sa["P"].cpu().numpy()), "augmented GPU vs filterpy", cond=cond)
