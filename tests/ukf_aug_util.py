"""Shared set-up for the augmented-state UKF tests."""
import numpy as np

from oracle import ukf as ou


def aug_case(F, L, seed=0, spread=3000.0):
    """F filters with L landmarks in the state: the true pose near the origin, landmarks 1-4 m away,
    the filter's state = truth + prior-sized noise, z = the true (range, bearing) + sensor noise."""
    rng = np.random.default_rng(seed)
    n = 3 + 2 * L
    pose = np.column_stack([rng.normal(0, 50, F), rng.normal(0, 50, F), rng.uniform(-np.pi, np.pi, F)])
    ang = rng.uniform(-np.pi, np.pi, (F, L))
    rad = rng.uniform(1000.0, 1000.0 + spread, (F, L))
    lm = np.stack([pose[:, :1] + rad * np.cos(ang), pose[:, 1:2] + rad * np.sin(ang)], axis=2)     # [F, L, 2]
    x = np.zeros((F, n))
    x[:, :3] = pose + rng.normal(0, 1, (F, 3)) * np.array([0.3, 0.3, 0.02])
    x[:, 3:] = (lm + rng.normal(0, 5.0, lm.shape)).reshape(F, -1)
    d = np.array([0.1, 0.1, 0.05] + [25.0] * (2 * L))
    P = np.broadcast_to(np.diag(d), (F, n, n)).copy()
    # a little correlation so P is not diagonal
    A = rng.normal(0, 0.05, (F, n, n))
    P = P + np.einsum("fij,fkj->fik", A, A)
    u = rng.uniform(0.5, 2.0, (F, 2))
    dx, dy = lm[..., 0] - pose[:, :1], lm[..., 1] - pose[:, 1:2]
    rg = np.hypot(dx, dy) + rng.normal(0, 0.5, (F, L))
    br = np.arctan2(dy, dx) - pose[:, 2:3] + rng.normal(0, 0.05, (F, L))
    br = (br + np.pi) % (2 * np.pi) - np.pi
    z = np.stack([rg, br], axis=2)
    return x, P, u, z


def oracle_aug(x, P, u, z, flags=7, **kw):
    F = x.shape[0]
    xs, Ps = np.empty_like(x), np.empty_like(P)
    for f in range(F):
        xs[f], Ps[f] = ou.step_literal_augmented(x[f], P[f], u[f], z[f], redraw=bool(flags & 4),
                                                 do_predict=bool(flags & 1), do_update=bool(flags & 2), **kw)
    return xs, Ps


def _errors(want, got):
    """x: |dx| / max(|x|, 10) (10 mm / rad floor); P: |dP_ij| / sqrt(P_ii P_jj) of the posterior."""
    xw, Pw = want
    xg, Pg = got
    ex = (np.abs(xg - xw) / np.maximum(np.abs(xw), 10.0)).max()
    dg = np.sqrt(np.abs(np.einsum("fii->fi", Pw)))
    scale = dg[:, :, None] * dg[:, None, :]
    eP = (np.abs(Pg - Pw) / np.maximum(scale, 1e-300)).max()
    return ex, eP


def oracle_conditioning(x, P, u, z, want, flags=7, **kw):
    """How far the ORACLE's own result moves when every input state entry moves by one ulp (up, down):
    the floor below which no other evaluation order can be told from the oracle.  With the reference's
    alpha = 1e-4 (systemClass.py:20) the central weight is -n / (alpha^2 n) = -1e8, so the weighted mean
    of coordinates of a few thousand mm carries ~1e8 * 3e3 * 2^-53 = 3e-5 mm of rounding, which the
    covariance of sigma-point spreads of ~1e-3 mm then amplifies; the movement grows with n."""
    ex = eP = 0.0
    for tgt in (np.inf, -np.inf):
        got = oracle_aug(np.nextafter(x, tgt), P, u, z, flags, **kw)
        a, b = _errors(want, got)
        ex, eP = max(ex, a), max(eP, b)
    return ex, eP


def assert_aug_close(want, got, what="", rtol=1e-5, cond=None, cond_factor=8.0):
    """north_star's 1e-5 relative on x and P, widened - only where the filter itself is ill-conditioned -
    to cond_factor times the oracle's own one-ulp sensitivity (``cond`` = oracle_conditioning(...))."""
    ex, eP = _errors(want, got)
    tx = max(rtol, cond_factor * cond[0]) if cond is not None else rtol
    tP = max(rtol, cond_factor * cond[1]) if cond is not None else rtol
    assert ex <= tx, "%s x: max rel err %.3g > %.3g" % (what, ex, tx)
    assert eP <= tP, "%s P: max scaled err %.3g > %.3g" % (what, eP, tP)
    return ex, eP


def truth_errors(g, x, P):
    """Per-filter errors of (x [F, n], P [F, n, n]) against the extended-precision truth of tests/golden/ukf_aug_kat.npz
    (tests/golden/make_ukf_aug_kat.py): x in full, P on its diagonal and the stored sample of entries; same scaling as
    _errors (x: |dx| / max(|x|, 10); P: |dP_ij| / sqrt(P_ii P_jj))."""
    xt, pd, ps, si, sj = g["x"], g["P_diag"], g["P_sample"], g["sample_i"], g["sample_j"]
    F = xt.shape[0]
    ex = (np.abs(x - xt) / np.maximum(np.abs(xt), 10.0)).max(axis=1)
    dg = np.sqrt(np.abs(pd))
    e_diag = np.abs(np.einsum("fii->fi", P) - pd) / (dg * dg)
    e_samp = np.abs(P[:, si, sj] - ps) / (dg[:, si] * dg[:, sj])
    return ex, np.maximum(e_diag.max(axis=1), e_samp.max(axis=1)).reshape(F)
