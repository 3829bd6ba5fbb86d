"""Generates tests/golden/ukf_aug_kat.npz: a TRUTH value for BASELINE.json configs[3] - the augmented UKF step at the
reference's alpha = 1e-4 (/root/reference/systemClass.py:20) with L = 200 landmarks in the state (n = 403).

At alpha = 1e-4 the central sigma weight is -1e8, so ANY fp64 evaluation of the filter carries rounding noise far above
fp64's 1e-16 (tests/ukf_aug_util.py::oracle_conditioning).  To tell "the kernel is as close to the true answer as the
fp64 oracle is" from "the kernel is wrong", this script evaluates the same equations (oracle/ukf.py: filterpy's
MerweScaledSigmaPoints + UnscentedKalmanFilter.predict/update with the repaired UKFMethods callbacks, re-draw on) in
x87 extended precision (np.longdouble: 64-bit significand, eps = 1.1e-19 - 2,048 times finer than fp64; 60-digit
mpmath as used for the n = 3 KAT would take hours at n = 403) for F filters of tests/ukf_aug_util.aug_case(F, 200,
seed).  Inputs are regenerated from the seed by the test, so only the results are stored: x [F, n] in full, and of
P [F, n, n] the diagonal plus a fixed random sample of off-diagonal entries (a full P is 1.3 MB per filter).

    python tests/golden/make_ukf_aug_kat.py            # ~ 5 minutes
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ukf as ou  # noqa: E402  (constants only)
from tests import ukf_aug_util as au  # noqa: E402

LD = np.longdouble
F, L, SEED, NSAMPLE = 16, 200, 200, 4096


def chol_upper(A):
    """U^T U = A in extended precision (row-oriented, vectorised over the row)."""
    n = A.shape[0]
    U = np.zeros((n, n), LD)
    for j in range(n):
        s = A[j, j] - U[:j, j] @ U[:j, j]
        assert s > 0
        U[j, j] = np.sqrt(s)
        U[j, j + 1:] = (A[j, j + 1:] - U[:j, j] @ U[:j, j + 1:]) / U[j, j]
    return U


def norm_angle(a):
    two_pi = LD(2) * np.arctan2(LD(0), LD(-1))                     # pi in extended precision
    a = a - two_pi * np.floor(a / two_pi)
    return np.where(a > two_pi / 2, a - two_pi, a)


def sigma_points(x, P, c):
    U = chol_upper(c * P)
    return np.concatenate([x[None], x[None] + U, x[None] - U], 0)


def fx(S, dt, u, wheel_r=LD(50), base_l=LD(200)):
    S = S.copy()
    th = S[:, 2].copy()
    S[:, 0] = S[:, 0] + dt * (wheel_r / 2 * np.cos(th) * (u[0] + u[1]))
    S[:, 1] = S[:, 1] + dt * (wheel_r / 2 * np.sin(th) * (u[0] + u[1]))
    S[:, 2] = th + dt * (-wheel_r / base_l * u[0] + wheel_r / base_l * u[1])
    return S


def hx(S):
    lm = S[:, 3:].reshape(S.shape[0], -1, 2)
    dx, dy = lm[..., 0] - S[:, 0:1], lm[..., 1] - S[:, 1:2]
    z = np.empty((S.shape[0], lm.shape[1] * 2), LD)
    z[:, 0::2] = np.sqrt(dx * dx + dy * dy)
    z[:, 1::2] = norm_angle(np.arctan2(dy, dx) - S[:, 2:3])
    return z


def mean_x(S, Wm):
    m = Wm @ S
    m[2] = np.arctan2(Wm @ np.sin(S[:, 2]), Wm @ np.cos(S[:, 2]))
    return m


def mean_z(Z, Wm):
    m = Wm @ Z
    m[1::2] = np.arctan2(Wm @ np.sin(Z[:, 1::2]), Wm @ np.cos(Z[:, 1::2]))
    return m


def resid_x(S, m):
    y = S - m
    y[:, 2] = norm_angle(y[:, 2])
    return y


def resid_z(Z, m):
    e = Z - m
    e[:, 1::2] = norm_angle(e[:, 1::2])
    return e


def step(x, P, u, z, alpha=LD("1e-4"), beta=LD(2), kappa=LD(0)):
    x, P, u, z = (np.asarray(v, LD) for v in (x, P, u, z.reshape(-1)))
    n = x.shape[0]
    lam = alpha * alpha * (n + kappa) - n
    c = n + lam
    Wm = np.full(2 * n + 1, LD(0.5) / c)
    Wc = Wm.copy()
    Wm[0] = lam / c
    Wc[0] = lam / c + (1 - alpha * alpha + beta)
    dt = LD(ou.DT)
    Q = np.zeros((n, n), LD)
    Q[:3, :3] = np.eye(3, dtype=LD) * LD("0.001")
    R = np.diag(np.tile(np.array([LD(ou.VAR_DIST), LD(ou.VAR_ANGLE)]), (n - 3) // 2))
    # predict
    Sf = fx(sigma_points(x, P, c), dt, u)
    xb = mean_x(Sf, Wm)
    Y = resid_x(Sf, xb)
    Pb = (Y * Wc[:, None]).T @ Y + Q
    Sf = sigma_points(xb, Pb, c)                                   # filterpy >= 1.4.5: re-draw from the prior
    # update
    Y = resid_x(Sf, xb)
    Z = hx(Sf)
    zb = mean_z(Z, Wm)
    E = resid_z(Z, zb)
    S = (E * Wc[:, None]).T @ E + R
    Pxz = (Y * Wc[:, None]).T @ E
    Us = chol_upper(S)
    # K = Pxz S^-1 by two triangular solves (extended precision has no LAPACK)
    W = np.zeros_like(Pxz.T)
    for i in range(Us.shape[0]):                                   # Us^T W = Pxz^T
        W[i] = (Pxz.T[i] - Us[:i, i] @ W[:i]) / Us[i, i]
    Kt = np.zeros_like(W)
    for i in range(Us.shape[0] - 1, -1, -1):                       # Us Kt = W
        Kt[i] = (W[i] - Us[i, i + 1:] @ Kt[i + 1:]) / Us[i, i]
    K = Kt.T
    y = z - zb
    y[1::2] = norm_angle(y[1::2])
    return xb + K @ y, Pb - K @ (S @ K.T)


def sample_index(n, count, seed=1):
    rng = np.random.default_rng(seed)
    i = rng.integers(0, n, count)
    j = rng.integers(0, n, count)
    return i, j


def main():
    n = 3 + 2 * L
    x0, P0, u, z = au.aug_case(F, L, seed=SEED)
    si, sj = sample_index(n, NSAMPLE)
    xs = np.zeros((F, n))
    Pd = np.zeros((F, n))
    Ps = np.zeros((F, NSAMPLE))
    for f in range(F):
        xt, Pt = step(x0[f], P0[f], u[f], z[f])
        xs[f] = xt.astype(np.float64)
        Pd[f] = np.diag(Pt).astype(np.float64)
        Ps[f] = Pt[si, sj].astype(np.float64)
        print("filter %d done" % f, flush=True)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ukf_aug_kat.npz")
    np.savez_compressed(out, F=F, L=L, seed=SEED, x=xs, P_diag=Pd, P_sample=Ps, sample_i=si, sample_j=sj,
                        precision="np.longdouble (x87 extended, eps 1.08e-19)")
    print("wrote", out)


if __name__ == "__main__":
    main()
