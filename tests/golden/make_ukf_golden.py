"""Generates tests/golden/ukf_kat7.npz: the reference's own configuration
(systemClass.py:20-29: x0 = 0, P0 = diag(.1, .1, .05), R = diag([.25, .09] * 8), Q = .001 I,
alpha = 1e-4, beta = 2, kappa = 0, DT = .005) with u = [1.0, 1.1] and 8 landmarks on a 2 m square,
five predict+update steps of the LITERAL NumPy restatement of filterpy (oracle/ukf.py).  filterpy is
not installable here (parity unpinned), so the frozen values are the restatement's, cross-checked in
this script against an evaluation of the same equations in 60-digit arithmetic (mpmath, stored as
x_mp / P_mp) to show that the fp64 values are the filter's and not rounding noise of the 1e8-scale
sigma weights (alpha = 1e-4).

    python tests/golden/make_ukf_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ukf  # noqa: E402


def mp_step(x, P, u, z, lm, dt, alpha, beta, kappa, Q, var_dist, var_angle, wheel_r=50, base_l=200):
    """One predict + update in 60-digit arithmetic (same equations as oracle/ukf.py, redraw on)."""
    import mpmath as mp
    mp.mp.dps = 60
    n = 3
    lam = mp.mpf(alpha) ** 2 * (n + kappa) - n
    c = n + lam
    Wm = [mp.mpf(1) / (2 * c)] * 7
    Wc = list(Wm)
    Wm[0] = lam / c
    Wc[0] = lam / c + (1 - mp.mpf(alpha) ** 2 + beta)

    def norm_angle(a):
        a = a - 2 * mp.pi * mp.floor(a / (2 * mp.pi))
        return a - 2 * mp.pi if a > mp.pi else a

    def sigmas(x, P):
        U = mp.cholesky(c * P).T                      # upper, U^T U = cP
        out = [x.copy()]
        for k in range(n):
            out.append(x + U[k, :].T)
        for k in range(n):
            out.append(x - U[k, :].T)
        return out

    def fx(s):
        h, w = mp.mpf(wheel_r) / 2, mp.mpf(wheel_r) / base_l
        return mp.matrix([s[0] + dt * (h * mp.cos(s[2]) * (u[0] + u[1])), s[1] + dt * (h * mp.sin(s[2]) * (u[0] + u[1])),
                          s[2] + dt * (-w * u[0] + w * u[1])])

    def hx(s):
        out = []
        for px, py in lm:
            out += [mp.sqrt((px - s[0]) ** 2 + (py - s[1]) ** 2), norm_angle(mp.atan2(py - s[1], px - s[0]) - s[2])]
        return mp.matrix(out)

    sf = [fx(s) for s in sigmas(x, P)]
    xm = mp.matrix([sum(Wm[k] * sf[k][0] for k in range(7)), sum(Wm[k] * sf[k][1] for k in range(7)),
                    mp.atan2(sum(Wm[k] * mp.sin(sf[k][2]) for k in range(7)),
                             sum(Wm[k] * mp.cos(sf[k][2]) for k in range(7)))])

    def res_x(a, b):
        y = a - b
        y[2] = norm_angle(y[2])
        return y

    Pm = mp.matrix(Q)
    for k in range(7):
        y = res_x(sf[k], xm)
        Pm += Wc[k] * y * y.T
    sf = sigmas(xm, Pm)
    sh = [hx(s) for s in sf]
    dz = len(z)
    zp = mp.matrix(dz, 1)
    for j in range(0, dz, 2):
        zp[j] = sum(Wm[k] * sh[k][j] for k in range(7))
        zp[j + 1] = mp.atan2(sum(Wm[k] * mp.sin(sh[k][j + 1]) for k in range(7)),
                             sum(Wm[k] * mp.cos(sh[k][j + 1]) for k in range(7)))

    def res_h(a, b):
        y = a - b
        for j in range(1, dz, 2):
            y[j] = norm_angle(y[j])
        return y

    S = mp.diag([var_dist, var_angle] * (dz // 2))
    Pxz = mp.matrix(3, dz)
    for k in range(7):
        e = res_h(sh[k], zp)
        S += Wc[k] * e * e.T
        Pxz += Wc[k] * res_x(sf[k], xm) * e.T
    K = Pxz * mp.inverse(S)
    xn = xm + K * res_h(mp.matrix(z), zp)
    Pn = Pm - K * S * K.T
    return xn, Pn


def main():
    lm = np.array([[2000., 0.], [2000., 2000.], [0., 2000.], [-2000., 2000.], [-2000., 0.], [-2000., -2000.],
                   [0., -2000.], [2000., -2000.]])
    sysm = ukf.System([ukf._Pos(*p) for p in lm])
    f = sysm.ukf
    u = np.array([1.0, 1.1])
    rng = np.random.default_rng(7)
    import mpmath as mp
    xs, Ps, zs, xs_mp, Ps_mp = [], [], [], [], []
    x_mp, P_mp = mp.matrix([0, 0, 0]), mp.diag([mp.mpf("0.1"), mp.mpf("0.1"), mp.mpf("0.05")])
    truth = np.zeros(3)
    for step in range(5):
        truth = ukf.transition_function(truth, ukf.DT, u)
        z = ukf.transfer_function(truth, sysm.landmarks) + rng.normal(size=16) * np.tile([0.5, 0.3], 8) * 0.2
        f.predict(u=u)
        f.update(z, landmarks=sysm.landmarks)
        xs.append(f.x.copy())
        Ps.append(f.P.copy())
        zs.append(z)
        x_mp, P_mp = mp_step(x_mp, P_mp, [mp.mpf(float(v)) for v in u], [mp.mpf(float(v)) for v in z],
                             [(mp.mpf(float(a)), mp.mpf(float(b))) for a, b in lm], mp.mpf(ukf.DT), 1e-4, 2, 0,
                             mp.diag([mp.mpf("0.001")] * 3), mp.mpf(ukf.VAR_DIST), mp.mpf(ukf.VAR_ANGLE))
        xs_mp.append(np.array([float(v) for v in x_mp]))
        Ps_mp.append(np.array([[float(P_mp[i, j]) for j in range(3)] for i in range(3)]))
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ukf_kat7.npz")
    np.savez(out, lm=lm, u=u, z=np.array(zs), x=np.array(xs), P=np.array(Ps), x_mp=np.array(xs_mp),
             P_mp=np.array(Ps_mp))
    print("wrote", out, "\nx5 =", xs[-1], "\ndiag P5 =", np.diag(Ps[-1]))
    print("fp64 literal vs 60-digit: max |dx| = %.3g, max |dP|/max|P| = %.3g" % (
        np.abs(np.array(xs) - np.array(xs_mp)).max(),
        max(np.abs(a - b).max() / np.abs(b).max() for a, b in zip(Ps, Ps_mp))))


if __name__ == "__main__":
    main()
