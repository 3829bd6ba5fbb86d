"""Generates tests/golden/filterpy_golden.npz from the REAL filterpy - run it wherever filterpy is installed (it is
not in this project's image, which is why parity of the UKF restatement is "unpinned"):

    python tests/golden/make_filterpy_golden.py         # writes tests/golden/filterpy_golden.npz

It wires filterpy's own engine the way /root/reference/systemClass.py:14-29 does -

    MerweScaledSigmaPoints(n=3, alpha=.0001, beta=2, kappa=0)
    UnscentedKalmanFilter(dim_x=3, dim_z=16, fx=transition_function, hx=transfer_function, dt=DT, points=...,
                          x_mean_fn=state_mean, z_mean_fn=z_mean, residual_x=residual_x, residual_z=residual_h)
    x = 0, P = diag(.1, .1, .05), R = diag([.25, .09] * 8), Q = .001 I

- around the reference's callbacks (/root/reference/UKFMethods.py:10-71 with its syntax defects repaired exactly as
oracle/ukf.py documents; the callbacks are imported from there so both sides run the SAME Python callbacks and the
only thing under test is the engine: sigma points, unscented transform, gain, covariance update), then runs

  case "kat7"  the five predict(u=[1.0, 1.1]) + update(z, landmarks=8 on a 2 m square) steps of tests/golden/ukf_kat7.npz
               with the z recorded there
  case "aug"   one predict + update of the augmented filter (BASELINE.json configs[3] at n = 3 + 2 * 12) on
               tests/ukf_aug_util.aug_case(2, 12, seed=2): hx reads the landmarks from the state

and records x, P after every step.  tests/test_oracle_pins.py then holds oracle/ukf.py (1e-9) and, on a GPU, the CUDA
kernels (north_star's 1e-5) to these values.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ukf as ou  # noqa: E402  (callbacks only; the engine below is filterpy's)
from tests import ukf_aug_util as au  # noqa: E402


def main():
    import filterpy
    from filterpy.kalman import MerweScaledSigmaPoints, UnscentedKalmanFilter
    here = os.path.dirname(os.path.abspath(__file__))
    k7 = np.load(os.path.join(here, "ukf_kat7.npz"))
    lm, u = k7["lm"], k7["u"]
    landmarks = [ou._Pos(*p) for p in lm]
    pts = MerweScaledSigmaPoints(n=3, alpha=.0001, beta=2, kappa=0)
    f = UnscentedKalmanFilter(dim_x=3, dim_z=16, fx=ou.transition_function, hx=ou.transfer_function, dt=ou.DT, points=pts,
                              x_mean_fn=ou.state_mean, z_mean_fn=ou.z_mean, residual_x=ou.residual_x,
                              residual_z=ou.residual_h)
    f.x = np.array([0., 0., 0.])
    f.P = np.diag([.1, .1, 0.05])
    f.R = np.diag([ou.VAR_DIST, ou.VAR_ANGLE] * 8)
    f.Q = np.eye(3) * 0.001
    xs, Ps, xpri, Ppri = [], [], [], []
    for k in range(5):
        f.predict(u=u)
        xpri.append(f.x.copy()); Ppri.append(f.P.copy())
        f.update(k7["z"][k], landmarks=landmarks)
        xs.append(f.x.copy()); Ps.append(f.P.copy())

    x0, P0, ua, za = au.aug_case(2, 12, seed=2)
    n = x0.shape[1]
    xa, Pa = [], []
    for i in range(x0.shape[0]):
        pa = MerweScaledSigmaPoints(n=n, alpha=.0001, beta=2, kappa=0)
        g = UnscentedKalmanFilter(dim_x=n, dim_z=24, fx=ou.transition_function, hx=ou.transfer_function_augmented, dt=ou.DT,
                                  points=pa, x_mean_fn=ou.state_mean, z_mean_fn=ou.z_mean, residual_x=ou.residual_x,
                                  residual_z=ou.residual_h)
        g.x = x0[i].copy()
        g.P = P0[i].copy()
        g.R = np.diag([ou.VAR_DIST, ou.VAR_ANGLE] * 12)
        g.Q = np.zeros((n, n))
        g.Q[:3, :3] = np.eye(3) * 0.001
        g.predict(u=ua[i])
        g.update(za[i].reshape(-1))
        xa.append(g.x.copy()); Pa.append(g.P.copy())
    out = os.path.join(here, "filterpy_golden.npz")
    np.savez(out, lm=lm, u=u, z=k7["z"], x=np.array(xs), P=np.array(Ps), x_prior=np.array(xpri), P_prior=np.array(Ppri),
             aug_x0=x0, aug_P0=P0, aug_u=ua, aug_z=za, aug_x=np.array(xa), aug_P=np.array(Pa),
             filterpy_version=filterpy.__version__, numpy_version=np.__version__)
    print("wrote", out, "from filterpy", filterpy.__version__)


if __name__ == "__main__":
    main()
