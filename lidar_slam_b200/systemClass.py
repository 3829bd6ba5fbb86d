"""Drop-in for the reference's ``systemClass`` module (/root/reference/systemClass.py).

``System(landmarks)`` keeps the reference's attributes (robot, dt, varDist, varAngle, landmarks,
sigmas, ukf) and ``config_ukf()`` (:14-29).  ``ukf`` exposes the filterpy surface the reference
touches - assignable ``x, P, R, Q`` (:26-29) and ``predict(**fx_args)`` / ``update(z, **hx_args)`` -
but each call is one launch of the CUDA kernel (``lslam_ukf_predict`` / ``lslam_ukf_update``) on a
batch of one filter; ``DeviceUKF`` with ``filters=F`` is the batched form of the same object.  The
reference file itself does not parse (:4-5) and references ``dt`` where only ``DT`` exists (:21);
the repair is the one documented in oracle/ukf.py.  No CPU fallback.
"""
import numpy as np

from . import batch
from .robot import Robot
from .UKFMethods import residual_h, residual_x, state_mean, transfer_function, transition_function, z_mean  # noqa: F401

LANDMARK_NUMBER = 8     # :7
VAR_DIST = 0.5 ** 2     # :8
VAR_ANGLE = 0.3 ** 2    # :9
DT = 0.005              # :10


class MerweSigmas:
    """The parameters of filterpy's MerweScaledSigmaPoints(n, alpha, beta, kappa) (:20) and its
    weights; the points themselves are drawn inside the kernel."""

    def __init__(self, n, alpha, beta, kappa):
        if n != 3:
            raise ValueError("the CUDA filter is the reference's n = 3 pose filter")
        self.n, self.alpha, self.beta, self.kappa = n, alpha, beta, kappa
        lam = alpha ** 2 * (n + kappa) - n
        c = .5 / (n + lam)
        self.Wm = np.full(2 * n + 1, c)
        self.Wc = np.full(2 * n + 1, c)
        self.Wm[0] = lam / (n + lam)
        self.Wc[0] = lam / (n + lam) + (1 - alpha ** 2 + beta)

    def num_sigmas(self):
        return 2 * self.n + 1


class DeviceUKF:
    """filterpy-shaped front of the batched CUDA filter.  With ``filters=1`` (the reference's use)
    x is [3], P [3,3]; with ``filters=F`` they are [F,3] / [F,3,3].  State lives on the device
    between calls; reading ``x`` / ``P`` copies it back."""

    def __init__(self, dim_x, dim_z, dt, points, filters=1, device=None, redraw=True):
        import torch
        if dim_x != 3 or dim_z % 2:
            raise ValueError("dim_x must be 3 and dim_z even (range/bearing pairs)")
        self._torch = torch
        self.eng = batch.engine(device)
        self.F, self.single = int(filters), int(filters) == 1
        self._dim_x, self._dim_z, self._dt = dim_x, dim_z, dt
        self.points_fn = points
        self.Wm, self.Wc = points.Wm, points.Wc
        self.redraw = redraw
        self.state = self.eng.alloc_ukf(self.F, dim_z // 2)
        self._sig = torch.zeros((self.F, 7, 3), dtype=torch.float64, device=self.eng.device)
        self._have_sig = False
        self._R = np.eye(dim_z)
        self._Q = np.eye(dim_x)

    # -- state: numpy on the way in and out --
    def _get(self, key):
        v = self.state[key].cpu().numpy()
        return v[0] if self.single else v

    def _set(self, key, value, shape):
        t = self._torch
        v = np.broadcast_to(np.asarray(value, dtype=np.float64), (self.F,) + shape)
        self.state[key] = t.from_numpy(np.array(v, order="C")).to(self.eng.device)
        self._have_sig = False

    x = property(lambda self: self._get("x"), lambda self, v: self._set("x", v, (3,)))
    P = property(lambda self: self._get("P"), lambda self, v: self._set("P", v, (3, 3)))

    @property
    def Q(self):
        return self._Q

    @Q.setter
    def Q(self, v):
        self._Q = np.array(v, dtype=np.float64).reshape(3, 3)

    @property
    def R(self):
        return self._R

    @R.setter
    def R(self, v):
        v = np.array(v, dtype=np.float64)
        d = np.diag(v)
        if v.ndim != 2 or np.any(v != np.diag(d)) or np.any(d[0::2] != d[0]) or np.any(d[1::2] != d[1]):
            raise ValueError("R must be diag([var_dist, var_angle] * landmarks) as systemClass.py:28 builds it")
        self._R = v

    def _params(self, dt):
        p = self.points_fn
        from . import UKFMethods                                  # R, L read at call time: user edits of the module constants count
        return self.eng.ukf_params(dt=dt, wheel_r=float(UKFMethods.R), base_l=float(UKFMethods.L), alpha=p.alpha, beta=p.beta,
                                   kappa=p.kappa, var_dist=self._R[0, 0], var_angle=self._R[1, 1], Q=self._Q)

    def _dev(self, a, shape):
        a = np.broadcast_to(np.asarray(a, dtype=np.float64), (self.F,) + shape)
        return self._torch.from_numpy(np.array(a, order="C")).to(self.eng.device)     # a writable copy (broadcast views are not)

    def predict(self, dt=None, u=None, **fx_args):
        """filterpy's predict(dt=None, **fx_args); the reference's fx takes ``u = [vl, vr]``."""
        if fx_args:
            raise TypeError("transition_function takes only u (UKFMethods.py:17)")
        self.state["params"] = self._params(self._dt if dt is None else dt)
        uu = self._dev([0.0, 0.0] if u is None else u, (2,))
        E = self.eng
        E.ukf_step(self.state, None, None, u=uu, flags=E.UKF_PREDICT | (E.UKF_REDRAW if self.redraw else 0),
                   sigmas_f=self._sig)
        self._have_sig = True

    def update(self, z, landmarks=None, R=None, **hx_args):
        """filterpy's update(z, R=None, **hx_args); the reference's hx takes ``landmarks`` (objects
        with get_pos(), UKFMethods.py:26-30).  z = [d0, b0, d1, b1, ...]."""
        if hx_args:
            raise TypeError("transfer_function takes only landmarks (UKFMethods.py:26)")
        if z is None:
            return
        if R is not None:
            self.R = R
        if self.single:
            pos = np.array([lm.get_pos()[:2] for lm in landmarks], dtype=np.float64).reshape(-1, 2)
        else:
            pos = np.asarray(landmarks, dtype=np.float64)           # batched form: [F, L, 2] positions
        Lm = pos.shape[-2]
        zz = self._dev(np.asarray(z, dtype=np.float64).reshape((-1, Lm, 2) if not self.single else (Lm, 2)), (Lm, 2))
        self.state["params"] = self._params(self._dt)
        E = self.eng
        E.ukf_step(self.state, zz, self._dev(pos, (Lm, 2)), flags=E.UKF_UPDATE,
                   sigmas_f=self._sig if self._have_sig else None)
        self._have_sig = False

    @property
    def status(self):
        return self._get("status")


class System:
    def __init__(self, landmarks, filters=1, device=None):
        self.robot = Robot()
        self.dt = DT
        self.varDist = VAR_DIST
        self.varAngle = VAR_ANGLE
        self.landmarks = landmarks
        self.sigmas = MerweSigmas(n=self.robot.get_dim_x(), alpha=0.0001, beta=2, kappa=0)
        self.ukf = DeviceUKF(dim_x=self.robot.get_dim_x(), dim_z=2 * LANDMARK_NUMBER, dt=self.dt,
                             points=self.sigmas, filters=filters, device=device)
        self.config_ukf()

    def config_ukf(self):                                   # :25-29
        self.ukf.x = self.robot.get_pos()
        self.ukf.P = np.diag([.1, .1, 0.05])
        self.ukf.R = np.diag([self.varDist, self.varAngle] * LANDMARK_NUMBER)
        self.ukf.Q = np.eye(3) * 0.001
