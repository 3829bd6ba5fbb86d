"""NumPy restatement of the UKF half of the path: filterpy's engine + the reference's callbacks.

TEST INFRASTRUCTURE (see oracle/__init__.py).  PARITY UNPINNED: the engine lives in filterpy
(absent here, not version-pinned by the reference) and the reference's own callback module does
not parse (/root/reference/UKFMethods.py:40,47 SyntaxError; systemClass.py:4-5 SyntaxError), so
nothing of this half has ever run upstream.  This file restates

  * the callbacks of /root/reference/UKFMethods.py:10-71 with ONLY the defects listed in
    SURVEY.md 8(a) U3/U5/U6 repaired (each repair is marked "repair:" below), and
  * the published algorithm of filterpy 1.4.x ``MerweScaledSigmaPoints``,
    ``unscented_transform`` and ``UnscentedKalmanFilter.predict/update`` (Labbe, "Kalman and
    Bayesian Filters in Python", ch. 10; van der Merwe's scaled sigma points), with the
    attribute surface the reference touches (systemClass.py:20-29: x, P, R, Q, predict, update).

This is the LITERAL form (np.dot, np.linalg.inv, scipy-style upper Cholesky, libm sin/cos):
what the reference would compute if it ran.  The operation ORDER the CUDA kernel is held to
(sequential sums, fused multiply-adds, Cholesky solve instead of inv) is pinned in
oracle/ukf.c; tests/test_oracle_ukf.py checks the two against each other.
"""
from math import atan2, cos, sin, sqrt

import numpy as np

# /root/reference/UKFMethods.py:6-8
R = 50        # wheel radius, mm
L = 200       # base length, mm
dt = 0.01     # 10 ms

# /root/reference/systemClass.py:7-10
LANDMARK_NUMBER = 8
VAR_DIST = 0.5 ** 2
VAR_ANGLE = 0.3 ** 2
DT = 0.005


def normalize_angle(angle):                         # UKFMethods.py:10-14
    angle = angle % (2 * np.pi)                     # Python floor-mod: result in [0, 2pi)
    if angle > np.pi:
        angle -= 2 * np.pi
    return angle


def transition_function(x, dt, u):                  # UKFMethods.py:17-24
    n = x.shape[0]
    fx = np.identity(n)
    # repair: the three rows were passed as three positional arguments of np.array (:20-22)
    bx = np.zeros((n, 2))
    bx[0] = [R / 2.0 * cos(x[2]), R / 2.0 * cos(x[2])]
    bx[1] = [R / 2.0 * sin(x[2]), R / 2.0 * sin(x[2])]
    bx[2] = [-1.0 * R / L, 1.0 * R / L]
    # (augmented state, n = 3 + 2*landmarks: the landmark rows of bx stay zero - SURVEY.md 8(a) C4)
    xBar = np.dot(fx, x) + dt * np.dot(bx, u)
    return xBar


class _Pos:
    """Stand-in for landmarking.Landmark in hx: only get_pos() is used (UKFMethods.py:30)."""

    def __init__(self, x, y):
        self._p = np.array([x, y], dtype=np.float64)

    def get_pos(self):
        return self._p


def transfer_function(x, landmarks):                # UKFMethods.py:26-34
    hx = []
    for lndmrk in landmarks:
        (px, py) = lndmrk.get_pos()[0], lndmrk.get_pos()[1]
        dist = sqrt((px - x[0]) ** 2 + (py - x[1]) ** 2)
        angle = atan2(py - x[1], px - x[0])
        hx.extend([dist, normalize_angle(angle - x[2])])
    return np.array(hx)


def transfer_function_augmented(x):
    """hx for the augmented state of BASELINE.json configs[3]: the landmark positions are
    x[3 + 2l], x[4 + 2l] instead of Landmark.get_pos() ('C4 is an extension')."""
    nl = (x.shape[0] - 3) // 2
    return transfer_function(x, [_Pos(x[3 + 2 * l], x[4 + 2 * l]) for l in range(nl)])


def state_mean(sigmas, Wm):                         # UKFMethods.py:37-45
    x = np.zeros(sigmas.shape[1])                   # repair: np.zeros(3) generalised for the augmented state
    # repair: unbalanced parentheses, math.sin/cos on arrays, sum_sim/sum_sin and no/np typos (:40-44)
    sum_sin = np.sum(np.dot(np.sin(sigmas[:, 2]), Wm))
    sum_cos = np.sum(np.dot(np.cos(sigmas[:, 2]), Wm))
    x[0] = np.sum(np.dot(sigmas[:, 0], Wm))
    x[1] = np.sum(np.dot(sigmas[:, 1], Wm))
    x[2] = atan2(sum_sin, sum_cos)
    for i in range(3, sigmas.shape[1]):
        x[i] = np.sum(np.dot(sigmas[:, i], Wm))
    return x


def z_mean(sigmas, Wm):                             # UKFMethods.py:47-57  (repair: ';' -> ':')
    nbrLmkrs = sigmas.shape[1]                      # really dim_z = 2 * landmarks
    x = np.zeros(nbrLmkrs)
    for z in range(0, nbrLmkrs, 2):
        # repair: misplaced parentheses np.sin(sigmas[:, z+1], Wm) (:52-53)
        sum_sin = np.sum(np.dot(np.sin(sigmas[:, z + 1]), Wm))
        sum_cos = np.sum(np.dot(np.cos(sigmas[:, z + 1]), Wm))
        x[z] = np.sum(np.dot(sigmas[:, z], Wm))
        x[z + 1] = atan2(sum_sin, sum_cos)
    return x


def residual_x(a, b):                               # UKFMethods.py:60-63
    x = a - b
    x[2] = normalize_angle(x[2])
    return x


def residual_h(a, b):                               # UKFMethods.py:66-71
    y = a - b
    for i in range(0, len(y), 2):
        y[i + 1] = normalize_angle(y[i + 1])
    return y


# ---- filterpy 1.4.x engine (restated) --------------------------------------------------------
def cholesky_upper(A):
    """scipy.linalg.cholesky(A) (upper, U^T U = A; only the upper triangle of A is read).
    Raises np.linalg.LinAlgError when A is not positive definite, like scipy."""
    A = np.asarray(A, dtype=np.float64)
    n = A.shape[0]
    U = np.zeros((n, n))
    for j in range(n):
        s = A[j, j] - np.dot(U[:j, j], U[:j, j])
        if not s > 0:
            raise np.linalg.LinAlgError("%d-th leading minor of the array is not positive definite" % (j + 1))
        U[j, j] = sqrt(s)
        for i in range(j + 1, n):
            U[j, i] = (A[j, i] - np.dot(U[:j, j], U[:j, i])) / U[j, j]
    return U


class MerweScaledSigmaPoints:
    def __init__(self, n, alpha, beta, kappa, sqrt_method=None, subtract=None):
        self.n = n
        self.alpha = alpha
        self.beta = beta
        self.kappa = kappa
        self.sqrt = cholesky_upper if sqrt_method is None else sqrt_method
        self.subtract = np.subtract if subtract is None else subtract
        self._compute_weights()

    def num_sigmas(self):
        return 2 * self.n + 1

    def sigma_points(self, x, P):
        n = self.n
        x = np.atleast_1d(np.asarray(x, dtype=np.float64))
        P = np.atleast_2d(P)
        lambda_ = self.alpha ** 2 * (n + self.kappa) - n
        U = self.sqrt((lambda_ + n) * P)
        sigmas = np.zeros((2 * n + 1, n))
        sigmas[0] = x
        for k in range(n):
            sigmas[k + 1] = self.subtract(x, -U[k])
            sigmas[n + k + 1] = self.subtract(x, U[k])
        return sigmas

    def _compute_weights(self):
        n = self.n
        lambda_ = self.alpha ** 2 * (n + self.kappa) - n
        c = .5 / (n + lambda_)
        self.Wc = np.full(2 * n + 1, c)
        self.Wm = np.full(2 * n + 1, c)
        self.Wc[0] = lambda_ / (n + lambda_) + (1 - self.alpha ** 2 + self.beta)
        self.Wm[0] = lambda_ / (n + lambda_)


def unscented_transform(sigmas, Wm, Wc, noise_cov=None, mean_fn=None, residual_fn=None):
    kmax, n = sigmas.shape
    x = np.dot(Wm, sigmas) if mean_fn is None else mean_fn(sigmas, Wm)
    if residual_fn is np.subtract or residual_fn is None:
        y = sigmas - x[np.newaxis, :]
        P = np.dot(y.T, np.dot(np.diag(Wc), y))
    else:
        P = np.zeros((n, n))
        for k in range(kmax):
            y = residual_fn(sigmas[k], x)
            P += Wc[k] * np.outer(y, y)
    if noise_cov is not None:
        P += noise_cov
    return (x, P)


class UnscentedKalmanFilter:
    """filterpy.kalman.UnscentedKalmanFilter, the subset the path uses.

    ``redraw=True`` (default) is the filterpy >= 1.4.5 behaviour: predict() re-draws the sigma
    points from the prior (x, P) so update() sees sigma points of the prior covariance including Q.
    """

    def __init__(self, dim_x, dim_z, dt, hx, fx, points, sqrt_fn=None, x_mean_fn=None, z_mean_fn=None,
                 residual_x=None, residual_z=None, redraw=True):
        self.x = np.zeros(dim_x)
        self.P = np.eye(dim_x)
        self.Q = np.eye(dim_x)
        self.R = np.eye(dim_z)
        self._dim_x = dim_x
        self._dim_z = dim_z
        self.points_fn = points
        self._dt = dt
        self._num_sigmas = points.num_sigmas()
        self.hx = hx
        self.fx = fx
        self.x_mean = x_mean_fn
        self.z_mean = z_mean_fn
        self.Wm, self.Wc = points.Wm, points.Wc
        self.residual_x = np.subtract if residual_x is None else residual_x
        self.residual_z = np.subtract if residual_z is None else residual_z
        self.sigmas_f = np.zeros((self._num_sigmas, self._dim_x))
        self.sigmas_h = np.zeros((self._num_sigmas, self._dim_z))
        self.K = np.zeros((dim_x, dim_z))
        self.y = np.zeros(dim_z)
        self.S = np.zeros((dim_z, dim_z))
        self.SI = np.zeros((dim_z, dim_z))
        self.inv = np.linalg.inv
        self.redraw = redraw
        self.x_prior = self.x.copy()
        self.P_prior = self.P.copy()
        self.x_post = self.x.copy()
        self.P_post = self.P.copy()

    def compute_process_sigmas(self, dt, fx=None, **fx_args):
        if fx is None:
            fx = self.fx
        sigmas = self.points_fn.sigma_points(self.x, self.P)
        for i, s in enumerate(sigmas):
            self.sigmas_f[i] = fx(s, dt, **fx_args)

    def predict(self, dt=None, UT=None, fx=None, **fx_args):
        if dt is None:
            dt = self._dt
        if UT is None:
            UT = unscented_transform
        self.compute_process_sigmas(dt, fx, **fx_args)
        self.x, self.P = UT(self.sigmas_f, self.Wm, self.Wc, self.Q, self.x_mean, self.residual_x)
        if self.redraw:
            self.sigmas_f = self.points_fn.sigma_points(self.x, self.P)
        self.x_prior = np.copy(self.x)
        self.P_prior = np.copy(self.P)

    def cross_variance(self, x, z, sigmas_f, sigmas_h):
        Pxz = np.zeros((sigmas_f.shape[1], sigmas_h.shape[1]))
        N = sigmas_f.shape[0]
        for i in range(N):
            dx = self.residual_x(sigmas_f[i], x)
            dz = self.residual_z(sigmas_h[i], z)
            Pxz += self.Wc[i] * np.outer(dx, dz)
        return Pxz

    def update(self, z, R=None, UT=None, hx=None, **hx_args):
        if z is None:
            self.x_post = self.x.copy()
            self.P_post = self.P.copy()
            return
        if hx is None:
            hx = self.hx
        if UT is None:
            UT = unscented_transform
        if R is None:
            R = self.R
        elif np.isscalar(R):
            R = np.eye(self._dim_z) * R
        sigmas_h = []
        for s in self.sigmas_f:
            sigmas_h.append(hx(s, **hx_args))
        self.sigmas_h = np.atleast_2d(sigmas_h)
        zp, self.S = UT(self.sigmas_h, self.Wm, self.Wc, R, self.z_mean, self.residual_z)
        self.SI = self.inv(self.S)
        Pxz = self.cross_variance(self.x, zp, self.sigmas_f, self.sigmas_h)
        self.K = np.dot(Pxz, self.SI)
        self.y = self.residual_z(z, zp)
        self.x = self.x + np.dot(self.K, self.y)
        self.P = self.P - np.dot(self.K, np.dot(self.S, self.K.T))
        self.x_post = self.x.copy()
        self.P_post = self.P.copy()


UKF = UnscentedKalmanFilter


class Robot:                                        # /root/reference/robot.py:3-14
    def __init__(self):
        self.pos = np.array([0., 0., 0.])
        self.dimX = self.pos.shape[0]
        self.name = "Main robot"

    def get_dim_x(self):
        return self.dimX

    def get_pos(self):
        return self.pos


class System:
    """/root/reference/systemClass.py:12-29 with its defects repaired: the two
    ``import filterpy.kalman import`` lines (:4-5), ``Robot()`` and the callbacks referenced
    without their module (:15,21), and ``dt=dt`` where only DT is defined (:21; repair: dt=DT)."""

    def __init__(self, landmarks, landmark_number=LANDMARK_NUMBER, alpha=0.0001, beta=2, kappa=0, redraw=True):
        self.robot = Robot()
        self.dt = DT
        self.varDist = VAR_DIST
        self.varAngle = VAR_ANGLE
        self.landmarks = landmarks
        self.landmark_number = landmark_number
        self.sigmas = MerweScaledSigmaPoints(n=self.robot.get_dim_x(), alpha=alpha, beta=beta, kappa=kappa)
        self.ukf = UKF(dim_x=self.robot.get_dim_x(), dim_z=2 * landmark_number, fx=transition_function,
                       hx=transfer_function, dt=self.dt, points=self.sigmas, x_mean_fn=state_mean,
                       z_mean_fn=z_mean, residual_x=residual_x, residual_z=residual_h, redraw=redraw)
        self.config_ukf()

    def config_ukf(self):                           # systemClass.py:25-29
        self.ukf.x = self.robot.get_pos()
        self.ukf.P = np.diag([.1, .1, 0.05])
        self.ukf.R = np.diag([self.varDist, self.varAngle] * self.landmark_number)
        self.ukf.Q = np.eye(3) * 0.001


def step_literal(x, P, u, z, lm_pos, dt=DT, alpha=1e-4, beta=2.0, kappa=0.0, Q=None,
                 var_dist=VAR_DIST, var_angle=VAR_ANGLE, redraw=True, do_predict=True, do_update=True):
    """One predict(u=u) + update(z, landmarks=...) of an n=3 filter with the literal engine.
    lm_pos [L, 2]; z [2L].  Returns (x, P)."""
    nl = len(lm_pos)
    pts = MerweScaledSigmaPoints(3, alpha, beta, kappa)
    f = UKF(3, 2 * nl, dt, transfer_function, transition_function, pts, x_mean_fn=state_mean,
            z_mean_fn=z_mean, residual_x=residual_x, residual_z=residual_h, redraw=redraw)
    f.x = np.array(x, dtype=np.float64)
    f.P = np.array(P, dtype=np.float64)
    f.Q = np.eye(3) * 0.001 if Q is None else np.array(Q, dtype=np.float64)
    f.R = np.diag([var_dist, var_angle] * nl)
    if do_predict:
        f.predict(u=np.asarray(u, dtype=np.float64))
    else:
        f.sigmas_f = pts.sigma_points(f.x, f.P)
    if do_update and nl > 0:
        f.update(np.asarray(z, dtype=np.float64), landmarks=[_Pos(p[0], p[1]) for p in lm_pos])
    return f.x, f.P


def step_literal_augmented(x, P, u, z, dt=DT, alpha=1e-4, beta=2.0, kappa=0.0, Q3=None, var_dist=VAR_DIST,
                           var_angle=VAR_ANGLE, redraw=True, do_predict=True, do_update=True):
    """One predict(u=u) + update(z) of an augmented filter (BASELINE.json configs[3])
    with the literal engine: x [3 + 2L] = pose + landmark positions, z [2L], hx reads the landmarks from
    the state (transfer_function_augmented), Q = the pose's 3x3 on the top-left block, zero elsewhere.
    Returns (x, P).  parity unpinned: the reference never builds this filter (it is a localiser)."""
    x = np.array(x, dtype=np.float64)
    n = x.shape[0]
    nl = (n - 3) // 2
    pts = MerweScaledSigmaPoints(n, alpha, beta, kappa)
    f = UKF(n, 2 * nl, dt, transfer_function_augmented, transition_function, pts, x_mean_fn=state_mean,
            z_mean_fn=z_mean, residual_x=residual_x, residual_z=residual_h, redraw=redraw)
    f.x = x
    f.P = np.array(P, dtype=np.float64)
    f.Q = np.zeros((n, n))
    f.Q[:3, :3] = np.eye(3) * 0.001 if Q3 is None else np.array(Q3, dtype=np.float64)
    f.R = np.diag([var_dist, var_angle] * nl)
    if do_predict:
        f.predict(u=np.asarray(u, dtype=np.float64))
    else:
        f.sigmas_f = pts.sigma_points(f.x, f.P)
    if do_update:
        f.update(np.asarray(z, dtype=np.float64).reshape(-1))
    return f.x, f.P
