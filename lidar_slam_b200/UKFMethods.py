"""Drop-in for the reference's ``UKFMethods`` module (/root/reference/UKFMethods.py).

The reference file does not parse (SyntaxError at :40 and :47) and has further defects
(SURVEY.md 8(a) U3/U5/U6); the functions below are the intended computations with only those
defects repaired, under the reference's names and argument meaning, for user code that calls them
directly.  The FILTER does not call them: ``systemClass.System(...).ukf`` evaluates the same
motion model, measurement model, circular means and wrapped residuals inside the CUDA kernel
(csrc/ukf.cuh, one launch per predict/update), which is what makes them the hot path.
"""
import math

import numpy as np

R = 50       # :6  wheel radius, mm
L = 200      # :7  wheel base, mm
dt = 0.01    # :8

_TWO_PI = 2 * np.pi


def normalize_angle(angle):                               # :10-14  -> (-pi, pi]
    angle = angle % _TWO_PI
    return angle - _TWO_PI if angle > np.pi else angle


def transition_function(x, dt, u):                        # :17-24  differential drive, u = [vl, vr]
    x = np.asarray(x, dtype=np.float64)
    th = x[2]
    step = np.zeros_like(x)
    step[0] = R / 2.0 * math.cos(th) * u[0] + R / 2.0 * math.cos(th) * u[1]
    step[1] = R / 2.0 * math.sin(th) * u[0] + R / 2.0 * math.sin(th) * u[1]
    step[2] = -1.0 * R / L * u[0] + 1.0 * R / L * u[1]
    return x + dt * step


def transfer_function(x, landmarks):                      # :26-34  [d0, b0, d1, b1, ...]
    out = np.empty(2 * len(landmarks))
    for k, lm in enumerate(landmarks):
        p = lm.get_pos()
        dx, dy = p[0] - x[0], p[1] - x[1]
        out[2 * k] = math.sqrt(dx ** 2 + dy ** 2)
        out[2 * k + 1] = normalize_angle(math.atan2(dy, dx) - x[2])
    return out


def state_mean(sigmas, Wm):                               # :37-45
    sigmas = np.asarray(sigmas)
    x = np.dot(Wm, sigmas)
    x[2] = math.atan2(np.dot(np.sin(sigmas[:, 2]), Wm), np.dot(np.cos(sigmas[:, 2]), Wm))
    return x


def z_mean(sigmas, Wm):                                   # :47-57
    sigmas = np.asarray(sigmas)
    z = np.dot(Wm, sigmas)
    for j in range(1, sigmas.shape[1], 2):
        z[j] = math.atan2(np.dot(np.sin(sigmas[:, j]), Wm), np.dot(np.cos(sigmas[:, j]), Wm))
    return z


def residual_x(a, b):                                     # :60-63
    y = np.asarray(a, dtype=np.float64) - b
    y[2] = normalize_angle(y[2])
    return y


def residual_h(a, b):                                     # :66-71
    y = np.asarray(a, dtype=np.float64) - b
    for j in range(1, len(y), 2):
        y[j] = normalize_angle(y[j])
    return y
