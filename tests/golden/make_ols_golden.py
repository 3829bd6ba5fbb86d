"""Generates tests/golden/ols_golden.npz with the REAL scikit-learn RANSACRegressor, called exactly as
/root/reference/line_detection.py:57-58 calls it (residual_threshold=3, everything else default) on
seeded 50-point windows, plus the index pairs its RandomState draws (replayed with the same seed:
fit() uses the random state for sample_without_replacement only).

    python tests/golden/make_ols_golden.py        (needs scikit-learn; the .npz is committed)
"""
import os
import sys
import warnings

import numpy as np
from sklearn import linear_model
from sklearn.utils.random import sample_without_replacement

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from lidar_slam_b200 import synth  # noqa: E402

T = 100


def windows():
    """(x, y) windows: walls of random rooms in line_detection's convention (x = d cos(radians(a)),
    zero ranges skipped, 50-point windows, short last window), and a few constructed edge cases."""
    out = []
    rng = np.random.default_rng(7)
    for s in range(12):
        w, h = rng.uniform(2000, 6000, 2)
        pose = (rng.uniform(-w / 2 + 300, w / 2 - 300), rng.uniform(-h / 2 + 300, h / 2 - 300), rng.uniform(-np.pi, np.pi))
        ang = np.arange(360.0)
        d = synth.room_ranges(pose, w, h, ang) + 2.0 * rng.standard_normal(360)
        d = np.where(rng.random(360) < 0.03, 0.0, d)
        keep = d != 0
        x = d[keep] * np.cos(np.radians(ang[keep]))
        y = d[keep] * np.sin(np.radians(ang[keep]))
        for i in range(0, len(x), 50):
            k = min(i + 50, len(x))
            if k - i >= 2:
                out.append((x[i:k].copy(), y[i:k].copy()))
    xs = np.linspace(0, 490, 50)
    out.append((xs, 0.5 * xs + 10))                                   # exact line: perfect score
    out.append((xs, np.full(50, 7.0)))                                # horizontal: zero total variance
    out.append((xs, 0.25 * xs + np.where(np.arange(50) % 5 == 0, 40.0, 0.0)))   # 20 % gross outliers
    out.append((np.repeat(xs[:25], 2), np.tile([0.0, 2.0], 25) + 0.1 * np.repeat(xs[:25], 2)))   # duplicate x
    out.append((xs[:2], np.array([1.0, 5.0])))                        # minimal window
    out.append((xs[:7], 3 * xs[:7] + rng.standard_normal(7)))         # short window
    return out


def main():
    ws = windows()
    W = len(ws)
    X = np.zeros((W, 50)); Y = np.zeros((W, 50)); N = np.zeros(W, np.int32)
    pairs = np.zeros((W, T, 2), np.int32)
    coef = np.zeros(W); icpt = np.zeros(W); ntr = np.zeros(W, np.int32); mask = np.zeros((W, 50), bool)
    for k, (x, y) in enumerate(ws):
        n = len(x)
        X[k, :n], Y[k, :n], N[k] = x, y, n
        seed = 1000 + k
        rs = np.random.RandomState(seed)
        for t in range(T):
            pairs[k, t] = sample_without_replacement(n, 2, random_state=rs)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r = linear_model.RANSACRegressor(residual_threshold=3, random_state=seed)      # line_detection.py:57
            r.fit(x.reshape(-1, 1), y)                                                      # :58
        coef[k], icpt[k], ntr[k] = r.estimator_.coef_[0], r.estimator_.intercept_, r.n_trials_
        mask[k, :n] = r.inlier_mask_
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ols_golden.npz")
    np.savez(out, x=X, y=Y, n=N, pairs=pairs, coef=coef, intercept=icpt, n_trials=ntr, inlier_mask=mask)
    print("wrote", out, W, "windows; trials:", ntr.tolist())


if __name__ == "__main__":
    main()
