"""Shared helpers of the OLS-flavour RANSAC tests (line_detection.py:51-73)."""
import os

import numpy as np

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ols_golden.npz"))
KEYS = ("fit", "ninl", "n_trials", "best_trial", "inl_mask")


def golden_batch():
    """The golden windows as a batch of one-window scans: xy [W, 50, 2], npts [W], pairs [W, 1, 100, 2]."""
    xy = np.stack([GOLD["x"], GOLD["y"]], -1)
    return np.ascontiguousarray(xy), GOLD["n"].astype(np.int32), np.ascontiguousarray(GOLD["pairs"][:, None])


def mask_bool(words, n):
    k = np.arange(n)
    return ((words[k >> 5] >> (k & 31).astype(np.uint32)) & 1).astype(bool)


def assert_ols_equal(want, got, what=""):
    """Bit-exact: integers, masks and the float records (NaN patterns compared as NaN)."""
    for k in KEYS:
        a, b = np.asarray(want[k]), np.asarray(got[k])
        if a.dtype.kind == "f":
            same = (a == b) | (np.isnan(a) & np.isnan(b))
            assert same.all(), "%s %s differs at %s: %r vs %r" % (what, k, np.argwhere(~same)[:3].tolist(),
                                                                   a[~same][:3], b[~same][:3])
        else:
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "%s %s differs" % (what, k)


def wall_windows(B, N=360, seed=0, noise=2.0, outliers=0.05):
    """B scans of N points along random lines with noise and gross outliers, in x order."""
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(-3000, 3000, (B, N)), axis=1)
    a = rng.uniform(-2, 2, (B, 1))
    b = rng.uniform(-1000, 1000, (B, 1))
    y = a * x + b + noise * rng.standard_normal((B, N))
    y = np.where(rng.random((B, N)) < outliers, y + rng.uniform(-300, 300, (B, N)), y)
    return np.ascontiguousarray(np.stack([x, y], -1))
