"""OLS-flavour RANSAC of the reference's offline script.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Restates what
/root/reference/line_detection.py:51-73 asks of scikit-learn for every window of PNT_NBR = 50
points: ``RANSACRegressor(residual_threshold=3).fit(x[i:k], y[i:k])`` with every other parameter at
its default - base estimator LinearRegression (y on x, with intercept), min_samples = 2,
max_trials = 100, loss |y - y_hat|, inlier iff loss <= 3 (NON-strict), stop_probability = 0.99,
stop_n_inliers = stop_score = max_skips = inf.

PINNED: scikit-learn (1.9.0 here) is importable, so tests/golden/ols_golden.npz holds outputs of the
REAL ``RANSACRegressor`` on seeded windows together with the index pairs its RandomState drew
(generator: tests/golden/make_ols_golden.py); tests/test_oracle_ols.py checks this restatement and
its C twin (oracle/ransac_ols.c) against them: inlier masks, trial counts exactly, coefficients to 1e-9.

The loop (sklearn/linear_model/_ransac.py, RANSACRegressor.fit):
    n_best = 1; score_best = -inf; max_trials = 100; n_trials = 0
    while n_trials < max_trials:
        n_trials += 1
        pair = 2 distinct indices                            (sample_without_replacement)
        model = OLS through the 2 points                     (LinearRegression.fit)
        r = |y - (x*coef + intercept)| over ALL points;  inl = r <= thr;  cnt = sum(inl)
        if cnt < n_best: continue
        score = R^2 of the model on the inliers              (LinearRegression.score -> r2_score)
        if cnt == n_best and score < score_best: continue
        accept; max_trials = min(max_trials, dynamic_max_trials(cnt, n, 2, 0.99))
    no accepted trial -> ValueError;  final model = OLS on the accepted inliers
OLS with one feature in closed form: x0 = mean(x), y0 = mean(y), coef = sum((x-x0)(y-y0)) /
sum((x-x0)^2) (0 when the denominator is 0: the minimum-norm solution LAPACK returns),
intercept = y0 - x0*coef.  sklearn reaches the same numbers through scipy.linalg.lstsq (SVD); the two
differ in the last bits, which is why coefficients are compared to 1e-9 and decisions exactly.
Sums use the fixed 32-lane tree of oracle/ransac_line.py (``tree32_sum``) so that the C twin and the
kernel can be bit-identical to each other.
"""
import math

import numpy as np

from .philox import sample_pairs
from .ransac_line import tree32_sum

RESIDUAL_THRESHOLD = 3        # line_detection.py:57
MAX_TRIALS = 100              # sklearn default
STOP_PROBABILITY = 0.99       # sklearn default
PNT_NBR = 50                  # line_detection.py:9
_EPSILON = float(np.spacing(1))


def dynamic_max_trials(n_inliers, n_samples, min_samples=2, probability=STOP_PROBABILITY):
    """sklearn _dynamic_max_trials."""
    inlier_ratio = n_inliers / float(n_samples)
    nom = max(_EPSILON, 1 - probability)
    denom = max(_EPSILON, 1 - inlier_ratio ** min_samples)
    if nom == 1:
        return 0
    if denom == 1:
        return float("inf")
    return abs(float(math.ceil(math.log(nom) / math.log(denom))))


def _msum(v, mask):
    """tree32 sum of v over the positions selected by mask, each element staying in the lane of its
    ORIGINAL index (the others contribute +0.0) - the order the kernel's warp uses."""
    return tree32_sum(np.where(mask, v, 0.0))


def ols(x, y, mask):
    """Least squares y ~ coef*x + intercept over the masked points, closed form (module docstring)."""
    n = int(mask.sum())
    x0 = _msum(x, mask) / n
    y0 = _msum(y, mask) / n
    xc = x - x0
    yc = y - y0
    den = _msum(xc * xc, mask)
    coef = _msum(xc * yc, mask) / den if den != 0 else 0.0
    return coef, y0 - x0 * coef


def ols_pair(xi, yi, xj, yj):
    """The same through two points (what LinearRegression.fit does with a 2-sample subset)."""
    x0, y0 = (xi + xj) / 2.0, (yi + yj) / 2.0
    xci, xcj, yci, ycj = xi - x0, xj - x0, yi - y0, yj - y0
    den = xci * xci + xcj * xcj
    coef = (xci * yci + xcj * ycj) / den if den != 0 else 0.0
    return coef, y0 - x0 * coef


def r2_on(x, y, mask, coef, intercept):
    """LinearRegression.score on the masked points: r2_score with force_finite (1 for a perfect fit, 0
    when only the denominator vanishes); NaN for fewer than two samples."""
    n = int(mask.sum())
    if n < 2:
        return float("nan")
    pred = x * coef + intercept
    num = _msum((y - pred) * (y - pred), mask)
    y0 = _msum(y, mask) / n
    den = _msum((y - y0) * (y - y0), mask)
    if num != 0 and den != 0:
        return 1.0 - num / den
    if num != 0:
        return 0.0
    return 1.0


def fit_window(x, y, pairs, thr=RESIDUAL_THRESHOLD, max_trials=MAX_TRIALS, probability=STOP_PROBABILITY):
    """One window.  x, y [n] float64; pairs [T, 2] the index pairs of trials 0.. in order.  Returns a dict
    (coef, intercept, inlier_mask, n_inliers, n_trials, best_trial, score) or raises ValueError like
    sklearn when no trial is accepted."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n = len(x)
    if n < 2:
        raise ValueError("`min_samples` may not be larger than number of samples: n_samples = %d." % n)
    n_best, score_best, mask_best, best_trial = 1, -math.inf, None, -1
    limit, n_trials = max_trials, 0
    while n_trials < limit:
        i, j = int(pairs[n_trials][0]), int(pairs[n_trials][1])
        n_trials += 1
        coef, intercept = ols_pair(x[i], y[i], x[j], y[j])
        r = np.abs(y - (x * coef + intercept))
        inl = r <= thr
        cnt = int(inl.sum())
        if cnt < n_best:
            continue
        score = r2_on(x, y, inl, coef, intercept)
        if cnt == n_best and score < score_best:
            continue
        n_best, score_best, mask_best, best_trial = cnt, score, inl, n_trials - 1
        limit = min(limit, dynamic_max_trials(n_best, n, 2, probability))
    if mask_best is None:
        raise ValueError("RANSAC could not find a valid consensus set.")
    coef, intercept = ols(x, y, mask_best)
    return {"coef": coef, "intercept": intercept, "inlier_mask": mask_best, "n_inliers": n_best,
            "n_trials": n_trials, "best_trial": best_trial, "score": score_best}


def philox_pairs(n, trials, seed, scan, window):
    """The library's own sampler (same stream layout as the LineModelND flavour)."""
    i, j = sample_pairs(seed, scan, window, trials, n)
    return np.stack([i, j], -1)
