"""Drop-in for the reference's offline script ``line_detection`` (/root/reference/line_detection.py).

Same names: ``CHUNK_SIZE``, ``PNT_NBR`` (:8-9), ``get_data(file_name)`` (:12-23).  The body of the
script's ``__main__`` block (:26-77) is the function ``detect_lines``: polar text records ->
cartesian (:31-35) -> one robust line per window of PNT_NBR points (:51-73).  The per-window
``RANSACRegressor(residual_threshold=3).fit`` (:57-58) of every window runs in ONE launch of
``lslam_ransac_ols`` (csrc/ransac_ols.cuh); there is no CPU fallback.  The matplotlib figure (:44-45,
67-70,76) is left out (headless: no graphics stack): what the script draws is returned instead - per
window the inlier/outlier masks, ``line_X = arange(min x, max x)`` and ``line_y = predict(line_X)``.

Differences at the edges: samples come from the library's counter-based Philox stream (``SEED``) instead
of an unseeded MT19937 (pass ``pairs`` to replay recorded draws); the script's debugging ``print`` calls
(:30,41,43,52,62-64,75) only happen with ``COMPAT_PRINTS = True``.
"""
import re

import numpy as np

from . import batch

CHUNK_SIZE = 36      # :8 (unused by the reference too)
PNT_NBR = 50         # :9
RESIDUAL_THRESHOLD = 3   # :57

SEED = 0
COMPAT_PRINTS = False
DEVICE = None


def _say(*a):
    if COMPAT_PRINTS:
        print(*a)


def get_data(file_name):
    """:12-23 - every number in the text file, in order, as floats (distance, angle, distance, ...).
    Returns None after printing the reference's message when the file cannot be opened."""
    values = []
    try:
        file_descriptor = open(file_name, 'r')
    except IOError:
        print("Could not open file. Exiting...\n")
    else:
        lines = file_descriptor.readlines()
        for line in lines:
            values += re.findall(r"[-+]?\d*\.\d+|\d+", line)
        file_descriptor.close()
        return [float(item) for item in values]


def polar_records_to_points(data):
    """:31-35 - (distance, angle_deg) pairs, zero distances skipped, x = d cos(radians(a)), y = d sin(...)."""
    d = np.asarray(data[0::2], dtype=np.float64)
    a = np.asarray(data[1::2], dtype=np.float64)
    m = min(len(d), len(a))
    d, a = d[:m], a[:m]
    keep = d != 0
    return d[keep] * np.cos(np.radians(a[keep])), d[keep] * np.sin(np.radians(a[keep]))


def _mask_to_bool(words, n):
    k = np.arange(n)
    return ((words[k >> 5].astype(np.uint32) >> (k & 31).astype(np.uint32)) & 1).astype(bool)


def detect_lines(x, y, pairs=None, seed=None, window=None):
    """One robust line per window of PNT_NBR consecutive points (:51-73).  Returns a list, one dict per
    window: coef, intercept (``estimator_.coef_[0]``, ``estimator_.intercept_``), inlier_mask,
    outlier_mask (:59-60), n_trials, line_X, line_y (:66-67), inliers_x (what the script appends to
    ``finalPointsX``, :73).  A window for which scikit-learn would raise ValueError raises it here."""
    import torch
    eng = batch.engine(DEVICE)
    window = PNT_NBR if window is None else int(window)
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    n = x.shape[0]
    if n == 0:
        return []
    xy = torch.from_numpy(np.ascontiguousarray(np.stack([x, y], -1)[None])).to(eng.device)
    if pairs is not None:
        pairs = torch.from_numpy(np.ascontiguousarray(pairs, dtype=np.int32)[None]).to(eng.device)
    r = eng.ransac_ols(xy, window=window, thr=RESIDUAL_THRESHOLD, seed=SEED if seed is None else seed, pairs=pairs)
    fit = r["fit"][0].cpu().numpy()
    ninl = r["ninl"][0].cpu().numpy()
    ntr = r["n_trials"][0].cpu().numpy()
    masks = r["inl_mask"][0].cpu().numpy().view(np.uint32)
    out = []
    for c, i in enumerate(range(0, n, window)):
        k = min(i + window, n)
        _say(i)
        if ninl[c] == batch._lib.LSLAM_ERR_TOO_FEW_POINTS:
            raise ValueError("`min_samples` may not be larger than number of samples: n_samples = %d." % (k - i))
        if ninl[c] == batch._lib.LSLAM_ERR_NO_MODEL:
            raise ValueError("RANSAC could not find a valid consensus set. All `max_trials` iterations were "
                             "skipped because each randomly chosen sub-sample failed the passing criteria.")
        inl = _mask_to_bool(masks[c], k - i)
        coef, icpt = float(fit[c, 0]), float(fit[c, 1])
        line_X = np.arange(fit[c, 3], fit[c, 4])[:, np.newaxis]
        _say("Param: ", np.array([coef]))
        out.append({"coef": coef, "intercept": icpt, "inlier_mask": inl, "outlier_mask": np.logical_not(inl),
                    "n_trials": int(ntr[c]), "score": float(fit[c, 2]), "line_X": line_X,
                    "line_y": line_X[:, 0] * coef + icpt, "inliers_x": x[i:k][inl]})
    return out


def main(file_name="./data/scan_prop.txt"):
    """The script's ``__main__`` flow without the figure; returns (windows, finalPointsX)."""
    data = get_data(file_name)
    if data is None:
        return None
    _say(data)
    x, y = polar_records_to_points(data)
    windows = detect_lines(x, y)
    finalPointsX = [w["inliers_x"] for w in windows]
    _say(finalPointsX)
    return windows, finalPointsX


if __name__ == '__main__':
    main()
