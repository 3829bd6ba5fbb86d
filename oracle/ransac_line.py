"""NumPy restatement of the RANSAC line / landmark extraction path (oracle side).

TEST INFRASTRUCTURE (see oracle/__init__.py).  PARITY UNPINNED: the arithmetic lives in
scikit-image (absent, un-pinned); this restates the published algorithm of
``skimage.measure.ransac`` + ``LineModelND`` (<= 0.15 era, matching the reference's
cpython-35 bytecode) driven exactly as /root/reference/ransac_functions.py:20-31 drives
it.  The same semantics are restated in C in oracle/ransac_line.c (checked against this
file in tests/test_oracle_ransac.py).

Operation order (the DEFINITION the CUDA kernels are bit-exact against for inlier
counts, masks, winning trial and sum of squared residuals):

  2-point model (LineModelND.estimate, 2 samples)
      o   = (p_i + p_j) / 2
      c_i = p_i - o ; c_j = p_j - o ; d = c_j - c_i
      n   = sqrt(d_x*d_x + d_y*d_y) ; if n != 0: d = d / n          (n == 0 keeps d = 0)
  residual (LineModelND.residuals)                       every op rounded, NO fma
      v = p - o ; t = v_x*d_x + v_y*d_y
      r = sqrt((v_x - t*d_x)^2 + (v_y - t*d_y)^2)
  consensus (skimage.measure.ransac)
      inlier = r < residual_threshold                    strict
      ssum   = sum_k r_k*r_k over ALL points             summation tree: tree32_sum()
      accept trial if count > best or (count == best and ssum < best_ssum)
      break after an accepted trial with ssum <= 0       (stop_residuals_sum=0; the
                                                         other two stop rules cannot fire
                                                         with stop_sample_num=inf,
                                                         stop_probability=1)
  final model: LineModelND.estimate(data[inliers])       SVD principal axis if > 2 inliers

Where upstream calls BLAS/einsum (``x.dot(x)``, ``@``, ``einsum('ij,ij->i')``), whose use of
FMA depends on the CPU it runs on, the restatement writes the products and sums out
explicitly so the result is platform-independent.  ``np.sum`` (pairwise, an
implementation detail of NumPy) is replaced by the fixed 32-lane tree below; it only
matters for breaking exact ties on inlier count between trials whose sums differ by a
rounding error.
"""
import numpy as np

from .philox import sample_pairs

# /root/reference/ransac_functions.py:9-12
THRESHOLD = 20
MAX_TRIALS = 100
MIN_SAMPLES = 2
MIN_POINTS = 100


def tree32_sum(v):
    """Sum of a 1-D float64 array with the fixed tree shared by oracle and kernels.

    lane l accumulates v[l], v[l+32], v[l+64], ... sequentially from 0.0; the 32 lane
    partials are then combined by the xor-butterfly 16, 8, 4, 2, 1 (result of lane 0).
    """
    v = np.asarray(v, dtype=np.float64)
    n = v.shape[0]
    k = (n + 31) // 32
    pad = np.zeros(k * 32, dtype=np.float64)
    pad[:n] = v
    rows = pad.reshape(k, 32)
    p = np.zeros(32, dtype=np.float64)
    for r in range(k):
        p = p + rows[r]
    for half in (16, 8, 4, 2, 1):
        p = p[:half] + p[half:2 * half]
    return float(p[0])


def line_from_pair(p_i, p_j):
    """LineModelND.estimate for exactly two points -> (origin[2], direction[2])."""
    p_i = np.asarray(p_i, dtype=np.float64)
    p_j = np.asarray(p_j, dtype=np.float64)
    o = (p_i + p_j) / 2.0
    c_i = p_i - o
    c_j = p_j - o
    d = c_j - c_i
    n = np.sqrt(d[0] * d[0] + d[1] * d[1])
    if n != 0:
        d = d / n
    return o, d


def line_residuals(data, o, d):
    """LineModelND.residuals: orthogonal distance of every point to the line (o, d)."""
    vx = data[:, 0] - o[0]
    vy = data[:, 1] - o[1]
    t = vx * d[0] + vy * d[1]
    rx = vx - t * d[0]
    ry = vy - t * d[1]
    return np.sqrt(rx * rx + ry * ry)


def line_estimate(pts):
    """LineModelND.estimate for k >= 2 points (2-point branch or SVD principal axis)."""
    pts = np.asarray(pts, dtype=np.float64)
    k = pts.shape[0]
    if k == 2:
        return line_from_pair(pts[0], pts[1])
    if k < 2:
        raise ValueError("At least 2 input points needed.")
    o = pts.mean(axis=0)
    _, _, v = np.linalg.svd(pts - o, full_matrices=False)
    return o, v[0]


def ransac_line(data, residual_threshold=THRESHOLD, max_trials=MAX_TRIALS,
                seed=0, scan=0, chunk=0, pairs=None, early_exit=True, details=False):
    """skimage.measure.ransac(data, LineModelND, 2, residual_threshold, max_trials).

    ``pairs`` overrides the Philox sampler with explicit (i[T], j[T]) index arrays.
    Returns ``((origin, direction), inliers[N] bool, info)``.  Like upstream: returns
    ``(None, None, info)`` when no trial was ever accepted (the reference then raises
    AttributeError at ransac_functions.py:25) and raises ValueError from the final
    ``estimate`` when the accepted trial has fewer than 2 inliers (a trial with ZERO inliers
    is accepted by the rule ``count == best and ssum < inf``).  Both cases are
    LSLAM_ERR_NO_MODEL in the C ABI.  ``info`` holds the winning trial, its pair,
    inlier count and ssum, and with ``details=True`` the per-trial counts/ssums.
    """
    data = np.asarray(data, dtype=np.float64)
    n = data.shape[0]
    if n < MIN_SAMPLES:
        raise ValueError("`min_samples` must be in range (0, <number-of-samples>)")
    if pairs is None:
        pi, pj = sample_pairs(seed, scan, chunk, max_trials, n)
    else:
        pi, pj = pairs
    best = None
    best_n = 0
    best_ssum = np.inf
    best_inl = None
    win = -1
    counts = np.zeros(max_trials, dtype=np.int32)
    ssums = np.full(max_trials, np.nan)
    executed = 0
    for t in range(max_trials):
        o, d = line_from_pair(data[pi[t]], data[pj[t]])
        r = np.abs(line_residuals(data, o, d))
        inl = r < residual_threshold
        ssum = tree32_sum(r * r)
        cnt = int(np.sum(inl))
        counts[t] = cnt
        ssums[t] = ssum
        executed = t + 1
        if cnt > best_n or (cnt == best_n and ssum < best_ssum):
            best, best_n, best_ssum, best_inl, win = (o, d), cnt, ssum, inl, t
            if early_exit and best_ssum <= 0:
                break
    info = {"win_trial": win, "ninl": best_n, "ssum": best_ssum, "trials_run": executed,
            "win_pair": (int(pi[win]), int(pj[win])) if win >= 0 else (-1, -1)}
    if details:
        info.update(counts=counts, ssums=ssums, pairs=(pi, pj), sample_model=best)
    if best_inl is None:
        return None, None, info
    model = line_estimate(data[best_inl])
    return model, best_inl, info


def extract_line(data, landmark_number=0, **kw):
    """RANSAC + slope/intercept/tip exactly as /root/reference/ransac_functions.py:20-31.

    Returns a dict with a, b, origin, direction, tip, inliers, xBase and the ransac info;
    raises AttributeError like the reference when no model is found.
    """
    data = np.asarray(data, dtype=np.float64)
    model, inliers, info = ransac_line(data, **kw)
    if model is None:
        raise AttributeError("'NoneType' object has no attribute 'params'")
    o, d = model
    with np.errstate(divide="ignore", invalid="ignore"):
        a = d[1] / d[0]                      # :26  (no guard for d[0] == 0)
        b = o[1] - a * o[0]                  # :27
    x_base = np.array(data[inliers, 0])      # :28
    tip_x = x_base[-1]                       # :29  x of the LAST inlier in data order
    with np.errstate(invalid="ignore"):
        tip_y = tip_x * a + b                # :30
    return {"a": float(a), "b": float(b), "origin": o, "direction": d,
            "tip": np.array([tip_x, tip_y]), "inliers": inliers, "xBase": x_base,
            "id": landmark_number, "info": info}


def project_inliers(data, inliers, a, b):
    """The points landmark_extraction returns (ransac_functions.py:28,46,49,53,55): ``xBase = data[inliers, 0]``,
    ``yBase = a * xBase + b`` with (a, b) the matched stored landmark's coefficients on a hit (:46) or the new
    line's (:49, :53) - a vertical projection; NumPy rounds the product and the sum separately."""
    xBase = np.asarray(data, dtype=np.float64)[np.asarray(inliers, dtype=bool), 0]
    yBase = a * xBase + b
    return np.stack([xBase, yBase], axis=-1)


def chunk_bounds(npts, chunk):
    """Chunking of one revolution as /root/reference/functions.py:64-76 does it.

    Full chunks of ``chunk`` points; the remainder is kept only if it has MORE than 2
    points (``len(distancesList) > 2``, functions.py:71).  Returns [(start, stop), ...].
    """
    out = []
    s = 0
    while s + chunk <= npts:
        out.append((s, s + chunk))
        s += chunk
    if npts - s > 2:
        out.append((s, npts))
    return out


def canonical_direction(d):
    """Direction up to sign -> the representative with d_x > 0 (or d_x == 0, d_y >= 0)."""
    d = np.asarray(d, dtype=np.float64)
    if d[0] < 0 or (d[0] == 0 and d[1] < 0):
        return -d
    return d
