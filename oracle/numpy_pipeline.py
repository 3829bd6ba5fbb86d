"""The whole per-scan step in Python/NumPy, the way the reference would run it (TEST INFRASTRUCTURE / CPU baseline):

  per chunk   landmark_extraction: ransac + slope/intercept/tip (/root/reference/ransac_functions.py:20-31 ->
              oracle/ransac_line.py), Landmark(...) (:31), the association loop (:34-54 -> oracle/assoc.py), the
              projected inlier points (:46-55), the caller's append (:75-78)
  per scan    one UKF predict(u) + update(z, landmarks=matched) with the filterpy engine restated in oracle/ukf.py
              (/root/reference/systemClass.py:14-29, UKFMethods.py:10-71); z of a matched line = range / bearing of
              its inlier centroid (the glue the reference never wrote; lslam_ukf_step zmode 1 has the same)

bench.py times this as ``cpu_baseline.numpy_1core`` / ``numpy_pool``: it is the CPU line
closest to what the reference's own Python costs - per-trial NumPy calls - next to the compiled C port.
"""
import math

import numpy as np

from . import assoc, ransac_line as rl, ukf


def scan_step(xy, landmarks, x, P, u, scan_index=0, seed=0, chunk=100, trials=100, thr=20.0, number0=0):
    """One revolution of one sensor.  xy [N, 2]; landmarks: the stream's Python list (mutated like the reference's).
    Returns (points [n, 2], x, P, lines)."""
    pts, z, lm, lines = [], [], [], []
    number = number0
    for c, (s, e) in enumerate(rl.chunk_bounds(xy.shape[0], chunk)):
        try:
            f = rl.extract_line(xy[s:e], number, residual_threshold=thr, max_trials=trials, seed=seed, scan=scan_index, chunk=c)
        except (AttributeError, ValueError):
            number += 1
            continue
        new = assoc.Landmark(f["a"], f["b"], number, f["origin"][0], f["origin"][1], f["tip"][0], f["tip"][1])
        hit, _ = assoc.associate(landmarks, new)
        a, b = (hit.get_a(), hit.get_b()) if hit is not None else (f["a"], f["b"])
        pts.append(rl.project_inliers(xy[s:e], f["inliers"], a, b))
        if hit is None:
            landmarks.append(new)
        else:
            o = f["origin"]
            z += [math.hypot(o[0], o[1]), math.atan2(o[1], o[0])]
            lm.append(hit.get_pos())
        lines.append(f)
        number += 1
    x, P = ukf.step_literal(x, P, u, z, lm)
    return (np.concatenate(pts, 0) if pts else np.zeros((0, 2))), x, P, lines


def _known_map(xy, scan_index, seed, chunk, trials, thr):
    """The bench's set-up for one stream: its own lines with the tip on the centroid (so the re-observation hits)."""
    lms = []
    for c, (s, e) in enumerate(rl.chunk_bounds(xy.shape[0], chunk)):
        try:
            f = rl.extract_line(xy[s:e], c, residual_threshold=thr, max_trials=trials, seed=seed, scan=scan_index, chunk=c)
        except (AttributeError, ValueError):
            continue
        lms.append(assoc.Landmark(f["a"], f["b"], c, f["origin"][0], f["origin"][1], f["origin"][0], f["origin"][1]))
    return lms


def run_scans(args):
    """Worker (also the multiprocessing.Pool target): (xy [n, N, 2], first scan index) -> number of scans stepped.
    The known map is built first (untimed by the caller only in the sense that both arms do it in set-up; here it is
    returned separately)."""
    xy, first, seed = args
    import time
    maps = [_known_map(xy[k], first + k, seed, 100, 100, 20.0) for k in range(xy.shape[0])]
    t0 = time.perf_counter()
    for k in range(xy.shape[0]):
        scan_step(xy[k], maps[k], np.zeros(3), np.diag([0.1, 0.1, 0.05]), np.ones(2), first + k, seed)
    return xy.shape[0], time.perf_counter() - t0
