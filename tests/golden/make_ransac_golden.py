"""Generates tests/golden/ransac_c1.npz: the C1 scan (BASELINE.json configs[0]: 360 beams,
4 m x 3 m room, robot at (0.3 m, -0.2 m, 10 deg), N(0, 5 mm) noise, default_rng(0)) and what
the NumPy oracle (oracle/ransac_line.py) extracts from it with the reference's constants
(THRESHOLD=20, MAX_TRIALS=100, chunk=100), seed 0.  The reference cannot produce these
itself (skimage absent, unseeded sampler); the fixture pins the oracle + sampler contract so
that any later change to either shows up as a diff.  Run from the repo root:
    python tests/golden/make_ransac_golden.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lidar_slam_b200.synth import room_scan          # noqa: E402
from oracle import ransac_line as rl                  # noqa: E402

xy = room_scan()
bounds = rl.chunk_bounds(len(xy), 100)
out = {k: [] for k in ("a", "b", "origin", "direction", "tip", "ninl", "win_trial", "win_pair", "ssum", "mask")}
for c, (s, e) in enumerate(bounds):
    r = rl.extract_line(xy[s:e], seed=0, scan=0, chunk=c)
    out["a"].append(r["a"]); out["b"].append(r["b"]); out["origin"].append(r["origin"])
    out["direction"].append(rl.canonical_direction(r["direction"])); out["tip"].append(r["tip"])
    out["ninl"].append(r["info"]["ninl"]); out["win_trial"].append(r["info"]["win_trial"])
    out["win_pair"].append(r["info"]["win_pair"]); out["ssum"].append(r["info"]["ssum"])
    m = np.zeros(100, bool); m[:e - s] = r["inliers"]; out["mask"].append(m)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ransac_c1.npz")
np.savez_compressed(path, xy=xy, bounds=np.array(bounds), **{k: np.array(v) for k, v in out.items()})
print("wrote", path)
