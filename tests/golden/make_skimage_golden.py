"""Generates tests/golden/skimage_golden.npz from the REAL scikit-image - run it wherever scikit-image is installed
(it is not in this project's image, which is why parity of the RANSAC restatement is "unpinned"):

    python tests/golden/make_skimage_golden.py          # writes tests/golden/skimage_golden.npz

It runs the reference's own call (/root/reference/ransac_functions.py:23-24)

    model, inliers = skimage.measure.ransac(data, LineModelND, min_samples=2, residual_threshold=20,
                                            max_trials=100, random_state=<recording generator seeded with s>)

on a set of chunks (four-wall room scans cut like functions.py:64-76, a corner chunk, a short remainder, gross
outliers, a vertical wall, duplicate points) and records, per case: the data, the index pairs the generator handed
to ransac (in draw order; upstream stops early when the sum of squared residuals reaches 0, so fewer than
max_trials may be drawn), the returned inlier mask and model parameters.  tests/test_pairs_replay.py then holds
oracle/ransac_line.py and the CUDA kernels to these results through the explicit-pairs path.

The generator is a subclass of the class upstream expects (np.random.RandomState for scikit-image <= 0.18,
np.random.Generator from 0.19 on) whose choice() records what it returns - no patching of upstream code.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from lidar_slam_b200 import synth  # noqa: E402

THR, TRIALS, NMAX = 20.0, 100, 100


class RecordingRandomState(np.random.RandomState):
    def __init__(self, seed):
        super().__init__(seed)
        self.log = []

    def choice(self, *a, **k):
        r = super().choice(*a, **k)
        self.log.append(np.array(r, dtype=np.int64).copy())
        return r


class RecordingGenerator(np.random.Generator):
    def __init__(self, seed):
        super().__init__(np.random.PCG64(seed))
        self.log = []

    def choice(self, *a, **k):
        r = super().choice(*a, **k)
        self.log.append(np.array(r, dtype=np.int64).copy())
        return r


def cases():
    out = []
    room = synth.room_scan()                                   # BASELINE.json configs[0]: 360 beams, 4 walls
    for s, e in ((0, 100), (100, 200), (200, 300), (300, 360)):
        out.append(room[s:e])
    for k, xy in enumerate(synth.random_room_scans(6, seed=31)):
        out.append(xy[100 * (k % 3):100 * (k % 3) + 100])
    rng = np.random.default_rng(5)
    t = np.linspace(-900, 900, 100)
    wall = np.column_stack([t, 0.002 * t + 1500 + rng.normal(0, 5, 100)])
    gross = wall.copy()
    gross[::9, 1] += rng.uniform(200, 900, gross[::9].shape[0])
    out.append(gross)
    out.append(np.column_stack([np.full(80, 2000.0) + rng.normal(0, 4, 80), np.linspace(-700, 800, 80)]))   # vertical
    dup = wall[:60].copy()
    dup[10:20] = 0.0                                           # dropouts: duplicate (0, 0) points
    out.append(dup)
    out.append(np.column_stack([t[:7], 3.0 * t[:7] + 40.0]))   # exact line: upstream leaves at trial 0 (sum r^2 == 0)
    return out


def main():
    import skimage
    from skimage.measure import LineModelND, ransac
    legacy = tuple(int(v) for v in skimage.__version__.split(".")[:2]) < (0, 19)
    cs = cases()
    K = len(cs)
    data = np.zeros((K, NMAX, 2))
    n = np.zeros(K, np.int32)
    pairs = np.zeros((K, TRIALS, 2), np.int32)
    drawn = np.zeros(K, np.int32)
    inliers = np.zeros((K, NMAX), bool)
    origin = np.zeros((K, 2))
    direction = np.zeros((K, 2))
    for k, d in enumerate(cs):
        rs = RecordingRandomState(100 + k) if legacy else RecordingGenerator(100 + k)
        kw = {"random_state": rs} if legacy or "random_state" in ransac.__code__.co_varnames else {"rng": rs}
        model, inl = ransac(d, LineModelND, min_samples=2, residual_threshold=THR, max_trials=TRIALS, **kw)
        log = [p for p in rs.log if p.shape == (2,)]
        n[k] = d.shape[0]
        data[k, :n[k]] = d
        drawn[k] = len(log)
        pairs[k, :len(log)] = np.array(log)
        inliers[k, :n[k]] = inl
        origin[k], direction[k] = model.params
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "skimage_golden.npz")
    np.savez(out, data=data, n=n, pairs=pairs, trials_drawn=drawn, inliers=inliers, origin=origin, direction=direction,
             thr=THR, skimage_version=skimage.__version__, numpy_version=np.__version__, legacy=legacy)
    print("wrote", out, "from scikit-image", skimage.__version__)


if __name__ == "__main__":
    main()
