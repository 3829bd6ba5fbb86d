"""The drop-in modules (lidar_slam_b200/{ransac_functions,landmarking,systemClass}.py) used the way
SLAM.py / ransac_functions.py use the reference's, checked against the oracle: same return values,
same mutation of the caller's lists, same landmark life cycle."""
import os
import threading
import time

import numpy as np
import pytest

from lidar_slam_b200 import synth
from oracle import assoc as oassoc, ransac_line as oransac, ukf as oukf

pytestmark = pytest.mark.gpu


def as_xy(points):
    if isinstance(points, np.ndarray):
        return points
    return np.array([[p.x(), p.y()] for p in points])


def oracle_extraction(chunk, number, landmarks, seed=0):
    """ransac_functions.py:15-59 on the oracle's pieces."""
    e = oransac.extract_line(np.array(chunk), number, seed=seed, scan=number, chunk=0)
    new = oassoc.Landmark(e["a"], e["b"], number, e["origin"][0], e["origin"][1], e["tip"][0], e["tip"][1])
    hit, _ = oassoc.associate(landmarks, new)
    a, b = (hit.a, hit.b) if hit is not None else (e["a"], e["b"])
    return np.stack([e["xBase"], a * e["xBase"] + b], -1), new, hit is None


def test_landmark_extraction_sequence(gpu_engine):
    from lidar_slam_b200 import ransac_functions as rf
    rng = np.random.default_rng(0)
    # three revolutions of a robot creeping along a corridor: re-observations, misses and expiries
    revs = [synth.room_scan(pose=(300.0 + 40 * k, -200.0, np.deg2rad(10.0)), rng=rng) for k in range(3)]
    mine, theirs = [], []
    number = 0
    old_life = rf.LIFE
    for xy in revs:
        for chunk in synth.chunk_revolution(xy)[:-1]:
            box = [chunk]
            pts, fitted, is_new = rf.landmark_extraction(box, number, mine)
            assert box == []                                           # the caller's list is emptied (:22)
            wpts, wfit, wnew = oracle_extraction(chunk, number, theirs)
            assert is_new == wnew
            np.testing.assert_allclose(as_xy(pts), wpts, rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose([fitted.a, fitted.b, *fitted.pos, *fitted.end],
                                       [wfit.a, wfit.b, *wfit.pos, *wfit.end], rtol=1e-5, atol=1e-6)
            assert fitted.id == number and fitted.life == old_life
            if is_new:
                mine.append(fitted)
                theirs.append(wfit)
            number += 1
            assert [(l.id, l.life) for l in mine] == [(l.id, l.life) for l in theirs]
    assert len(mine) >= 4 and number == 12


def test_expiry_and_skip_semantics_on_the_python_list(gpu_engine):
    from lidar_slam_b200 import ransac_functions as rf
    from lidar_slam_b200.landmarking import Landmark
    t = np.linspace(0, 990, 100)
    chunk = np.stack([t, 0.0 * t + 100.0], -1).tolist()              # the line y = 100
    def stored():
        A = Landmark(5.0, 0.0, 0, 0, 0, 10, 50); A.life = 1          # expires on this query
        B = Landmark(0.0, 100.0, 1, 450.0, 100.0, 495.0, 100.0)      # would match, but is skipped (mode 0)
        C = Landmark(-5.0, 0.0, 2, 0, 0, 10, -50)
        return [A, B, C]
    lms = stored()
    pts, fitted, is_new = rf.landmark_extraction([chunk], 3, lms)
    ref = [oassoc.Landmark(l.a, l.b, l.id, l.pos[0], l.pos[1], l.end[0], l.end[1]) for l in stored()]
    ref[0].life = 1
    hit, _ = oassoc.associate(ref, oassoc.Landmark(fitted.a, fitted.b, 3, *fitted.pos, *fitted.end))
    assert is_new == (hit is None)
    assert [(l.id, l.life) for l in lms] == [(l.id, l.life) for l in ref]
    assert [l.id for l in lms] == [1, 2] and lms[0].life == 40 and lms[1].life == 39 and is_new


def test_no_model_and_bad_input(gpu_engine):
    from lidar_slam_b200 import ransac_functions as rf
    with pytest.raises(ValueError):
        rf.landmark_extraction([[[0.0, 0.0]]], 0, [])
    far = [[0.0, 0.0], [1e5, 0.0], [0.0, 1e5], [1e5, 1e5], [5e4, 7e4]]
    got = rf.landmark_extraction([far], 0, [])                          # every trial has its 2 sample inliers
    assert len(as_xy(got[0])) >= 2
    with pytest.raises(ValueError):
        rf.landmark_extraction([[[0.0, 0.0], [1.0, np.nan], [2.0, 2.0]]], 0, [])


def test_check_ransac_thread_and_adapter(gpu_engine):
    from lidar_slam_b200 import ransac_functions as rf
    xy = synth.room_scan()
    items = synth.chunk_revolution(xy)
    pair, temp, allp, todo, lms = [], [], [], [], []
    ev, stop = threading.Event(), threading.Event()
    th = threading.Thread(target=rf.check_ransac, args=(pair, temp, allp, todo, lms, ev, stop), daemon=True)
    th.start()
    for it in items:                       # feed one item at a time (the reference drops chunks that race)
        todo.append(it)
        t0 = time.time()
        while todo and time.time() - t0 < 30:
            time.sleep(0.001)
    assert ev.wait(30)
    stop.set()
    th.join(10)
    assert len(pair) == 1 and len(lms) == 4
    pts_thread = as_xy(pair[0])
    ad = rf.RevolutionAdapter(capacity=16)
    out = None
    for it in items:
        out = ad.put(it)
    assert out is not None and list(out["match"]) == [-1, -1, -1, -1]
    # the per-chunk path seeds scan = landmarkNumber, chunk 0; the adapter batches the revolution as
    # scans 0..3 of a [4, 100] batch - the same Philox streams, so the same lines and points
    np.testing.assert_allclose(out["points"], pts_thread, rtol=1e-9, atol=1e-9)


def test_system_filter_against_literal_oracle(gpu_engine):
    from lidar_slam_b200 import systemClass
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ukf_kat7.npz"))
    lms = [oukf._Pos(*p) for p in g["lm"]]
    s = systemClass.System(lms)
    assert s.dt == 0.005 and s.sigmas.Wm[0] == pytest.approx(1 - 1e8, rel=1e-6)
    np.testing.assert_array_equal(s.ukf.x, [0, 0, 0])
    np.testing.assert_array_equal(s.ukf.P, np.diag([.1, .1, .05]))
    for k in range(5):
        s.ukf.predict(u=g["u"])
        s.ukf.update(g["z"][k], landmarks=lms)
        np.testing.assert_allclose(s.ukf.x, g["x"][k], rtol=1e-5, atol=1e-4)
        assert np.abs(s.ukf.P - g["P"][k]).max() <= 1e-5 * np.abs(g["P"][k]).max()
    with pytest.raises(ValueError):
        s.ukf.R = np.ones((16, 16))
