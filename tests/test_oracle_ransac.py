"""Oracle self-checks: NumPy restatement vs C restatement, known-answer cases, properties.

The reference has no tests for this path and its arithmetic lives in
scikit-image, which is absent: parity is UNPINNED against upstream.  What is pinned here is
that the two independent restatements agree bit for bit, that they reproduce analytically
known answers, and the committed golden fixture.
"""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from lidar_slam_b200 import synth
from oracle import ransac_line as rl
from tests.parity_util import mask_to_bool

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ransac_c1.npz")


def numpy_batch(xy, chunk=100, trials=100, thr=20.0, seed=0, scan_base=0):
    """oracle/ransac_line.py over a batch, in the output layout of lslam_ransac_lines."""
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    W = (chunk + 31) // 32
    out = {"line": np.full((B, C, 8), np.nan), "ninl": np.zeros((B, C), np.int32),
           "win_trial": np.full((B, C), -1, np.int32), "win_pair": np.full((B, C, 2), -1, np.int32),
           "ssum": np.full((B, C), np.nan), "inl_mask": np.zeros((B, C, W), np.uint32)}
    for b in range(B):
        for c, (s, e) in enumerate(rl.chunk_bounds(N, chunk)):
            r = rl.extract_line(xy[b, s:e], residual_threshold=thr, max_trials=trials, seed=seed,
                                scan=scan_base + b, chunk=c)
            d = rl.canonical_direction(r["direction"])
            out["line"][b, c] = [r["a"], r["b"], r["origin"][0], r["origin"][1], d[0], d[1],
                                 r["tip"][0], r["tip"][1]]
            out["ninl"][b, c] = r["info"]["ninl"]
            out["win_trial"][b, c] = r["info"]["win_trial"]
            out["win_pair"][b, c] = r["info"]["win_pair"]
            out["ssum"][b, c] = r["info"]["ssum"]
            idx = np.nonzero(r["inliers"])[0]
            np.bitwise_or.at(out["inl_mask"][b, c], idx >> 5, (1 << (idx & 31)).astype(np.uint32))
    return out


def test_numpy_vs_c_random_rooms(cport):
    from tests.parity_util import assert_ransac_equal
    xy = synth.random_room_scans(24, seed=11)
    assert_ransac_equal(numpy_batch(xy, seed=5, scan_base=1000),
                        cport.ransac_lines(xy, seed=5, scan_base=1000), "numpy-vs-C")


def test_numpy_vs_c_highres(cport):
    from tests.parity_util import assert_ransac_equal
    xy = synth.polygon_scan(beams=2050, nsides=5)[None]
    assert_ransac_equal(numpy_batch(xy, chunk=410, trials=64, seed=3),
                        cport.ransac_lines(xy, chunk=410, trials=64, seed=3), "numpy-vs-C 410")


def test_golden_c1(cport):
    g = np.load(GOLD)
    xy = g["xy"]
    assert np.array_equal(xy, synth.room_scan()), "scan generator drifted from the fixture"
    res = cport.ransac_lines(xy[None])
    for c, (s, e) in enumerate(g["bounds"]):
        r = rl.extract_line(xy[s:e], seed=0, scan=0, chunk=c)
        assert r["info"]["ninl"] == g["ninl"][c] == res["ninl"][0, c]
        assert r["info"]["win_trial"] == g["win_trial"][c] == res["win_trial"][0, c]
        assert tuple(r["info"]["win_pair"]) == tuple(g["win_pair"][c]) == tuple(res["win_pair"][0, c])
        assert r["info"]["ssum"] == g["ssum"][c] == res["ssum"][0, c]
        assert np.array_equal(r["inliers"], g["mask"][c][:e - s])
        assert np.array_equal(mask_to_bool(res["inl_mask"][0, c], e - s), r["inliers"])
        np.testing.assert_allclose([r["a"], r["b"]], [g["a"][c], g["b"][c]], rtol=1e-12)
        np.testing.assert_allclose(res["line"][0, c, :2], [g["a"][c], g["b"][c]], rtol=1e-9)


def test_four_wall_room_analytic():
    """KAT-1: noise-free room, robot axis-aligned at the centre: chunk k of 90 beams sees
    mostly one wall; the extracted line must be that wall (x = +-W/2 or y = +-H/2)."""
    xy = synth.room_scan(pose=(0.0, 0.0, 0.0), room_w_mm=4000.0, room_h_mm=3000.0, noise_mm=0.0)
    # beams 0..359 clockwise from +y: beam 0 hits the top wall (y = 1500)
    r = rl.extract_line(xy[0:40], seed=1)          # top wall: y = 1500 -> a = 0, b = 1500
    assert abs(r["a"]) < 1e-9 and abs(r["b"] - 1500.0) < 1e-6 and r["info"]["ninl"] == 40
    r = rl.extract_line(xy[160:200], seed=1)       # bottom wall: y = -1500
    assert abs(r["a"]) < 1e-9 and abs(r["b"] + 1500.0) < 1e-6
    r = rl.extract_line(xy[70:110], seed=1)        # right wall x = 2000: vertical, a = +-inf/huge
    d = rl.canonical_direction(r["direction"])
    assert abs(d[0]) < 1e-9 and abs(abs(d[1]) - 1.0) < 1e-12
    assert abs(r["origin"][0] - 2000.0) < 1e-6


def test_collinear_chunk_early_exit():
    """KAT-2: exactly collinear points with exactly representable coordinates: the first
    trial already has all points at distance 0 -> ssum == 0 -> break after one trial."""
    t = np.arange(100, dtype=np.float64)
    data = np.stack([t * 8.0, np.full(100, 64.0)], -1)      # horizontal line y = 64
    model, inl, info = rl.ransac_line(data, seed=9)
    assert info["trials_run"] == 1 and info["win_trial"] == 0 and info["ssum"] == 0.0
    assert inl.all() and info["ninl"] == 100
    # evaluating all trials (what the GPU does) selects the same winner
    _, _, info_all = rl.ransac_line(data, seed=9, early_exit=False)
    assert info_all["trials_run"] == 100 and info_all["win_trial"] == 0


def test_vertical_wall_slope_is_not_guarded():
    """KAT-3: dir_x == 0 -> a = +-inf, b = -+inf or nan exactly like ransac_functions.py:26-27."""
    t = np.arange(50, dtype=np.float64)
    data = np.stack([np.full(50, 1500.0), t * 16.0], -1)
    r = rl.extract_line(data, seed=2)
    assert np.isinf(r["a"]) and not np.isfinite(r["b"])
    assert r["info"]["ninl"] == 50


def test_duplicate_points_zero_norm_model():
    """KAT-4: both samples identical -> norm 0 -> direction stays (0, 0); residual = distance
    to that point; the model is still 'successful' upstream."""
    o, d = rl.line_from_pair([0.0, 0.0], [0.0, 0.0])
    assert tuple(d) == (0.0, 0.0)
    r = rl.line_residuals(np.array([[3.0, 4.0], [0.0, 0.0]]), o, d)
    assert tuple(r) == (5.0, 0.0)
    data = np.zeros((100, 2))
    data[60:] = [1000.0, 0.0]
    model, inl, info = rl.ransac_line(data, seed=0)
    assert info["ninl"] == 100        # a pair with one point in each cluster explains everything


def test_too_few_points_raises():
    with pytest.raises(ValueError):
        rl.ransac_line(np.zeros((1, 2)))


def test_no_model_cases(cport):
    data = np.array([[0.0, 0.0], [10.0, 0.0], [20.0, 5.0]])
    # zero inliers everywhere: upstream still ACCEPTS the first trial (0 == 0 and ssum < inf)
    # and its final estimate() on an empty selection raises ValueError
    with pytest.raises(ValueError):
        rl.ransac_line(data, residual_threshold=0.0)
    res = cport.ransac_lines(data[None], chunk=3, trials=10, thr=0.0)
    assert res["ninl"][0, 0] == -3 and res["win_trial"][0, 0] == -1
    # never accepted (NaN sums): (None, None) -> the reference raises AttributeError at :25
    bad = np.full((3, 2), np.nan)
    model, inl, info = rl.ransac_line(bad)
    assert model is None and inl is None and info["win_trial"] == -1
    with pytest.raises(AttributeError):
        rl.extract_line(bad)


def test_chunk_bounds_follow_scanning():
    assert rl.chunk_bounds(360, 100) == [(0, 100), (100, 200), (200, 300), (300, 360)]
    assert rl.chunk_bounds(302, 100) == [(0, 100), (100, 200), (200, 300)]      # remainder 2 dropped
    assert rl.chunk_bounds(303, 100) == [(0, 100), (100, 200), (200, 300), (300, 303)]
    assert rl.chunk_bounds(8192, 410)[-1] == (7790, 8192) and len(rl.chunk_bounds(8192, 410)) == 20
    assert synth.chunk_revolution(np.zeros((360, 2)))[-1] == 0
    assert [len(c) for c in synth.chunk_revolution(np.zeros((360, 2)))[:-1]] == [100, 100, 100, 60]


@settings(max_examples=30, deadline=None)
@given(st.lists(st.floats(0.0, 1e6, allow_nan=False), min_size=0, max_size=200))
def test_tree32_sum_matches_reference_tree(vals):
    v = np.array(vals, dtype=np.float64)
    lanes = [0.0] * 32
    for k, x in enumerate(v):
        lanes[k % 32] = lanes[k % 32] + float(x)
    h = 16
    while h >= 1:
        for l in range(h):
            lanes[l] = lanes[l] + lanes[l + h]
        h //= 2
    assert rl.tree32_sum(v) == lanes[0]


@settings(max_examples=10, deadline=None)
@given(st.integers(0, 2**31), st.floats(-3.0, 3.0))
def test_rotating_the_scene_rotates_the_line(seed, phi):
    """Property: with the same sample pairs, a rigid rotation of the
    points rotates the fitted direction and keeps the inlier set (up to rounding at the
    threshold, avoided by a noise level far below it)."""
    rng = np.random.default_rng(seed)
    t = rng.uniform(-1000, 1000, 80)
    pts = np.stack([t, 0.3 * t + 50 + rng.normal(0, 1.0, 80)], -1)
    pts[::9] += rng.uniform(200, 400, (len(pts[::9]), 2))       # gross outliers
    c, s = np.cos(phi), np.sin(phi)
    rot = pts @ np.array([[c, s], [-s, c]])
    pairs = rl.sample_pairs(seed, 0, 0, 100, 80)
    m0, i0, info0 = rl.ransac_line(pts, pairs=pairs)
    m1, i1, info1 = rl.ransac_line(rot, pairs=pairs)
    assert np.array_equal(i0, i1)
    d0 = rl.canonical_direction(m0[1])
    d1 = m1[1]
    want = np.array([c * d0[0] - s * d0[1], s * d0[0] + c * d0[1]])
    assert min(np.abs(d1 - want).max(), np.abs(d1 + want).max()) < 1e-9


@settings(max_examples=10, deadline=None)
@given(st.integers(0, 2**31))
def test_trial_order_permutation_keeps_lexicographic_winner(seed):
    """Property: the sequential accept rule selects the lexicographic optimum
    (count desc, ssum asc, first occurrence), so permuting DISTINCT hypotheses changes at most
    which of several identical-scoring trials is reported, never its score."""
    rng = np.random.default_rng(seed)
    t = rng.uniform(-500, 500, 60)
    pts = np.stack([t, -0.7 * t + rng.normal(0, 4.0, 60)], -1)
    pi, pj = rl.sample_pairs(seed, 1, 2, 50, 60)
    perm = rng.permutation(50)
    _, _, a = rl.ransac_line(pts, max_trials=50, pairs=(pi, pj), details=True)
    _, _, b = rl.ransac_line(pts, max_trials=50, pairs=(pi[perm], pj[perm]), details=True)
    assert a["ninl"] == b["ninl"] and a["ssum"] == b["ssum"]
    key = sorted(zip(-a["counts"], a["ssums"], range(50)))[0]
    assert a["win_trial"] == key[2]
