"""oracle/assoc.py against golden vectors produced by the REAL /root/reference/landmarking.py
(tests/golden/make_landmarking_golden.py): this part of the oracle is PINNED to the reference."""
import os

import numpy as np
import pytest

from oracle import assoc

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "landmarking_golden.npz"))


def mk(row, ident=0):
    return assoc.Landmark(row[0], row[1], ident, row[2], row[3], row[4], row[5])


def test_constants():
    assert tuple(G["consts"]) == (assoc.LIFE, assoc.TOLERANCE_A, assoc.TOLERANCE_B, assoc.TOLERANCE)


def test_is_equal_matches_reference():
    got = np.array([mk(s).is_equal(mk(n)) for s, n in zip(G["stored"], G["new"])])
    assert np.array_equal(got, G["equal"])
    assert 0.05 < got.mean() < 0.5


def test_life_bookkeeping_matches_reference():
    lm = assoc.Landmark(0, 0, 0, 0, 0, 1, 1)
    for want_ret, want_life in zip(G["ret_trace"], G["life_trace"]):
        r = lm.decrease_life()
        assert (-1 if r is None else int(bool(r))) == want_ret
        assert lm.get_life() == want_life
    lm.reset_life()
    assert lm.get_life() == int(G["life_after_reset"]) == assoc.LIFE


def replay(stream_q, reference_exact=True):
    landmarks, number, out = [], 0, []
    for row in stream_q:
        f = mk(row, number)
        hit, idx = assoc.associate(landmarks, f, reference_exact)
        if hit is None:
            landmarks.append(f)
            f.life = int(row[6])
        number += 1
        out.append((idx, len(landmarks), [l.life for l in landmarks]))
    return out


def test_association_loop_matches_reference_replay():
    removed = 0
    for s in range(G["stream_q"].shape[0]):
        out = replay(G["stream_q"][s])
        for step, (idx, cnt, lives) in enumerate(out):
            assert idx == G["stream_match"][s, step]
            assert cnt == G["stream_count"][s, step]
            assert lives == list(G["stream_lives"][s, step, :cnt])
        counts = G["stream_count"][s]
        removed += int((np.diff(counts) < 0).sum())
    assert removed > 20            # the fixture really exercises expiry


def test_skip_bug_is_observable():
    """KAT-5: stored [A(life 1), B, C]; a query equal only to B.  A expires and is removed;
    the reference then skips B (it slid into A's slot) and tests C -> miss.  Without the bug
    B is hit."""
    def table():
        A = assoc.Landmark(5.0, 0.0, 0, 0, 0, 10, 50)
        A.life = 1
        B = assoc.Landmark(0.0, 100.0, 1, 0, 100, 50, 100)
        C = assoc.Landmark(-5.0, 0.0, 2, 0, 0, 10, -50)
        return [A, B, C]
    q = assoc.Landmark(0.01, 101.0, 9, 60, 100, 120, 100)
    lms = table()
    hit, idx = assoc.associate(lms, q, reference_exact=True)
    assert hit is None and idx == -1 and [l.id for l in lms] == [1, 2]
    assert lms[0].life == assoc.LIFE and lms[1].life == assoc.LIFE - 1     # B untouched, C aged
    lms = table()
    hit, idx = assoc.associate(lms, q, reference_exact=False)
    assert hit is not None and hit.id == 1 and idx == 0


def test_table_roundtrip_and_stream():
    t = assoc.LandmarkTable(8)
    q = np.array([[0.0, 100.0, 0, 100, 50, 100], [0.01, 101.0, 60, 100, 120, 100], [3.0, 0.0, 0, 0, 10, 30]])
    match, mid, proj = assoc.associate_stream(t, q)
    assert list(match) == [-1, 0, -1] and list(mid) == [0, 0, 2] and t.count == 2 and t.next_id == 3
    assert tuple(proj[1]) == (0.0, 100.0)        # a hit projects with the STORED landmark's a, b (:46)
    assert list(t.life[:2]) == [assoc.LIFE - 1, assoc.LIFE]


@pytest.mark.parametrize("mode", [0, 1])
def test_c_twin_matches_python(cport, mode):
    """oracle/assoc.c (the fast CPU baseline) against oracle/assoc.py (pinned to the real module)."""
    from tests import assoc_util as au
    for T, M, pre, Q in [(50, 8, 0, 12), (50, 8, 8, 6), (20, 40, 37, 10), (10, 70, 66, 5)]:
        queries, valid, tables = au.random_streams(T, Q, M, seed=M + mode, prefill=pre)
        want = au.oracle_associate(queries, valid, tables, mode=mode)
        t2 = {k: v.copy() for k, v in tables.items()}
        o = cport.associate(queries, t2, q_valid=valid, mode=mode)
        au.assert_assoc_equal(want, (o["match"], o["match_id"], o["proj_ab"], t2), "C twin")
        hit = o["match"] >= 0
        assert np.isnan(o["hit_pos"][~hit]).all() and not np.isnan(o["hit_pos"][hit]).any()


def test_dropin_landmark_gap_has_the_pinned_rounding():
    """Landmark.is_equal on the host (lidar_slam_b200/landmarking.py) and lslam_associate on the device must agree on a
    landmark sitting exactly on the 100 mm gate: both use sqrt(fma(vy, vy, vx * vx)) (oracle/assoc.py::norm2, pinned to
    the reference's np.linalg.norm by tests/golden/landmarking_golden.npz)."""
    from lidar_slam_b200 import landmarking as dl
    from oracle import assoc as oa
    rng = np.random.default_rng(3)
    p = rng.normal(0, 3000, (2000, 2))
    q = p + rng.normal(0, 80, (2000, 2))
    for a, b in zip(p, q):
        assert dl.Landmark._gap(a, b) == oa.norm2(a[0] - b[0], a[1] - b[1])
    assert dl.Landmark._gap((0.0, 0.0), (60.0, 80.0)) == 100.0
