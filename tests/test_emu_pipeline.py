"""The three kernels chained the way lidar_slam_b200/pipeline.py chains them (line records ->
association with hit_pos -> UKF with zmode 1 / match validity), run through the emulated library and
compared with the C pipeline (oracle/cport.pipeline_step) over several steps."""
import numpy as np

from lidar_slam_b200 import synth
from tests import pipeline_util as pu
from tests.parity_util import assert_ransac_equal


def test_three_steps_with_a_known_map(emu):
    S, M = 6, 8
    xy = synth.random_room_scans(S, seed=4)
    u = np.full((S, 2), 1.0)
    want = pu.cpu_pipeline(xy, 3, M=M, u=u)
    r = emu.emu_ransac_lines(xy)
    assert_ransac_equal(want[0]["ransac"], r, "pipeline RANSAC")
    r = want[0]["ransac"]        # fitted parameters agree to 1e-5, not bitwise: chain both sides from the same lines
    tables = pu.map_from_lines(r["line"], r["ninl"], M)
    x = np.zeros((S, 3))
    P = np.tile(np.diag([0.1, 0.1, 0.05]), (S, 1, 1))
    for k in range(3):
        hit_pos = np.full((S, 4, 2), -7.0)
        m, mid, proj = emu.emu_associate(r["line"], tables, q_valid=r["ninl"], layout=1, hit_pos=hit_pos)
        assert np.array_equal(m, want[k]["assoc"]["match"]) and np.array_equal(mid, want[k]["assoc"]["match_id"])
        assert np.array_equal(np.nan_to_num(hit_pos), np.nan_to_num(want[k]["assoc"]["hit_pos"]))
        for key in ("rows", "life", "ids", "count", "next_id"):
            assert np.array_equal(np.nan_to_num(tables[key]), np.nan_to_num(want[k]["tables"][key])), key
        x, P, st = emu.emu_ukf_step(x, P, u, r["line"], hit_pos, m, valid_min=0, zmode=1)
        assert np.array_equal(x.view(np.uint64), want[k]["x"].view(np.uint64))
        assert np.array_equal(P.view(np.uint64), want[k]["P"].view(np.uint64))
        assert (st == 0).all() and (m >= 0).any()


def test_empty_map_builds_landmarks_then_fills_up(emu):
    S, M = 3, 8
    xy = synth.random_room_scans(S, seed=5)
    want = pu.cpu_pipeline(xy, 3, M=M, with_map=False)
    r = want[0]["ransac"]
    tables = pu.cport.empty_tables(S, M)
    for k in range(3):
        hit_pos = np.zeros((S, 4, 2))
        m, mid, _ = emu.emu_associate(r["line"], tables, q_valid=r["ninl"], layout=1, hit_pos=hit_pos)
        assert np.array_equal(m, want[k]["assoc"]["match"])
    assert (want[2]["assoc"]["match"] == -3).any()          # a static sensor never re-matches its own walls
