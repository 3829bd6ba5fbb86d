"""The association kernel SOURCE (csrc/assoc.cuh) run on the CPU through tests/emu against
oracle/assoc.py (itself pinned to the real landmarking.py): match indices, ids, projection
coefficients and the whole updated table must be identical."""
import numpy as np
import pytest

from tests import assoc_util as au


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("M,prefill,Q", [(8, 0, 12), (8, 8, 6), (5, 4, 9), (16, 13, 9), (12, 12, 7), (40, 37, 10), (70, 66, 5)])
def test_first_match_modes(emu, mode, M, prefill, Q):
    queries, valid, tables = au.random_streams(6, Q, M, seed=M + mode, prefill=prefill)
    want = au.oracle_associate(queries, valid, tables, mode=mode)
    got_t = {k: v.copy() for k, v in tables.items()}
    m, i, p = emu.emu_associate(queries, got_t, q_valid=valid, mode=mode)
    au.assert_assoc_equal(want, (m, i, p, got_t), "mode %d M %d" % (mode, M))
    assert (want[0] >= 0).any() and ((want[0] == -1).any() or prefill > 30)


def test_skip_bug_changes_the_outcome(emu):
    """KAT-5 on the kernel: A(life 1), B, C stored; a query equal only to B.  Mode 0 removes A
    and skips B (miss, appended); mode 1 hits B."""
    rows = np.zeros((1, 8, 6))
    rows[0, 0] = [5.0, 0.0, 0, 0, 10, 50]
    rows[0, 1] = [0.0, 100.0, 0, 100, 50, 100]
    rows[0, 2] = [-5.0, 0.0, 0, 0, 10, -50]
    q = np.array([[[0.01, 101.0, 60, 100, 120, 100]]])
    for mode, want_match, want_count in ((0, -1, 3), (1, 0, 2)):
        t = {"rows": rows.copy(), "life": np.array([[1, 40, 40, 0, 0, 0, 0, 0]], np.int32),
             "ids": np.array([[0, 1, 2, 0, 0, 0, 0, 0]], np.int32), "count": np.array([3], np.int32),
             "next_id": np.array([3], np.int32)}
        m, i, p = emu.emu_associate(q, t, mode=mode)
        assert m[0, 0] == want_match and t["count"][0] == want_count
        if mode == 0:
            assert list(t["ids"][0, :3]) == [1, 2, 3] and list(t["life"][0, :3]) == [40, 39, 40]
        else:
            assert i[0, 0] == 1 and tuple(p[0, 0]) == (0.0, 100.0)


def test_ransac_line_layout_and_full_table(emu):
    """layout 1 reads a, b, origin, tip straight out of lslam_ransac_lines' line records;
    a full table reports -3 and drops the new landmark."""
    rng = np.random.default_rng(3)
    line = rng.normal(size=(2, 5, 8)) * 1000
    q6 = line[..., [0, 1, 2, 3, 6, 7]]
    t6 = {"rows": np.zeros((2, 3, 6)), "life": np.zeros((2, 3), np.int32), "ids": np.zeros((2, 3), np.int32),
          "count": np.zeros(2, np.int32), "next_id": np.zeros(2, np.int32)}
    t8 = {k: v.copy() for k, v in t6.items()}
    a = emu.emu_associate(q6, t6)
    b = emu.emu_associate(line, t8, layout=1)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    assert list(a[0][0]) == [-1, -1, -1, -3, -3] and t6["count"][0] == 3
    want = au.oracle_associate(q6, None, {"rows": np.zeros((2, 3, 6)), "life": np.zeros((2, 3), np.int32),
                                         "ids": np.zeros((2, 3), np.int32), "count": np.zeros(2, np.int32),
                                         "next_id": np.zeros(2, np.int32)})
    au.assert_assoc_equal(want, (a[0], a[1], a[2], t6), "full table")


def test_nearest_mode_picks_the_closest_gated_landmark(emu):
    rows = np.zeros((1, 4, 6))
    rows[0, 0] = [0.0, 100.0, 0, 100, 90, 100]        # gate distance 40 to the query
    rows[0, 1] = [0.0, 101.0, 0, 101, 125, 101]       # gate distance 5 -> nearest
    rows[0, 2] = [2.0, 0.0, 0, 0, 10, 20]             # fails the slope gate
    t = {"rows": rows, "life": np.array([[40, 40, 1, 0]], np.int32), "ids": np.array([[7, 8, 9, 0]], np.int32),
         "count": np.array([3], np.int32), "next_id": np.array([10], np.int32)}
    q = np.array([[[0.0, 100.5, 130, 100.5, 200, 100.5]]])
    m, i, p = emu.emu_associate(q, t, mode=2)
    assert i[0, 0] == 8 and m[0, 0] == 1 and t["count"][0] == 2          # landmark 9 expired and was removed
    assert list(t["life"][0, :2]) == [39, 40]
    t0 = {"rows": rows.copy(), "life": np.array([[40, 40, 1, 0]], np.int32), "ids": np.array([[7, 8, 9, 0]], np.int32),
          "count": np.array([3], np.int32), "next_id": np.array([10], np.int32)}
    m0, i0, _ = emu.emu_associate(q, t0, mode=0)
    assert i0[0, 0] == 7                                                   # first match, not nearest


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_gate_decided_by_the_last_bit(emu, mode):
    """The first-match kernels compare the SQUARED gate sums with the largest double whose rounded root passes the tolerance
    (assoc_gate_sq) instead of taking the two square roots: distances within a few ulps of the tolerance, and exactly on it."""
    for tol in (100.0, 37.3):
        queries, valid, tables = au.gate_boundary_streams(384, seed=int(tol), tol=tol)
        if mode == 2:
            m, i, p = emu.emu_associate(queries, {k: v.copy() for k, v in tables.items()}, mode=2, tol=tol)
            m1, i1, p1 = emu.emu_associate(queries, {k: v.copy() for k, v in tables.items()}, mode=1, tol=tol)
            assert np.array_equal(m, m1)                    # one stored entry: nearest and first match agree
            continue
        if tol != 100.0:                                    # the oracle's tolerance is the reference's constant
            continue
        want = au.oracle_associate(queries, valid, tables, mode=mode)
        got_t = {k: v.copy() for k, v in tables.items()}
        m, i, p = emu.emu_associate(queries, got_t, mode=mode, tol=tol)
        au.assert_assoc_equal(want, (m, i, p, got_t), "gate mode %d" % mode)
        assert (want[0] >= 0).sum() > 50 and (want[0] < 0).sum() > 50 and (want[0][::7] >= 0).all()


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("M,prefill,Q,what", [(8, 6, 7, "misaligned"), (12, 9, 5, "misaligned"), (8, 6, 400, "unstaged"),
                                              (8, 6, 400, "unstaged-misaligned")])
def test_load_paths(emu, mode, M, prefill, Q, what):
    """The three ways the kernel brings a table and its queries in give the same answers: 16-byte asynchronous copies
    (every other test), plain loads because a buffer is not 16-byte aligned, and queries too many to be staged in shared
    memory with the table (read from global memory inside the walk)."""
    queries, valid, tables = au.random_streams(5, Q, M, seed=3 * M + mode, prefill=prefill)
    want = au.oracle_associate(queries, valid, tables, mode=mode)
    got_t = {k: v.copy() for k, v in tables.items()}
    if "misaligned" in what:
        queries = au.misaligned_copy(queries)
        for k in ("rows", "life", "ids"):
            got_t[k] = au.misaligned_copy(got_t[k])
    m, i, p = emu.emu_associate(queries, got_t, q_valid=valid, mode=mode)
    au.assert_assoc_equal(want, (m, i, p, got_t), "%s mode %d" % (what, mode))
