"""Fused UKF step + all-gather (lslam_ukf_step_gather) under CPU emulation: two 'ranks' run the kernel
source one after the other against NumPy arrays standing in for the peer-mapped gather buffers.  Every
rank's buffer must end up holding every rank's (x, P) and ids, bit-identical to what the plain
lslam_ukf_step leaves in x / P - on the unicast path (one store per peer) and on the multicast path."""
import numpy as np
import pytest

from tests import ukf_util as uu
from tests.emu_util import emu_ukf_step, emu_ukf_step_gather


@pytest.mark.parametrize("L,multicast", [(4, False), (4, True), (8, False), (20, False)])
def test_gather_records_equal_local_results(L, multicast):
    world, F = 2, 37
    ranks = [uu.random_filters(F, L, seed=10 + r, line_records=True) for r in range(world)]
    for d in ranks:
        d["lm_valid"][3] = 0                     # a filter without any valid slot: predict only
        d["P"][5] = -np.eye(3)                   # a filter whose Cholesky fails: status set, state kept
    ids = [np.arange(F * L, dtype=np.int32).reshape(F, L) * (r + 1) - 3 for r in range(world)]
    nbuf = 1 if multicast else world
    state = [np.full((world * F, 12), -7.0) for _ in range(nbuf)]
    idbuf = [np.full((world * F, L), -77, np.int32) for _ in range(nbuf)]
    for r, d in enumerate(ranks):
        want = emu_ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], valid_min=1, zmode=1)
        got = emu_ukf_step_gather(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], ids[r], r * F,
                                  state, idbuf, multicast=multicast, valid_min=1, zmode=1)
        for a, b in zip(want, got):
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8))         # the step itself is unchanged
        assert want[2][5] != 0 and want[2][3] == 0
        rec = np.concatenate([got[0], got[1].reshape(F, 9)], axis=1)
        for b in range(nbuf):
            assert np.array_equal(state[b][r * F:(r + 1) * F].view(np.uint64), rec.view(np.uint64))
            assert np.array_equal(idbuf[b][r * F:(r + 1) * F], ids[r])
    # nothing outside the two ranks' blocks was touched, and all 'ranks' hold the same gathered data
    for b in range(1, nbuf):
        assert np.array_equal(state[b].view(np.uint64), state[0].view(np.uint64)) and np.array_equal(idbuf[b], idbuf[0])


def test_sliced_calls_fill_the_same_records():
    world, F, L = 2, 24, 4
    d = uu.random_filters(F, L, seed=3, line_records=True)
    ids = np.arange(F * L, dtype=np.int32).reshape(F, L)
    whole = [np.zeros((world * F, 12)) for _ in range(world)]
    parts = [np.zeros((world * F, 12)) for _ in range(world)]
    wid = [np.zeros((world * F, L), np.int32) for _ in range(world)]
    pid = [np.zeros((world * F, L), np.int32) for _ in range(world)]
    emu_ukf_step_gather(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], ids, 1 * F, whole, wid)
    for s0, s1 in ((0, 7), (7, 24)):
        sl = slice(s0, s1)
        emu_ukf_step_gather(d["x"][sl], d["P"][sl], d["u"][sl], d["z"][sl], d["lm_pos"][sl], d["lm_valid"][sl],
                            ids[sl], 1 * F + s0, parts, pid)
    for r in range(world):
        assert np.array_equal(whole[r].view(np.uint64), parts[r].view(np.uint64)) and np.array_equal(wid[r], pid[r])
        assert not whole[r][:F].any() and whole[r][F:].any()


def test_bad_arguments():
    from lidar_slam_b200._lib import LslamError
    d = uu.random_filters(4, 4, seed=1, line_records=True)
    buf, ib = [np.zeros((8, 12))], [np.zeros((8, 4), np.int32)]
    with pytest.raises(LslamError):
        emu_ukf_step_gather(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], None, -1, buf, ib)
    with pytest.raises(LslamError):
        emu_ukf_step_gather(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"], None, 0, buf * 9, ib * 9)
