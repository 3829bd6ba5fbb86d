"""The augmented-state UKF kernels (lidar_slam_b200/csrc/ukf_aug.cuh behind lslam_ukf_aug_step), compiled
for the CPU with tests/emu/cuda_emu.h, against the literal filterpy restatement
(oracle/ukf.py step_literal_augmented).  Small state sizes: the emulator runs one OS thread per CUDA
thread.  The -m gpu twin (tests/test_gpu_ukf_aug.py) runs L up to 200."""
import numpy as np
import pytest

from tests import emu_util as eu, ukf_aug_util as au
from lidar_slam_b200 import batch


@pytest.mark.parametrize("L,F,flags", [(4, 3, 7), (4, 2, 3), (3, 2, 1), (3, 2, 2), (18, 1, 7), (48, 1, 7)],
                         ids=["L4", "L4-noredraw", "L3-predict", "L3-update", "L18-two-panels", "L48-skips-zero-k-tiles"])
def test_aug_step_matches_literal(L, F, flags):
    x, P, u, z = au.aug_case(F, L, seed=10 + L)
    want = au.oracle_aug(x, P, u, z, flags)
    cond = au.oracle_conditioning(x, P, u, z, want, flags)
    xg, Pg, st = eu.emu_ukf_aug_step(x, P, u, z, flags=flags)
    assert not st.any()
    au.assert_aug_close(want, (xg, Pg), "emu L=%d flags=%d" % (L, flags), cond=cond)
    assert np.abs(want[1] - P).max() > 1e-3          # the step did something


def test_aug_strict_tolerance_with_wider_sigma_spread():
    """alpha = 0.05 instead of the reference's 1e-4: the central weight is ~ -400, the filter is well
    conditioned and the plain 1e-5 bound holds without the conditioning term."""
    L, F = 4, 2
    x, P, u, z = au.aug_case(F, L, seed=3)
    params = batch.Engine.ukf_params(alpha=0.05)
    want = au.oracle_aug(x, P, u, z, 7, alpha=0.05)
    xg, Pg, st = eu.emu_ukf_aug_step(x, P, u, z, params=params)
    assert not st.any()
    ex, eP = au.assert_aug_close(want, (xg, Pg), "alpha 0.05")
    assert ex < 1e-8 and eP < 1e-8


def test_aug_batching_and_failed_filters_keep_state():
    """Workspace for one filter at a time gives the same bits as all at once; a filter whose P is not
    positive definite reports LSLAM_UKF_ST_SIGMA and keeps (x, P), its neighbours are unaffected."""
    L, F = 3, 3
    x, P, u, z = au.aug_case(F, L, seed=5)
    P[1, 4, 4] = -1.0
    a = eu.emu_ukf_aug_step(x, P, u, z)
    b = eu.emu_ukf_aug_step(x, P, u, z, workspace_filters=1)
    for p, q in zip(a, b):
        assert np.array_equal(p, q)
    assert a[2].tolist() == [0, 1, 0]
    assert np.array_equal(a[0][1], x[1]) and np.array_equal(a[1][1], P[1])
    with pytest.raises(np.linalg.LinAlgError):
        au.oracle_aug(x[1:2], P[1:2], u[1:2], z[1:2])
    keep = [0, 2]
    want = au.oracle_aug(x[keep], P[keep], u[keep], z[keep])
    cond = au.oracle_conditioning(x[keep], P[keep], u[keep], z[keep], want)
    au.assert_aug_close(want, (a[0][keep], a[1][keep]), "neighbours", cond=cond)


@pytest.mark.parametrize("flags", [7, 3, 2])
def test_aug_scratch_is_written_before_it_is_read(flags):
    """A workspace full of NaNs gives the same bits as a zeroed one (no stage reads scratch it did not write), in the
    re-draw, no-re-draw and update-only orders of stages."""
    x, P, u, z = au.aug_case(2, 18, seed=21)
    a = eu.emu_ukf_aug_step(x, P, u, z, flags=flags)
    b = eu.emu_ukf_aug_step(x, P, u, z, flags=flags, ws_fill=0xFF)
    for p, q in zip(a, b):
        assert np.array_equal(p, q)
    assert np.isfinite(b[0]).all() and np.isfinite(b[1]).all()



def test_aug_bad_arguments():
    h = eu.emu_handle()
    from lidar_slam_b200._lib import LslamError
    x, P, u, z = au.aug_case(1, 2, seed=1)
    st = np.zeros(1, np.int32)
    ws = np.zeros(h.ukf_aug_workspace_bytes(2, 1), np.uint8)
    par = batch.Engine.ukf_params()
    with pytest.raises(LslamError):      # workspace too small for one filter
        h.ukf_aug_step(eu.ptr(x), eu.ptr(P), eu.ptr(u), eu.ptr(z), 1, 2, par, 7, eu.ptr(st), eu.ptr(ws), 1024)
    with pytest.raises(LslamError):      # update without z
        h.ukf_aug_step(eu.ptr(x), eu.ptr(P), eu.ptr(u), None, 1, 2, par, 7, eu.ptr(st), eu.ptr(ws), ws.size)
    with pytest.raises(LslamError):      # L = 0
        h.ukf_aug_step(eu.ptr(x), eu.ptr(P), eu.ptr(u), eu.ptr(z), 1, 0, par, 7, eu.ptr(st), eu.ptr(ws), ws.size)
    with pytest.raises(LslamError):
        h.ukf_aug_workspace_bytes(0, 1)


def test_failed_redraw_or_S_leaves_the_prior_in_both_filters():
    """filterpy's predict() has stored the prior before the sigma points are re-drawn, and update() is what fails on a
    bad S: in both cases the filter is left at the PRIOR.  The n = 3 kernel and the augmented path must agree on that
    (status LSLAM_UKF_ST_REDRAW = 2 / LSLAM_UKF_ST_S = 4): the result equals a predict-only call bit for bit."""
    from oracle import cport
    bad_q = cport.ukf_params(Q=-1e3 * np.eye(3))            # prior covariance not positive definite -> re-draw fails
    bad_r = cport.ukf_params(var_dist=-1e12)                # S not positive definite
    x, P, u, z = au.aug_case(3, 3, seed=9)
    for params, bit in ((bad_q, 2), (bad_r, 4)):
        prior = eu.emu_ukf_aug_step(x, P, u, None, params=params, flags=1)
        got = eu.emu_ukf_aug_step(x, P, u, z, params=params, flags=7, workspace_filters=2)
        assert got[2].tolist() == [bit] * 3 and not prior[2].any()
        assert np.array_equal(got[0], prior[0]) and np.array_equal(got[1], prior[1])
        assert not np.array_equal(got[0], x)
    # the same two failures in the n = 3 filter
    rng = np.random.default_rng(4)
    F, L = 5, 4
    x3 = rng.normal(0, 1, (F, 3)); P3 = np.tile(np.diag([.1, .1, .05]), (F, 1, 1)); u3 = rng.uniform(.5, 1.5, (F, 2))
    lm = rng.normal(0, 1500, (F, L, 2))
    z3 = np.stack([np.hypot(lm[..., 0], lm[..., 1]), np.arctan2(lm[..., 1], lm[..., 0])], -1)
    for params, bit in ((bad_q, 2), (bad_r, 4)):
        prior = eu.emu_ukf_step(x3, P3, u3, None, None, params=params, flags=1)
        got = eu.emu_ukf_step(x3, P3, u3, z3, lm, params=params, flags=7)
        assert got[2].tolist() == [bit] * F and not prior[2].any()
        assert np.array_equal(got[0], prior[0]) and np.array_equal(got[1], prior[1])
