"""OLS-flavour RANSAC (line_detection.py:51-73): the restatement against the REAL
scikit-learn RANSACRegressor (tests/golden/ols_golden.npz, made by make_ols_golden.py), its C twin bit
for bit, and the kernel source under emulation bit for bit."""
import math

import numpy as np
import pytest

from oracle import cport, ransac_ols as ro
from tests.emu_util import emu_ransac_ols
from tests.ols_util import GOLD, assert_ols_equal, golden_batch, mask_bool, wall_windows


def test_restatement_matches_scikit_learn():
    for k in range(len(GOLD["n"])):
        n = int(GOLD["n"][k])
        r = ro.fit_window(GOLD["x"][k, :n], GOLD["y"][k, :n], GOLD["pairs"][k])
        assert np.array_equal(r["inlier_mask"], GOLD["inlier_mask"][k, :n]), k          # inlier_mask_
        assert r["n_trials"] == GOLD["n_trials"][k], k                                  # n_trials_
        # coefficients: sklearn solves the same least squares through LAPACK's SVD
        assert abs(r["coef"] - GOLD["coef"][k]) <= 1e-9 * max(1.0, abs(GOLD["coef"][k])), k
        assert abs(r["intercept"] - GOLD["intercept"][k]) <= 1e-9 * max(1.0, abs(GOLD["intercept"][k])), k
    assert len(set(GOLD["n_trials"].tolist())) > 10          # the dynamic stop rule is exercised


def test_c_twin_and_emulated_kernel_on_golden_windows():
    xy, npts, pairs = golden_batch()
    want = cport.ransac_ols(xy, npts=npts, pairs=pairs)
    for k in range(len(npts)):
        n = int(npts[k])
        r = ro.fit_window(xy[k, :n, 0], xy[k, :n, 1], pairs[k, 0])
        assert np.array_equal(mask_bool(want["inl_mask"][k, 0], n), r["inlier_mask"])
        assert (want["ninl"][k, 0], want["n_trials"][k, 0], want["best_trial"][k, 0]) == \
            (r["n_inliers"], r["n_trials"], r["best_trial"])
        assert want["fit"][k, 0, 0] == r["coef"] and want["fit"][k, 0, 1] == r["intercept"]
        assert want["fit"][k, 0, 2] == r["score"]
        assert want["fit"][k, 0, 3] == xy[k, :n, 0].min() and want["fit"][k, 0, 4] == xy[k, :n, 0].max()
    assert_ols_equal(want, emu_ransac_ols(xy, npts=npts, pairs=pairs), "golden")


@pytest.mark.parametrize("N,window", [(360, 50), (130, 64), (97, 97), (51, 50)])
def test_philox_sampler_batches(N, window):
    xy = wall_windows(24, N, seed=N)
    npts = np.full(24, N, np.int32)
    C = (N + window - 1) // window
    one = window + 1 if C > 1 else 1                            # a scan whose last window holds one point
    npts[3], npts[4], npts[5] = N - 7, one, 0                   # short last window, 1-point window, empty scan
    want = cport.ransac_ols(xy, npts=npts, window=window, seed=9, scan_base=100)
    got = emu_ransac_ols(xy, npts=npts, window=window, seed=9, scan_base=100)
    assert_ols_equal(want, got, "philox %d/%d" % (N, window))
    w1 = 1 if C > 1 else 0
    assert want["ninl"][4, w1] == -2 and want["n_trials"][4, w1] == 0           # sklearn: min_samples > n_samples
    assert (want["ninl"][5] == 0).all()
    assert want["ninl"][0, 0] >= 2 and want["ninl"].shape == (24, C)
    # the pairs the sampler drew are the LineModelND flavour's (same Philox stream layout)
    r = ro.fit_window(xy[0, :window, 0], xy[0, :window, 1], ro.philox_pairs(window, 100, 9, 100, 0))
    assert r["n_trials"] == want["n_trials"][0, 0] and r["coef"] == want["fit"][0, 0, 0]


@pytest.mark.parametrize("trials", [1, 5, 31, 32, 33, 64, 100, 257])
def test_trial_budgets_and_stop_probabilities(trials):
    """The kernel evaluates 32 trials at a time, one per lane, and replays upstream's accept rule over them: budgets that end
    inside, at and just past a group of 32, clean walls (every trial ties on the count: the bound drops to one trial), noisy
    windows (the bound stays at the budget), and other stop probabilities (incl. 1.0: max_trials = inf -> the budget)."""
    clean = wall_windows(4, 100, seed=trials, noise=0.0, outliers=0.0)
    noisy = wall_windows(4, 100, seed=trials + 1, noise=40.0, outliers=0.5)
    mixed = wall_windows(4, 100, seed=trials + 2)
    xy = np.concatenate([clean, noisy, mixed])
    for prob in (0.99, 0.5, 1.0):
        want = cport.ransac_ols(xy, trials=trials, probability=prob, seed=11)
        got = emu_ransac_ols(xy, trials=trials, probability=prob, seed=11)
        assert_ols_equal(want, got, "trials %d p %.2f" % (trials, prob))
        assert (want["n_trials"] <= trials).all() and (want["n_trials"][:4] >= 1).all()
    assert (cport.ransac_ols(xy, trials=trials, probability=0.99, seed=11)["n_trials"][:4] == 1).all()   # clean walls: one trial


def test_no_consensus_and_nonfinite():
    xy = wall_windows(3, 50, seed=1)
    xy[1, :, 1] = np.nan                                    # every residual NaN -> every trial skipped
    xy[2, 10, 0] = np.nan
    for fn in (cport.ransac_ols, emu_ransac_ols):
        r = fn(xy, seed=2)
        assert r["ninl"][1, 0] == -3 and r["best_trial"][1, 0] == -1 and r["n_trials"][1, 0] == 100
        assert np.isnan(r["fit"][1, 0, :3]).all() and not r["inl_mask"][1].any()
        assert r["ninl"][2, 0] >= 2 and np.isnan(r["fit"][2, 0, 3:5]).all() and np.isfinite(r["fit"][2, 0, :3]).all()
        assert not mask_bool(r["inl_mask"][2, 0], 50)[10]
    with pytest.raises(ValueError):
        ro.fit_window(xy[1, :, 0], xy[1, :, 1], ro.philox_pairs(50, 100, 2, 1, 0))


def test_dynamic_max_trials_is_far_from_a_rounding_boundary():
    """The kernel evaluates ceil(log(nom)/log(denom)) with the device's log; a last-bit difference
    from the host's can only matter if the quotient is within a few ulp of an integer.  It never is
    for any window size the entry point accepts."""
    nom = math.log(max(ro._EPSILON, 1 - ro.STOP_PROBABILITY))
    worst = 1.0
    for n in range(2, 1025):
        c = np.arange(1, n, dtype=np.float64)
        q = nom / np.log(np.maximum(ro._EPSILON, 1 - (c / n) ** 2))
        worst = min(worst, float(np.min(np.abs(q - np.rint(q)) / q)))
    assert worst > 1e-12
    assert ro.dynamic_max_trials(50, 50) == 1.0 and ro.dynamic_max_trials(0, 50) == float("inf")
