"""GPU parity of lslam_ransac_ols (csrc/ransac_ols.cuh) against the C oracle, which is pinned to the
real scikit-learn RANSACRegressor (tests/golden/ols_golden.npz): integers, masks and float records bit
for bit; and the drop-in line_detection module against the scikit-learn golden outputs."""
import numpy as np
import pytest

from oracle import cport
from tests.ols_util import GOLD, assert_ols_equal, golden_batch, wall_windows

pytestmark = pytest.mark.gpu


def gpu_ols(eng, xy, npts=None, pairs=None, **kw):
    import torch
    t = lambda a, dt: None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(eng.device)
    r = eng.ransac_ols(t(xy, np.float64), npts=t(npts, np.int32), pairs=t(pairs, np.int32), **kw)
    torch.cuda.synchronize()
    out = {k: v.cpu().numpy() for k, v in r.items()}
    out["inl_mask"] = out["inl_mask"].view(np.uint32)
    return out


def test_golden_windows_explicit_pairs(gpu_engine):
    xy, npts, pairs = golden_batch()
    want = cport.ransac_ols(xy, npts=npts, pairs=pairs)
    got = gpu_ols(gpu_engine, xy, npts, pairs)
    assert_ols_equal(want, got, "golden")
    # and straight against scikit-learn's own outputs
    assert np.array_equal(got["n_trials"][:, 0], GOLD["n_trials"])
    np.testing.assert_allclose(got["fit"][:, 0, 0], GOLD["coef"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(got["fit"][:, 0, 1], GOLD["intercept"], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("N,window,B", [(360, 50, 4096), (130, 64, 512), (1000, 1000, 32), (51, 50, 64)])
def test_philox_batches(gpu_engine, N, window, B):
    xy = wall_windows(B, N, seed=N + 1)
    npts = np.full(B, N, np.int32)
    npts[3], npts[5] = N - 7, 0
    xy[7, :, 1] = np.nan
    want = cport.ransac_ols(xy, npts=npts, window=window, seed=4, scan_base=1 << 33)
    got = gpu_ols(gpu_engine, xy, npts, window=window, seed=4, scan_base=1 << 33)
    assert_ols_equal(want, got, "philox")
    assert want["ninl"][7, 0] == -3 and (want["ninl"][5] == 0).all()


@pytest.mark.parametrize("trials", [1, 5, 31, 32, 33, 64, 100, 257])
def test_trial_budgets_and_stop_probabilities(gpu_engine, trials):
    """As tests/test_oracle_ols.py: budgets around the kernel's groups of 32 speculative trials, clean / noisy / mixed windows,
    three stop probabilities."""
    xy = np.concatenate([wall_windows(64, 100, seed=trials, noise=0.0, outliers=0.0),
                         wall_windows(64, 100, seed=trials + 1, noise=40.0, outliers=0.5), wall_windows(64, 100, seed=trials + 2)])
    for prob in (0.99, 0.5, 1.0):
        want = cport.ransac_ols(xy, trials=trials, probability=prob, seed=11)
        got = gpu_ols(gpu_engine, xy, trials=trials, stop_probability=prob, seed=11)
        assert_ols_equal(want, got, "trials %d p %.2f" % (trials, prob))


def test_bad_arguments(gpu_engine):
    import torch
    from lidar_slam_b200._lib import LslamError
    xy = torch.zeros((2, 100, 2), dtype=torch.float64, device=gpu_engine.device)
    with pytest.raises(LslamError):
        gpu_engine.ransac_ols(xy, window=1)
    with pytest.raises(LslamError):
        gpu_engine.ransac_ols(xy, window=2000)
    with pytest.raises(ValueError):
        gpu_engine.ransac_ols(xy, pairs=torch.zeros((2, 2, 5, 2), dtype=torch.int32, device=gpu_engine.device))


def test_line_detection_dropin(gpu_engine, tmp_path):
    """The reference script's flow: text file of (distance, angle) records -> windows of 50 -> lines."""
    from lidar_slam_b200 import line_detection as ld
    assert (ld.CHUNK_SIZE, ld.PNT_NBR) == (36, 50)
    rng = np.random.default_rng(3)
    ang = np.arange(0, 120, 0.5)
    d = 2000.0 / np.cos(np.radians(ang - 60)) + rng.standard_normal(len(ang))       # one wall
    d[::17] = 0                                                                      # dropped returns
    f = tmp_path / "scan_prop.txt"
    f.write_text("".join("%.3f, %.1f\n" % (a, b) for a, b in zip(d, ang)))
    data = ld.get_data(str(f))
    assert len(data) == 2 * len(ang)
    x, y = ld.polar_records_to_points(data)
    assert len(x) == int((d.round(3) != 0).sum())
    wins, final = ld.main(str(f))
    xy = np.stack([x, y], -1)[None]
    want = cport.ransac_ols(np.ascontiguousarray(xy), window=50, seed=ld.SEED)
    assert len(wins) == want["ninl"].shape[1]
    for c, w in enumerate(wins):
        assert w["coef"] == want["fit"][0, c, 0] and w["intercept"] == want["fit"][0, c, 1]
        assert int(w["inlier_mask"].sum()) == want["ninl"][0, c] and w["n_trials"] == want["n_trials"][0, c]
        assert np.array_equal(w["inliers_x"], x[50 * c:50 * c + 50][w["inlier_mask"]])
        assert len(w["line_X"]) == len(np.arange(x[50 * c:50 * c + 50].min(), x[50 * c:50 * c + 50].max()))
    assert ld.get_data(str(tmp_path / "missing.txt")) is None
    # replaying scikit-learn's own draws reproduces scikit-learn's masks (golden window 0)
    n = int(GOLD["n"][0])
    w = ld.detect_lines(GOLD["x"][0, :n], GOLD["y"][0, :n], pairs=GOLD["pairs"][0][None])
    assert np.array_equal(w[0]["inlier_mask"], GOLD["inlier_mask"][0, :n]) and w[0]["n_trials"] == GOLD["n_trials"][0]
