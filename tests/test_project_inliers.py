"""lslam_project_inliers (R6b): the projected inlier points landmark_extraction returns
(/root/reference/ransac_functions.py:28,46,49,53,55) against oracle/ransac_line.py::project_inliers - bit-exact
(one product and one sum per point)."""
import numpy as np
import pytest

from lidar_slam_b200 import synth
from oracle import ransac_line as rl
from tests.parity_util import mask_to_bool


def oracle_points(xy, npts, r, ab, chunk):
    """Per slot, in slot order: oracle projection with the slot's coefficients."""
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    parts = []
    for b in range(B):
        nb = N if npts is None else int(npts[b])
        for c in range(C):
            if r["ninl"][b, c] < 2:
                parts.append(np.zeros((0, 2)))
                continue
            s, e = c * chunk, min((c + 1) * chunk, nb)
            inl = mask_to_bool(np.asarray(r["inl_mask"][b, c]), e - s)
            parts.append(rl.project_inliers(xy[b, s:e], inl, ab[b, c, 0], ab[b, c, 1]))
    return parts


def check(xy, npts, r, ab_in, run, chunk=100):
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    ab = np.asarray(r["line"])[..., :2] if ab_in is None else ab_in
    parts = oracle_points(xy, npts, r, ab, chunk)
    pts, off = run(compact=True)
    want = np.concatenate(parts, 0)
    assert pts.shape == want.shape and off[-1] == want.shape[0]
    assert np.array_equal(pts.view(np.uint64), want.view(np.uint64)), "compact layout not bit-identical"
    pad = run(compact=False)
    for k, p in enumerate(parts):
        got = pad.reshape(B * C, chunk, 2)[k]
        assert np.array_equal(got[:len(p)].view(np.uint64), p.view(np.uint64))
        assert np.isnan(got[len(p):]).all()                         # rows past ninl untouched


def test_emulated_projection_matches_oracle(emu, cport):
    xy = synth.random_room_scans(9, seed=41)
    npts = np.array([360, 360, 301, 360, 2, 360, 259, 360, 0], np.int32)
    r = cport.ransac_lines(xy, npts=npts, seed=4)
    rng = np.random.default_rng(0)
    ab = rng.normal(size=(9, 4, 2)) * np.array([2.0, 500.0])       # e.g. a stored landmark's coefficients
    for coeff in (None, ab):
        check(xy, npts, r, coeff, lambda compact: emu.emu_project_inliers(xy, r, ab=coeff, npts=npts, compact=compact))


@pytest.mark.gpu
def test_gpu_projection_matches_oracle(gpu_engine, cport):
    import torch
    eng = gpu_engine
    xy = synth.random_room_scans(2000, seed=42)
    npts = np.full(2000, 360, np.int32)
    npts[::7] = 301
    npts[5::11] = 2
    dev = eng.device
    txy, tn = torch.from_numpy(xy).to(dev), torch.from_numpy(npts).to(dev)
    rg = eng.ransac_lines(txy, npts=tn, seed=4)
    r = {k: v.cpu().numpy() for k, v in rg.items()}
    rng = np.random.default_rng(1)
    ab = rng.normal(size=(2000, 4, 2)) * np.array([2.0, 500.0])

    def run(coeff):
        def f(compact):
            out = eng.project_inliers(txy, rg, ab=None if coeff is None else torch.from_numpy(coeff).to(dev), npts=tn, compact=compact)
            torch.cuda.synchronize()
            return (out[0].cpu().numpy(), out[1].cpu().numpy()) if compact else out.cpu().numpy()
        return f
    for coeff in (None, ab):
        check(xy, npts, r, coeff, run(coeff))
