"""GPU parity (run on the B200 box with -m gpu): the sm_100a kernels through the C ABI
against the C oracle on the same seeded inputs.  Bar (BASELINE.json north_star): bit-exact
inlier counts / masks / selected hypotheses / residual sums, 1e-5 relative on line
parameters (tests/parity_util.py)."""
import os

import numpy as np
import pytest

from lidar_slam_b200 import synth
from tests.parity_util import assert_ransac_equal

pytestmark = pytest.mark.gpu


WANT_SSUM = True


@pytest.fixture(params=[(2, True), (1, True), (1, False), (3, True), (3, False)],
                ids=["ws", "phased", "phased-nossum", "tile", "tile-nossum"], autouse=True)
def ransac_kernel(request, gpu_engine):
    """Both RANSAC kernels: 2 = warp-specialised persistent, 1 = phase-structured (also without the
    optional residual-sum output: the fast decide path)."""
    global WANT_SSUM
    gpu_engine.set_option("ransac.kernel", request.param[0])
    WANT_SSUM = request.param[1]
    yield request.param[0]
    gpu_engine.set_option("ransac.kernel", 0)
    WANT_SSUM = True


def test_ws_role_shapes(gpu_engine, cport, ransac_kernel):
    if ransac_kernel != 2:
        pytest.skip("options of kernel 2")
    xy = synth.random_room_scans(3000, seed=8)
    npts = np.full(3000, 360, np.int32)
    npts[::7] = 301
    npts[5::11] = 2
    want = cport.ransac_lines(xy, npts=npts, seed=2)
    try:
        for ns, np_, ring, g in ((1, 1, 2, 1), (3, 2, 2, 5), (8, 8, 4, 0), (2, 5, 6, 3), (12, 7, 3, 16), (10, 6, 3, 16), (16, 14, 3, 16), (8, 22, 2, 8)):
            for k, v in (("ransac.scorers", ns), ("ransac.posts", np_), ("ransac.ring", ring), ("ransac.slots_per_cta", g)):
                gpu_engine.set_option(k, v)
            got = run_gpu(gpu_engine, xy, npts=npts, seed=2)
            assert_ransac_equal(want, got, "ns=%d np=%d ring=%d g=%d" % (ns, np_, ring, g))
    finally:
        for k in ("ransac.scorers", "ransac.posts", "ransac.ring", "ransac.slots_per_cta"):
            gpu_engine.set_option(k, 0)


def to_np(out):
    import torch
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


def run_gpu(eng, xy, npts=None, **kw):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(xy)).to(eng.device)
    n = torch.from_numpy(npts).to(eng.device) if npts is not None else None
    out = None
    if not WANT_SSUM:
        out = eng.alloc_ransac_out(t.shape[0], t.shape[1], kw.get("chunk", 100), ssum=False)
    return to_np(eng.ransac_lines(t, npts=n, out=out, **kw))


def test_library_is_the_cuda_build(gpu_engine):
    from lidar_slam_b200 import _lib
    assert os.path.basename(_lib.DEFAULT_PATH) == "liblslam.so"
    assert gpu_engine.launch_count() >= 0


def test_c1_golden(gpu_engine, cport):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ransac_c1.npz"))
    got = run_gpu(gpu_engine, g["xy"][None])
    assert np.array_equal(got["ninl"][0], g["ninl"])
    assert np.array_equal(got["win_trial"][0], g["win_trial"])
    assert np.array_equal(got["win_pair"][0], g["win_pair"])
    if "ssum" in got:
        assert np.array_equal(got["ssum"][0].view(np.uint64), g["ssum"].view(np.uint64))
    np.testing.assert_allclose(got["line"][0, :, 0], g["a"], rtol=1e-5)
    np.testing.assert_allclose(got["line"][0, :, 1], g["b"], rtol=1e-5)
    np.testing.assert_allclose(got["line"][0, :, 6:8], g["tip"], rtol=1e-5)
    assert_ransac_equal(cport.ransac_lines(g["xy"][None]), got, "C1")


def test_c2_4096_scans_all_layouts(gpu_engine, cport):
    """BASELINE.json configs[1]: 4096 synthetic 360-point scans, RANSAC line extraction."""
    xy = synth.random_room_scans(4096, seed=2)
    want = cport.ransac_lines(xy, seed=11, scan_base=5 * 10**12)
    try:
        for tpt, slots, threads in ((0, 0, 0), (1, 0, 0), (2, 8, 256), (4, 16, 512), (2, 5, 768), (4, 32, 0)):
            gpu_engine.set_option("ransac.tpt", tpt)
            gpu_engine.set_option("ransac.slots_per_cta", slots)
            gpu_engine.set_option("ransac.threads", threads)
            got = run_gpu(gpu_engine, xy, seed=11, scan_base=5 * 10**12)
            assert_ransac_equal(want, got, "C2 tpt=%d slots=%d threads=%d" % (tpt, slots, threads))
    finally:
        for k in ("ransac.tpt", "ransac.slots_per_cta", "ransac.threads"):
            gpu_engine.set_option(k, 0)
    assert gpu_engine.ransac_counters()["faults"] == 0


def test_c3_highres(gpu_engine, cport):
    """BASELINE.json configs[2]: 8192-point scans, 1024 hypotheses/line, 20 lines/scan."""
    xy = np.stack([synth.polygon_scan(beams=8192, rng=np.random.default_rng(s)) for s in range(6)])
    want = cport.ransac_lines(xy, chunk=410, trials=1024, seed=3)
    got = run_gpu(gpu_engine, xy, chunk=410, trials=1024, seed=3)
    assert_ransac_equal(want, got, "C3")
    assert (got["ninl"] > 100).all()


def test_ragged_and_status(gpu_engine, cport):
    xy = synth.random_room_scans(12, seed=5)
    npts = np.array([360, 359, 303, 302, 301, 300, 100, 2, 1, 0, 57, 202], np.int32)
    assert_ransac_equal(cport.ransac_lines(xy, npts=npts), run_gpu(gpu_engine, xy, npts=npts), "ragged")
    bad = synth.random_room_scans(1, seed=1)
    bad[0, 150, 0] = np.nan
    got = run_gpu(gpu_engine, bad)
    assert got["ninl"][0, 1] == -5 and got["ninl"][0, 0] > 2


def test_degenerate_geometry(gpu_engine, cport):
    t = np.linspace(-1000, 1000, 100)
    y = np.zeros(100)
    y[5::7] = 20.0
    y[3::11] = np.nextafter(20.0, 0.0)
    y[2::13] = np.nextafter(20.0, 100.0)
    cases = {
        "collinear": np.stack([t, 0.5 * t + 100], -1),
        "vertical": np.stack([np.full(100, 1500.0), t], -1),
        "all-zero": np.zeros((100, 2)),
        "two-clusters": np.concatenate([np.zeros((50, 2)), np.tile([[100.0, 5.0]], (50, 1))]),
        "tiny-scale": np.stack([t, 0.5 * t], -1) * 1e-250,
        "threshold-edge": np.stack([np.arange(100.0) * 16.0, y], -1),
    }
    gpu_engine.ransac_counters()
    for name, data in cases.items():
        assert_ransac_equal(cport.ransac_lines(data[None], seed=4), run_gpu(gpu_engine, data[None], seed=4), name)
    c = gpu_engine.ransac_counters()
    assert c["faults"] == 0 and c["literal_rescored"] > 0


def test_word_tails_of_every_length(gpu_engine, cport):
    """As tests/test_emu_ransac.py: chunk lengths that leave tails of 1 .. 31 points in the tile kernel's two-bit words."""
    xy = synth.random_room_scans(64, seed=31)
    for chunk in (33, 34, 35, 47, 48, 62, 63, 64, 65, 95, 97):
        want = cport.ransac_lines(xy, chunk=chunk, trials=48, seed=chunk)
        assert_ransac_equal(want, run_gpu(gpu_engine, xy, chunk=chunk, trials=48, seed=chunk), "chunk %d" % chunk)
    npts = np.tile(np.array([90, 77, 360, 38], np.int32), 16)
    want = cport.ransac_lines(xy, npts=npts, chunk=37, trials=48, seed=5)
    assert_ransac_equal(want, run_gpu(gpu_engine, xy, npts=npts, chunk=37, trials=48, seed=5), "ragged chunk 37")
    assert gpu_engine.ransac_counters()["faults"] == 0


def test_screening_range_edges(gpu_engine, cport):
    """As tests/test_emu_ransac.py: the fp32 screening's range (tile_screen32) - chunks kilometres wide, a cluster far from
    the origin, thresholds the band is not small against, points inside the fp32 band but outside the fp64 one."""
    rng = np.random.default_rng(7)
    t = np.linspace(-1000, 1000, 100)
    y = 0.05 * t + rng.normal(0, 8, 100)
    near = np.zeros(100)
    near[4::9] = 20.0 + 5e-4
    near[5::9] = 20.0 - 5e-4
    near[6::9] = 20.0 + 1.5e-3
    near[7::9] = 20.0 - 1.5e-3
    cases = {
        "km-wide": (np.stack([t * 5000.0, y * 5000.0], -1), 20.0),
        "10-km-wide": (np.stack([t * 2e4, y], -1), 20.0),
        "far-from-origin": (np.stack([t * 0.1 + 1e9, y * 0.1 - 3e9], -1), 20.0),
        "band-fp32-only": (np.stack([np.arange(100.0) * 16.0, near], -1), 20.0),
        "small-threshold": (np.stack([t, y], -1), 1e-3),
        "threshold-0.5": (np.stack([t, y * 0.05], -1), 0.5),
        "large-threshold": (np.stack([t, y], -1), 5e3),
    }
    for name, (data, thr) in cases.items():
        want = cport.ransac_lines(data[None], thr=thr, seed=9)
        assert_ransac_equal(want, run_gpu(gpu_engine, data[None], thr=thr, seed=9), name)
    assert gpu_engine.ransac_counters()["faults"] == 0


def test_full_size_properties(gpu_engine):
    """At full size (262,144 device-generated scans) the oracle is too slow; check
    size-independent properties: determinism, batch-split invariance (scan_base), inlier count
    equals the mask population, winner pair inside the chunk."""
    import torch
    eng = gpu_engine
    B = 262144
    xy = eng.synth_scans(B, seed=9)
    a = eng.ransac_lines(xy, seed=1)
    b = eng.ransac_lines(xy, seed=1)
    torch.cuda.synchronize()
    for k in a:
        assert torch.equal(a[k].view(torch.uint8), b[k].view(torch.uint8)), k
    half = B // 2
    lo = eng.ransac_lines(xy[:half].contiguous(), seed=1, scan_base=0)
    hi = eng.ransac_lines(xy[half:].contiguous(), seed=1, scan_base=half)
    for k in a:
        assert torch.equal(torch.cat([lo[k], hi[k]]).view(torch.uint8), a[k].view(torch.uint8)), k
    ninl = a["ninl"]
    assert (ninl >= 2).all()
    m = a["inl_mask"]
    pop = torch.zeros_like(ninl)
    for bit in range(32):
        pop += ((m >> bit) & 1).sum(-1).to(torch.int32)
    assert torch.equal(pop, ninl)
    n_chunk = torch.tensor([100, 100, 100, 60], device=eng.device)
    assert (a["win_pair"] >= 0).all() and (a["win_pair"] < n_chunk[None, :, None]).all()
    assert (a["win_pair"][..., 0] != a["win_pair"][..., 1]).all()
    assert eng.ransac_counters()["faults"] == 0


def test_synth_and_polar(gpu_engine):
    import torch
    eng = gpu_engine
    xy = eng.synth_scans(64, seed=3).cpu().numpy()
    r = np.hypot(xy[..., 0], xy[..., 1])
    assert r.max() < 6000 * 1.5 and (r == 0).mean() < 0.06 and (r == 0).mean() > 0.002
    d = np.linspace(0, 6000, 1000)
    ang = np.linspace(0, 360, 1000)
    got = eng.polar_to_cart(torch.from_numpy(d).to(eng.device), torch.from_numpy(ang).to(eng.device)).cpu().numpy()
    np.testing.assert_allclose(got, synth.polar_to_cart(d, ang), rtol=1e-12, atol=1e-9)
