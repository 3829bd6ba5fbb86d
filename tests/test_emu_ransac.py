"""The RANSAC kernel SOURCE (lidar_slam_b200/csrc/ransac.cuh) executed on the CPU through
tests/emu/cuda_emu.h, compared with the C oracle through the same C ABI.  This is the
CPU-side proof of the kernel logic (fast pass + exact re-check + literal tie-break); the
`-m gpu` tests repeat it on the real sm_100a build."""
import numpy as np
import pytest

from lidar_slam_b200 import synth
from tests.parity_util import assert_ransac_equal


@pytest.fixture(params=[(2, True), (1, True), (1, False), (3, True), (3, False)],
                ids=["ws", "phased", "phased-nossum", "tile", "tile-nossum"], autouse=True)
def ransac_kernel(request, emu):
    """Every test runs on both RANSAC kernels: 2 = warp-specialised persistent (csrc/ransac_ws.cuh),
    1 = phase-structured (csrc/ransac.cuh) - the latter also without the optional residual-sum output,
    which switches the decide stage of unambiguous chunks to the fast model."""
    h = emu.emu_handle()
    h.set_option("ransac.kernel", request.param[0])
    emu.WANT_SSUM = request.param[1]
    yield request.param[0]
    h.set_option("ransac.kernel", 0)
    emu.WANT_SSUM = True


def test_ws_role_shapes(emu, cport, ransac_kernel):
    """Warp-specialised kernel: results do not depend on the number of scorer / post warps, the ring
    depth or the tile size (incl. a ring deeper than the tile count and a single post warp)."""
    if ransac_kernel != 2:
        pytest.skip("options of kernel 2")
    xy = synth.random_room_scans(9, seed=8)
    npts = np.array([360, 360, 301, 360, 2, 360, 360, 259, 360], np.int32)
    want = cport.ransac_lines(xy, npts=npts, seed=2)
    h = emu.emu_handle()
    try:
        for ns, np_, ring, g in ((1, 1, 2, 1), (3, 2, 2, 5), (8, 8, 4, 0), (2, 5, 6, 3), (12, 7, 3, 16), (14, 16, 2, 4)):
            h.set_option("ransac.scorers", ns)
            h.set_option("ransac.posts", np_)
            h.set_option("ransac.ring", ring)
            h.set_option("ransac.slots_per_cta", g)
            got = emu.emu_ransac_lines(xy, npts=npts, seed=2)
            assert_ransac_equal(want, got, "ns=%d np=%d ring=%d g=%d" % (ns, np_, ring, g))
    finally:
        for k in ("ransac.scorers", "ransac.posts", "ransac.ring", "ransac.slots_per_cta"):
            h.set_option(k, 0)


def test_random_rooms_all_thread_layouts(emu, cport):
    xy = synth.random_room_scans(12, seed=3)
    want = cport.ransac_lines(xy, seed=7, scan_base=123456789012)
    h = emu.emu_handle()
    try:
        for tpt in (1, 2, 4):
            h.set_option("ransac.tpt", tpt)
            got = emu.emu_ransac_lines(xy, seed=7, scan_base=123456789012)
            assert_ransac_equal(want, got, "tpt=%d" % tpt)
        h.set_option("ransac.tpt", 0)
        for g in (1, 3, 7):
            h.set_option("ransac.slots_per_cta", g)
            got = emu.emu_ransac_lines(xy, seed=7, scan_base=123456789012)
            assert_ransac_equal(want, got, "slots_per_cta=%d" % g)
    finally:
        h.set_option("ransac.tpt", 0)
        h.set_option("ransac.slots_per_cta", 0)
    assert h.ransac_counters()["faults"] == 0


def test_ragged_scans(emu, cport):
    xy = synth.random_room_scans(12, seed=5)
    npts = np.array([360, 359, 303, 302, 301, 300, 100, 2, 1, 0, 57, 202], np.int32)
    want = cport.ransac_lines(xy, npts=npts)
    got = emu.emu_ransac_lines(xy, npts=npts)
    assert_ransac_equal(want, got, "ragged")
    assert want["ninl"][3, 3] == 0 and want["ninl"][2, 3] == 3      # remainder of 2 dropped, of 3 kept


def test_highres_chunks(emu, cport):
    xy = synth.polygon_scan(beams=1640, nsides=5)[None]
    want = cport.ransac_lines(xy, chunk=410, trials=256, seed=1)
    got = emu.emu_ransac_lines(xy, chunk=410, trials=256, seed=1)
    assert_ransac_equal(want, got, "chunk 410")


def test_degenerate_geometry(emu, cport):
    t = np.linspace(-1000, 1000, 100)
    cases = {
        "collinear": np.stack([t, 0.5 * t + 100], -1),
        "vertical": np.stack([np.full(100, 1500.0), t], -1),
        "all-zero": np.zeros((100, 2)),
        "two-clusters": np.concatenate([np.zeros((50, 2)), np.tile([[100.0, 5.0]], (50, 1))]),
        "tiny-scale": np.stack([t, 0.5 * t], -1) * 1e-250,
    }
    h = emu.emu_handle()
    for name, data in cases.items():
        want = cport.ransac_lines(data[None], seed=4)
        got = emu.emu_ransac_lines(data[None], seed=4)
        assert_ransac_equal(want, got, name)
    assert h.ransac_counters()["faults"] == 0


def test_threshold_edge_points_are_rechecked(emu, cport):
    """Points placed within rounding distance of the threshold force the re-check band: the
    fast pass must hand those trials to the literal formula and still agree bit for bit."""
    rng = np.random.default_rng(0)
    t = np.arange(100, dtype=np.float64) * 16.0
    y = np.zeros(100)
    y[5::7] = 20.0                                    # exactly on the threshold of y = 0
    y[3::11] = np.nextafter(20.0, 0.0)
    y[2::13] = np.nextafter(20.0, 100.0)
    data = np.stack([t, y], -1)
    want = cport.ransac_lines(data[None], seed=2)
    h = emu.emu_handle()
    h.ransac_counters()
    got = emu.emu_ransac_lines(data[None], seed=2)
    assert_ransac_equal(want, got, "threshold edge")
    c = h.ransac_counters()
    assert c["literal_rescored"] > 0 and c["faults"] == 0


def test_word_tails_of_every_length(emu, cport):
    """The tile kernel's scoring loop packs two bits per point, the even and the odd points of a 32-point word in two registers:
    chunk lengths that leave tails of 1 .. 31 points (odd and even, with and without full words before them)."""
    xy = synth.random_room_scans(2, seed=31)
    for chunk in (33, 34, 35, 47, 48, 62, 63, 64, 65, 95, 97):
        want = cport.ransac_lines(xy, chunk=chunk, trials=48, seed=chunk)
        got = emu.emu_ransac_lines(xy, chunk=chunk, trials=48, seed=chunk)
        assert_ransac_equal(want, got, "chunk %d" % chunk)
    npts = np.array([90, 77], np.int32)                    # ragged: a last chunk shorter than the others
    want = cport.ransac_lines(xy, npts=npts, chunk=37, trials=48, seed=5)
    assert_ransac_equal(want, emu.emu_ransac_lines(xy, npts=npts, chunk=37, trials=48, seed=5), "ragged chunk 37")
    assert emu.emu_handle().ransac_counters()["faults"] == 0


def test_screening_range_edges(emu, cport):
    """fp32 screening (tile_screen32): the re-check band is scaled to a power of two; outside its range the slot is scored
    literally.  Chunks kilometres wide (band' > 1), a cluster far from the origin (coordinates 2^22 radii away), thresholds
    the band is not small against, and points inside the fp32 band but outside the fp64 one."""
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
        got = emu.emu_ransac_lines(data[None], thr=thr, seed=9)
        assert_ransac_equal(want, got, name)
    assert emu.emu_handle().ransac_counters()["faults"] == 0


def test_status_codes(emu):
    bad = synth.random_room_scans(1, seed=1)
    bad[0, 150, 0] = np.nan
    bad[0, 250, 1] = np.inf
    got = emu.emu_ransac_lines(bad)
    assert got["ninl"][0, 1] == -5 and got["ninl"][0, 2] == -5      # LSLAM_ERR_NONFINITE
    assert got["ninl"][0, 0] > 2 and got["ninl"][0, 3] > 2
    data = np.array([[0.0, 0.0], [10.0, 0.0], [20.0, 5.0]])
    got = emu.emu_ransac_lines(data[None], chunk=3, trials=10, thr=0.0)
    assert got["ninl"][0, 0] == -3 and got["win_trial"][0, 0] == -1   # LSLAM_ERR_NO_MODEL


def test_bad_arguments_raise(emu):
    from lidar_slam_b200._lib import LslamError
    xy = np.zeros((1, 10, 2))
    with pytest.raises(LslamError):
        emu.emu_ransac_lines(xy, chunk=1)
    with pytest.raises(LslamError):
        emu.emu_ransac_lines(xy, trials=0)
    with pytest.raises(LslamError):
        emu.emu_ransac_lines(xy, thr=float("nan"))
    with pytest.raises(LslamError):
        emu.emu_handle().set_option("no.such.option", 1)
