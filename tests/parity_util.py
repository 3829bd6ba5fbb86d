"""Shared comparison helpers: the parity bar of BASELINE.json north_star.

  bit-exact : inlier counts, inlier masks, selected hypothesis (trial + sample pair), and the
              winner's sum of squared residuals (same summation tree on both sides)
  1e-5 rel  : fitted line parameters (a, b, origin, direction, tip) - the refit sums run in a
              different order on the GPU.  A small absolute floor covers coordinates that are
              analytically zero (e.g. the centroid of a symmetric wall).
"""
import numpy as np

RTOL = 1e-5          # north_star: "within 1e-5 relative for fitted line parameters"
ATOL_MM = 1e-6       # absolute floor in millimetres for quantities that cancel to ~0


def assert_ransac_equal(want, got, what=""):
    for k in ("ninl", "win_trial", "win_pair"):
        assert np.array_equal(want[k], np.asarray(got[k])), "%s %s differs" % (what, k)
    wm = np.asarray(want["inl_mask"]).view(np.uint32)
    gm = np.asarray(got["inl_mask"]).view(np.uint32)
    assert np.array_equal(wm, gm), "%s inlier masks differ" % what
    if "ssum" in got:                      # optional output (NULL at the ABI: the fast decide path)
        ws = np.asarray(want["ssum"]).view(np.uint64)
        gs = np.asarray(got["ssum"]).view(np.uint64)
        valid = np.asarray(want["ninl"]) > 0
        assert np.array_equal(ws[valid], gs[valid]), "%s ssum not bit-identical" % what
    wl, gl = np.asarray(want["line"]), np.asarray(got["line"])
    assert wl.shape == gl.shape
    both_nan = np.isnan(wl) & np.isnan(gl)
    same_inf = np.isinf(wl) & (wl == gl)
    ok = both_nan | same_inf | np.isclose(wl, gl, rtol=RTOL, atol=ATOL_MM, equal_nan=True)
    # a (slope) of a near-vertical line is ill-conditioned (ransac_functions.py:26 has no guard):
    # there the direction vector, which is compared too, is the meaningful quantity
    steep = np.abs(wl[..., 4]) < 1e-6
    ok[..., 0] |= steep
    ok[..., 1] |= steep
    ok[..., 7] |= steep
    assert ok.all(), "%s line parameters differ: max abs diff %g" % (
        what, np.nanmax(np.abs(np.where(ok, 0, wl - gl))))


def mask_to_bool(words, n):
    words = np.asarray(words).view(np.uint32)
    k = np.arange(n)
    return ((words[k >> 5] >> (k & 31).astype(np.uint32)) & 1).astype(bool)
