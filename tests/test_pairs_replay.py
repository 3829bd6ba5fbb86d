"""Explicit sample pairs: a seeded upstream run replayed pair for pair.

skimage.measure.ransac draws ``random_state.choice(n, 2, replace=False)`` per trial (the call at
/root/reference/ransac_functions.py:23-24).  oracle/mt_sampler.py reproduces that draw sequence with the
NumPy generators upstream uses; the NumPy restatement, the C restatement, the kernel sources on the CPU
and (``-m gpu``) the sm_100a kernels must then agree on every output for the SAME pairs.  When a golden
file made where scikit-image is installed exists (tests/golden/make_skimage_golden.py ->
skimage_golden.npz), the restatements and the kernels are additionally held to upstream's own results.
"""
import os

import numpy as np
import pytest

from lidar_slam_b200 import synth
from oracle import mt_sampler, ransac_line as rl
from tests.parity_util import assert_ransac_equal, mask_to_bool

GOLD = os.path.join(os.path.dirname(__file__), "golden", "skimage_golden.npz")


def batch_pairs(xy, npts, chunk, trials, seed0, legacy=True):
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    n_slot = np.zeros((B, C), np.int64)
    for b in range(B):
        nb = N if npts is None else int(npts[b])
        for c in range(C):
            n_slot[b, c] = max(0, min(chunk, nb - c * chunk))
    seeds = seed0 + np.arange(B * C).reshape(B, C)
    return mt_sampler.upstream_pairs_batch(seeds, n_slot, trials, legacy), n_slot


def test_upstream_draws_are_distinct_and_reproducible():
    for legacy in (True, False):
        p = mt_sampler.upstream_pairs(7, 100, 100, legacy)
        assert p.shape == (100, 2) and (p[:, 0] != p[:, 1]).all() and p.min() >= 0 and p.max() < 100
        assert np.array_equal(p, mt_sampler.upstream_pairs(7, 100, 100, legacy))
    # the legacy stream is RandomState(7).choice(100, 2, replace=False) verbatim
    rs = np.random.RandomState(7)
    assert np.array_equal(mt_sampler.upstream_pairs(7, 100, 3), [rs.choice(100, 2, replace=False) for _ in range(3)])


@pytest.mark.parametrize("legacy", [True, False])
def test_numpy_and_c_restatements_agree_on_replayed_pairs(cport, legacy):
    xy = synth.random_room_scans(6, seed=21)
    pairs, n_slot = batch_pairs(xy, None, 100, 100, 1000, legacy)
    got = cport.ransac_lines(xy, pairs=pairs)
    for b in range(xy.shape[0]):
        for c, (s, e) in enumerate(rl.chunk_bounds(360, 100)):
            model, inl, info = rl.ransac_line(xy[b, s:e], pairs=(pairs[b, c, :, 0], pairs[b, c, :, 1]))
            assert info["ninl"] == got["ninl"][b, c] and info["win_trial"] == got["win_trial"][b, c]
            assert tuple(got["win_pair"][b, c]) == info["win_pair"]
            assert np.float64(info["ssum"]).view(np.uint64) == got["ssum"][b, c].view(np.uint64)
            assert np.array_equal(mask_to_bool(got["inl_mask"][b, c], e - s), inl)
            np.testing.assert_allclose(got["line"][b, c, 2:4], model[0], rtol=1e-9)
    # and the explicit pairs really replace the Philox sampler
    assert not np.array_equal(got["win_pair"], cport.ransac_lines(xy)["win_pair"])


@pytest.mark.parametrize("kernel", [1, 3, 0])
def test_emulated_kernels_replay_pairs(emu, cport, kernel):
    xy = synth.random_room_scans(7, seed=22)
    npts = np.array([360, 301, 360, 2, 259, 360, 360], np.int32)
    pairs, _ = batch_pairs(xy, npts, 100, 100, 5)
    pairs[2, 1, 5] = (400, -3)            # out-of-range indices are clamped into the chunk, on both sides
    want = cport.ransac_lines(xy, npts=npts, pairs=pairs)
    h = emu.emu_handle()
    try:
        h.set_option("ransac.kernel", kernel)
        for ssum in (True, False):
            got = emu.emu_ransac_lines(xy, npts=npts, pairs=pairs, want_ssum=ssum)
            assert_ransac_equal(want, got, "kernel %d" % kernel)
    finally:
        h.set_option("ransac.kernel", 0)


def test_emulated_big_chunks_with_pairs_use_a_kernel_that_supports_them(emu, cport):
    xy = synth.random_room_scans(1, beams=820, seed=23)
    pairs, _ = batch_pairs(xy, None, 410, 256, 9)
    want = cport.ransac_lines(xy, chunk=410, trials=256, pairs=pairs)
    got = emu.emu_ransac_lines(xy, chunk=410, trials=256, pairs=pairs)     # 410 x 256 >= 65,536: default would be kernel 2
    assert_ransac_equal(want, got, "big chunks with pairs")


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1, 3])
def test_gpu_replays_pairs(gpu_engine, cport, kernel):
    import torch
    xy = synth.random_room_scans(300, seed=24)
    npts = np.full(300, 360, np.int32)
    npts[::7] = 301
    npts[5::11] = 2
    pairs, _ = batch_pairs(xy, npts, 100, 100, 77)
    want = cport.ransac_lines(xy, npts=npts, pairs=pairs)
    eng = gpu_engine
    try:
        eng.set_option("ransac.kernel", kernel)
        dev = eng.device
        out = eng.ransac_lines(torch.from_numpy(xy).to(dev), npts=torch.from_numpy(npts).to(dev),
                               pairs=torch.from_numpy(pairs).to(dev))
        torch.cuda.synchronize()
        assert_ransac_equal(want, {k: v.cpu().numpy() for k, v in out.items()}, "gpu kernel %d" % kernel)
    finally:
        eng.set_option("ransac.kernel", 0)


# ---- upstream's own results, when someone with scikit-image has made the golden file ------------------------
needs_gold = pytest.mark.skipif(not os.path.exists(GOLD), reason="tests/golden/skimage_golden.npz absent: run "
                                "tests/golden/make_skimage_golden.py where scikit-image is installed")


@needs_gold
def test_restatement_matches_upstream_golden(cport):
    g = np.load(GOLD)
    for k in range(len(g["n"])):
        n, T = int(g["n"][k]), int(g["trials_drawn"][k])
        data = g["data"][k, :n]
        pairs = g["pairs"][k, :T]
        model, inl, info = rl.ransac_line(data, residual_threshold=float(g["thr"]), max_trials=T,
                                          pairs=(pairs[:, 0], pairs[:, 1]))
        assert np.array_equal(inl, g["inliers"][k, :n]), "case %d inlier mask" % k
        np.testing.assert_allclose(model[0], g["origin"][k], rtol=1e-9, atol=1e-9)
        d = rl.canonical_direction(model[1])
        np.testing.assert_allclose(d, rl.canonical_direction(g["direction"][k]), rtol=1e-9, atol=1e-12)


@needs_gold
@pytest.mark.gpu
def test_gpu_matches_upstream_golden(gpu_engine):
    import torch
    g = np.load(GOLD)
    K, nmax = g["data"].shape[0], g["data"].shape[1]
    T = int(g["pairs"].shape[1])
    pairs = g["pairs"].copy()
    for k in range(K):                    # pad a run upstream left early with repeats of its first pair
        pairs[k, int(g["trials_drawn"][k]):] = pairs[k, 0]
    dev = gpu_engine.device
    out = gpu_engine.ransac_lines(torch.from_numpy(np.ascontiguousarray(g["data"])).to(dev),
                                  npts=torch.from_numpy(g["n"].astype(np.int32)).to(dev), chunk=nmax, trials=T,
                                  thr=float(g["thr"]), pairs=torch.from_numpy(pairs.reshape(K, 1, T, 2).astype(np.int32)).to(dev))
    torch.cuda.synchronize()
    for k in range(K):
        n = int(g["n"][k])
        assert np.array_equal(mask_to_bool(out["inl_mask"][k, 0].cpu().numpy(), n), g["inliers"][k, :n])
        np.testing.assert_allclose(out["line"][k, 0, 2:4].cpu().numpy(), g["origin"][k], rtol=1e-5)
