"""GPU parity of the whole step (lidar_slam_b200/pipeline.py: RANSAC -> association -> UKF through
the C ABI) against the C pipeline on the same seeded scans, several steps with persistent tables
and filters, device-resident and through the host-buffer path."""
import numpy as np
import pytest

from lidar_slam_b200 import synth
from tests import pipeline_util as pu, ukf_util as uu
from tests.parity_util import assert_ransac_equal

pytestmark = pytest.mark.gpu


def test_pipeline_steps_match_cpu(gpu_engine):
    import torch
    from lidar_slam_b200 import pipeline
    eng = gpu_engine
    S, M, steps = 2048, 8, 3
    xy = synth.random_room_scans(S, seed=12)
    u = np.full((S, 2), 1.0)
    want = pu.cpu_pipeline(xy, steps, M=M, u=u, seed=3, scan_base=77)
    pipe = pipeline.ScanPipeline(eng, S, table_capacity=M, seed=3, scan_base=77)
    pipe.ukf["u"].fill_(1.0)
    xy_d = torch.from_numpy(xy).to(eng.device)
    r = eng.ransac_lines(xy_d, seed=3, scan_base=77)
    torch.cuda.synchronize()
    assert_ransac_equal(want[0]["ransac"], {k: v.cpu().numpy() for k, v in r.items()}, "pipeline RANSAC")
    # chain both sides from the oracle's line records for the map (they agree to 1e-5, not bitwise)
    line_o = torch.from_numpy(want[0]["ransac"]["line"]).to(eng.device)
    pipe.install_map_from_lines(line_o, r["ninl"])
    for k in range(steps):
        res = pipe.step(xy_d)
        torch.cuda.synchronize()
        m = res["match"].cpu().numpy()
        assert np.array_equal(m, want[k]["assoc"]["match"]), "step %d match" % k
        assert np.array_equal(res["match_id"].cpu().numpy(), want[k]["assoc"]["match_id"])
        assert np.array_equal(pipe.tables["life"].cpu().numpy(), want[k]["tables"]["life"])
        uu.assert_ukf_close((want[k]["x"], want[k]["P"], want[k]["status"]),
                            (res["x"].cpu().numpy(), res["P"].cpu().numpy(), pipe.ukf["status"].cpu().numpy()),
                            "pipeline step %d" % k)
    assert (m >= 0).mean() > 0.9


def test_host_path_equals_device_path(gpu_engine):
    import torch
    from lidar_slam_b200 import pipeline
    eng = gpu_engine
    S = 4096
    xy = eng.synth_scans(S, seed=21)
    a = pipeline.ScanPipeline(eng, S, seed=1)
    b = pipeline.ScanPipeline(eng, S, seed=1)
    r0 = eng.ransac_lines(xy, seed=1)
    a.install_map_from_lines(r0["line"], r0["ninl"])
    b.install_map_from_lines(r0["line"], r0["ninl"])
    xy_h = b.alloc_host_io()
    xy_h.copy_(xy)
    torch.cuda.synchronize()
    for _ in range(2):
        ra = a.step(xy)
        hb = b.step_host(xy_h)
        torch.cuda.synchronize()
        for k in b.result_keys():
            assert torch.equal(ra[k].cpu().view(torch.uint8), hb[k].view(torch.uint8)), k
    h2d, d2h = b.io_bytes()
    assert h2d == xy_h.numel() * 8 and d2h == sum(hb[k].numel() * hb[k].element_size() for k in b.result_keys())


@pytest.mark.parametrize("kind,slices", [("xy", 4), ("packets", 1), ("packets", 5)])
def test_overlapped_and_packet_host_paths(gpu_engine, kind, slices):
    """Sliced three-stream host path and the raw-packet input give exactly the single-stream results."""
    import torch
    from lidar_slam_b200 import pipeline
    eng = gpu_engine
    S = 3000                      # not a multiple of the slice count
    pk = eng.synth_express(S, 12, seed=8)
    xy, _, st = eng.express_to_cart(pk, 360)
    assert not st.any().item()
    a = pipeline.ScanPipeline(eng, S, seed=2, scan_base=5)
    b = pipeline.ScanPipeline(eng, S, seed=2, scan_base=5)
    r0 = eng.ransac_lines(xy, seed=2, scan_base=5)
    a.install_map_from_lines(r0["line"], r0["ninl"])
    b.install_map_from_lines(r0["line"], r0["ninl"])
    h_in = b.alloc_host_io(kind, 12)
    h_in.copy_(pk if kind == "packets" else xy)
    torch.cuda.synchronize()
    for _ in range(3):
        ra = a.step(xy)
        hb = b.step_host(h_in, slices=slices)
        torch.cuda.synchronize()
        for k in b.result_keys():
            assert torch.equal(ra[k].cpu().view(torch.uint8), hb[k].view(torch.uint8)), k
        if kind == "packets":
            assert not hb["status"].any().item()
    assert torch.equal(a.tables["life"], b.tables["life"])
    h2d, d2h = b.io_bytes(h_in)
    assert h2d == h_in.numel() * h_in.element_size()


@pytest.mark.parametrize("kind", ["xy", "packets"])
def test_pipelined_host_steps(gpu_engine, kind):
    """step_host(wait=False): steps queued back to back (step k + 1 before the host has read step k) give, step for
    step, the results of the device-resident path; a step's pinned results stay valid until the call after the next."""
    import torch
    from lidar_slam_b200 import pipeline
    eng = gpu_engine
    S, steps = 3000, 5
    pk = eng.synth_express(S, 12, seed=9)
    xy, _, st = eng.express_to_cart(pk, 360)
    assert not st.any().item()
    a = pipeline.ScanPipeline(eng, S, seed=4, scan_base=11)
    b = pipeline.ScanPipeline(eng, S, seed=4, scan_base=11)
    r0 = eng.ransac_lines(xy, seed=4, scan_base=11)
    a.install_map_from_lines(r0["line"], r0["ninl"])
    b.install_map_from_lines(r0["line"], r0["ninl"])
    h_in = b.alloc_host_io(kind, 12)
    h_in.copy_(pk if kind == "packets" else xy)
    want = []
    for _ in range(steps):
        ra = a.step(xy)
        torch.cuda.synchronize()
        want.append({k: ra[k].cpu().clone() for k in b.result_keys()})
    torch.cuda.synchronize()

    def check(i, res):
        res["done"].synchronize()
        for k in b.result_keys():
            assert torch.equal(want[i][k].view(torch.uint8), res[k].view(torch.uint8)), "step %d %s" % (i, k)

    pending = None
    for i in range(steps):
        res = b.step_host(h_in, slices=4, wait=False)
        if pending is not None:
            check(i - 1, pending)
        pending = res
    check(steps - 1, pending)
    b.drain()
    torch.cuda.synchronize()
    assert torch.equal(a.tables["life"], b.tables["life"])
    with pytest.raises(ValueError):
        b.step_host(h_in, slices=1, wait=False)


@pytest.mark.parametrize("slices", [1, 3])
def test_host_step_returns_the_projected_inlier_points(gpu_engine, slices):
    """step_host(points=True): what landmark_extraction itself returns (ransac_functions.py:28,46-55) for every chunk of
    every stream - x of the inliers, y = a' x + b' with the MATCHED stored landmark's coefficients - against the oracle
    projection of the same masks, bit for bit, in (stream, chunk, data) order."""
    import torch
    from lidar_slam_b200 import pipeline
    from oracle import ransac_line as rl
    from tests.parity_util import mask_to_bool
    eng = gpu_engine
    S = 500
    xy = eng.synth_scans(S, seed=31)
    p = pipeline.ScanPipeline(eng, S, seed=3)
    r0 = eng.ransac_lines(xy, seed=3)
    p.install_map_from_lines(r0["line"], r0["ninl"])
    h_in = p.alloc_host_io()
    h_in.copy_(xy)
    torch.cuda.synchronize()
    for _ in range(2):
        out = p.step_host(h_in, slices=slices, points=True)
        torch.cuda.synchronize()
        got = torch.cat(list(out["points"]), 0).numpy()
        ninl = out["ninl"].numpy()
        masks = p.rout["inl_mask"].cpu().numpy().view(np.uint32)
        ab = p.aout["proj_ab"].cpu().numpy()
        xyh = h_in.numpy()
        parts = []
        for s in range(S):
            for c in range(4):
                if ninl[s, c] < 2:
                    continue
                a0, a1 = c * 100, min(360, c * 100 + 100)
                parts.append(rl.project_inliers(xyh[s, a0:a1], mask_to_bool(masks[s, c], a1 - a0), ab[s, c, 0], ab[s, c, 1]))
        want = np.concatenate(parts, 0)
        assert got.shape == want.shape and np.array_equal(got.view(np.uint64), want.view(np.uint64))
        assert (out["match_id"].numpy() >= 0).mean() > 0.9           # the stored landmarks' coefficients were used
