"""Every kernel of liblslam.so once, at small sizes, for compute-sanitizer (memcheck / racecheck / synccheck / initcheck):

    compute-sanitizer --tool racecheck python tools/sanitize_cases.py            # tools/sanitize.sh runs all tools

Results are also compared with the C oracle where that is cheap, so a run under the sanitizer is still a parity run."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch, pipeline, synth  # noqa: E402
from oracle import cport  # noqa: E402
from tests import ukf_aug_util as au  # noqa: E402
from tests.parity_util import assert_ransac_equal  # noqa: E402

only = set(sys.argv[1:])
eng = batch.Engine(0)
dev = eng.device


def want(name):
    return not only or name in only


def ransac(kernel, xy, npts, **kw):
    eng.set_option("ransac.kernel", kernel)
    r = eng.ransac_lines(torch.from_numpy(xy).to(dev), npts=None if npts is None else torch.from_numpy(npts).to(dev), **kw)
    torch.cuda.synchronize()
    eng.set_option("ransac.kernel", 0)
    return {k: v.cpu().numpy() for k, v in r.items()}


xy = synth.random_room_scans(24, seed=1)
npts = np.full(24, 360, np.int32)
npts[3] = 301
npts[7] = 2
if want("ransac"):
    w = cport.ransac_lines(xy, npts=npts, seed=3)
    for kernel in (1, 2, 3):
        assert_ransac_equal(w, ransac(kernel, xy, npts, seed=3), "kernel %d" % kernel)
    big = synth.random_room_scans(2, beams=820, seed=2)
    w = cport.ransac_lines(big, chunk=410, trials=256, seed=1)
    for kernel in (1, 2, 3):
        assert_ransac_equal(w, ransac(kernel, big, None, chunk=410, trials=256, seed=1), "big kernel %d" % kernel)
    print("ransac ok", flush=True)

if want("pipeline"):
    S = 96
    pk = eng.synth_express(S, 12, seed=5)
    xyd, _, st = eng.express_to_cart(pk, 360)
    p = pipeline.ScanPipeline(eng, S, seed=2)
    r0 = eng.ransac_lines(xyd, seed=2)
    p.install_map_from_lines(r0["line"], r0["ninl"])
    p.ukf["u"].fill_(1.0)
    for _ in range(2):
        res = p.step(xyd)
    pts, off = p.project(xyd, res)
    h = p.alloc_host_io("packets", 12)
    h.copy_(pk)
    out = p.step_host(h, slices=3, points=True)
    torch.cuda.synchronize()
    assert int(p.ukf["status"].abs().sum()) == 0
    print("pipeline (express, ransac, associate, ukf L=4, project) ok", flush=True)

if want("ukf"):
    from tests import ukf_util
    for L in (1, 8, 16, 32):
        ukf_util.check_small_batch(eng, F=12, L=L, seed=L)
    print("ukf ok", flush=True)

if want("assoc"):
    T, M, Q = 40, 200, 50
    tb = eng.alloc_tables(T, M)
    g = torch.Generator().manual_seed(1)
    q = (torch.randn((T, Q, 6), generator=g, dtype=torch.float64) * 100).to(dev)
    for mode in (0, 1, 2):
        eng.associate(q, tb, mode=mode)
    torch.cuda.synchronize()
    print("associate ok", flush=True)

if want("ols"):
    o = eng.ransac_ols(torch.from_numpy(xy).to(dev), window=50, thr=3.0)
    torch.cuda.synchronize()
    print("ols ok", flush=True)

if want("aug"):
    for L, F in ((5, 6), (40, 3)):
        x, P, u, z = au.aug_case(F, L, seed=L)
        wx = au.oracle_aug(x, P, u, z)
        cond = au.oracle_conditioning(x, P, u, z, wx)
        st = eng.alloc_ukf_aug(F, L, workspace_filters=2)
        st["x"], st["P"] = torch.from_numpy(x).to(dev), torch.from_numpy(P).to(dev)
        eng.ukf_aug_step(st, torch.from_numpy(z).to(dev), u=torch.from_numpy(u).to(dev))
        torch.cuda.synchronize()
        au.assert_aug_close(wx, (st["x"].cpu().numpy(), st["P"].cpu().numpy()), "aug L=%d" % L, cond=cond)
    print("aug ok", flush=True)
print("all cases ok")
