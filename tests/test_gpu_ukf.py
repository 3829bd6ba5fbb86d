"""GPU parity of the UKF kernel (csrc/ukf.cuh through the C ABI) against the pinned-order C oracle
(oracle/ukf.c) and the frozen golden vector.  Bar: 1e-5 relative on mean and covariance
(tests/ukf_util.py); the operation order is identical (tests/test_emu_ukf.py shows bit-identity on
the CPU), so what remains is the CUDA math library's sin/cos/atan2 against glibc's."""
import os

import numpy as np
import pytest

from tests import ukf_util as uu

pytestmark = pytest.mark.gpu


def run_gpu(eng, d, zmode=None, valid_min=1, flags=None, params=None, use_valid=True, sigmas=None):
    import torch
    dev = eng.device
    F = len(d["x"])
    st = eng.alloc_ukf(F, d["z"].shape[1] if d.get("z") is not None else 0)
    st["x"] = torch.from_numpy(np.ascontiguousarray(d["x"])).to(dev)
    st["P"] = torch.from_numpy(np.ascontiguousarray(d["P"])).to(dev)
    if params is not None:
        st["params"] = list(params)
    t = lambda k: torch.from_numpy(np.ascontiguousarray(d[k])).to(dev) if d.get(k) is not None else None
    eng.ukf_step(st, t("z"), t("lm_pos"), lm_valid=t("lm_valid") if use_valid else None, valid_min=valid_min,
                 u=t("u"), zmode=zmode, flags=flags, sigmas_f=sigmas)
    torch.cuda.synchronize()
    return st["x"].cpu().numpy(), st["P"].cpu().numpy(), st["status"].cpu().numpy()


@pytest.mark.parametrize("L,F", [(1, 500), (4, 4000), (5, 333), (8, 2000), (13, 100), (16, 64), (32, 50)])
def test_step_matches_oracle(gpu_engine, L, F):
    d = uu.random_filters(F, L, seed=L)
    uu.assert_ukf_close(uu.oracle_step(d), run_gpu(gpu_engine, d), "L %d" % L)


def test_golden_kat7(gpu_engine):
    import torch
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ukf_kat7.npz"))
    eng = gpu_engine
    st = eng.alloc_ukf(1, 8)                         # systemClass.config_ukf state
    u = torch.from_numpy(g["u"][None]).to(eng.device)
    lm = torch.from_numpy(g["lm"][None]).to(eng.device)
    for k in range(5):
        eng.ukf_step(st, torch.from_numpy(g["z"][k].reshape(1, 8, 2)).to(eng.device), lm, u=u)
        torch.cuda.synchronize()
        got = (st["x"].cpu().numpy(), st["P"].cpu().numpy(), st["status"].cpu().numpy())
        uu.assert_ukf_close((g["x"][k][None], g["P"][k][None], np.zeros(1, np.int32)), got, "KAT-7 step %d" % k)
        uu.assert_ukf_close((g["x_mp"][k][None], g["P_mp"][k][None], np.zeros(1, np.int32)), got, "KAT-7 mp %d" % k)


def test_line_records_match_validity_and_status(gpu_engine):
    d = uu.random_filters(3000, 4, seed=5, line_records=True)
    d["lm_valid"] = np.where(d["lm_valid"] > 0, np.arange(4)[None, :], -1).astype(np.int32)
    d["P"][7] = -np.eye(3)
    want = uu.oracle_step(d, zmode=1, valid_min=0)
    got = run_gpu(gpu_engine, d, zmode=1, valid_min=0)
    assert want[2][7] == 1 and want[2].sum() == 1
    uu.assert_ukf_close(want, got, "line records")


def test_predict_and_update_halves(gpu_engine, cport):
    import torch
    eng = gpu_engine
    d = uu.random_filters(1000, 8, seed=9)
    sg_o = np.zeros((1000, 7, 3))
    x1, P1, s1 = cport.ukf_step(d["x"], d["P"], d["u"], None, None, flags=1, sigmas_f=sg_o)     # no redraw
    want = cport.ukf_step(x1, P1, None, d["z"], d["lm_pos"], d["lm_valid"], flags=2, sigmas_f=sg_o)
    sg = torch.zeros((1000, 7, 3), dtype=torch.float64, device=eng.device)
    g1 = run_gpu(eng, {"x": d["x"], "P": d["P"], "u": d["u"]}, flags=1, sigmas=sg)
    uu.assert_ukf_close((x1, P1, s1), g1, "predict half")
    np.testing.assert_allclose(sg.cpu().numpy(), sg_o, rtol=1e-9, atol=1e-9)
    g2 = run_gpu(eng, {"x": g1[0], "P": g1[1], "z": d["z"], "lm_pos": d["lm_pos"], "lm_valid": d["lm_valid"]},
                 flags=2, sigmas=sg)
    uu.assert_ukf_close(want, g2, "update half")


def test_full_size_properties(gpu_engine):
    """1M filters (BASELINE.json configs[4] size): determinism, covariance symmetric-positive and
    shrunk by the update, filters without a valid slot equal a predict-only run."""
    import torch
    eng = gpu_engine
    F, L = 1 << 20, 4
    g = torch.Generator(device="cpu").manual_seed(3)
    lm = (torch.randn((F, L, 2), generator=g, dtype=torch.float64) * 1500).to(eng.device)
    z = torch.stack([lm.norm(dim=-1), torch.atan2(lm[..., 1], lm[..., 0])], -1).contiguous()
    z += 0.01 * torch.randn(z.shape, generator=torch.Generator(device=eng.device).manual_seed(1), device=eng.device,
                            dtype=torch.float64)
    valid = (torch.rand((F, L), generator=g) < 0.7).to(torch.int32).to(eng.device)
    u = torch.ones((F, 2), dtype=torch.float64, device=eng.device)
    a, b, p = eng.alloc_ukf(F, L), eng.alloc_ukf(F, L), eng.alloc_ukf(F, L)
    eng.ukf_step(a, z, lm, lm_valid=valid, u=u)
    eng.ukf_step(b, z, lm, lm_valid=valid, u=u)
    eng.ukf_step(p, None, None, u=u, flags=5)
    torch.cuda.synchronize()
    assert torch.equal(a["x"].view(torch.uint8), b["x"].view(torch.uint8))
    assert torch.equal(a["P"].view(torch.uint8), b["P"].view(torch.uint8))
    assert (a["status"] == 0).all()
    none = valid.sum(1) == 0
    assert none.any() and torch.equal(a["P"][none], p["P"][none]) and torch.equal(a["x"][none], p["x"][none])
    some = ~none
    assert (torch.diagonal(a["P"], dim1=1, dim2=2) > 0).all()
    assert (torch.diagonal(a["P"][some], dim1=1, dim2=2).sum(-1) < torch.diagonal(p["P"][some], dim1=1, dim2=2).sum(-1)).all()
    assert (a["P"] - a["P"].transpose(1, 2)).abs().max() < 1e-9
