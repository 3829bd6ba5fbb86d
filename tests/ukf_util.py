"""Shared generators / checkers for the UKF tests (oracle vs kernel).

Parity bar (BASELINE.json north_star): UKF mean and covariance within 1e-5 relative.  The
covariance is compared entrywise relative to the largest entry of the matrix (an entry that
cancels to ~0 has no meaningful relative error of its own); the heading is compared as an angle
(absolute 1e-5 rad floor for headings near 0); the position has an absolute floor of 1e-4 mm:
with the reference's alpha = 1e-4 (systemClass.py:20) the sigma weights are -1e8 / +1.7e7, so two
fp64 evaluation orders of sum_k Wm_k h(sigma_k) over ~2 m ranges differ by about
eps * 1e8 * 2000 mm = 4e-5 mm whatever the implementation (measured: 5.7e-6 mm between the literal
fp64 restatement and 60-digit arithmetic, tests/golden/make_ukf_golden.py); 1e-4 mm is 2e-4 of the
0.5 mm range noise the filter is configured with."""
import numpy as np

from oracle import cport, ukf

RTOL = 1e-5
POS_FLOOR_MM = 1e-4


def random_filters(F, L, seed=0, pos_scale=300.0, valid_frac=0.8, line_records=False):
    """F filters: pose around the origin, L landmark slots 0.5-3 m away, measurements = hx(truth)
    + noise, a random subset of slots valid.  Returns dict of arrays."""
    rng = np.random.default_rng(seed)
    x = np.stack([rng.normal(0, pos_scale, F), rng.normal(0, pos_scale, F), rng.uniform(-np.pi, np.pi, F)], -1)
    A = rng.normal(size=(F, 3, 3)) * np.array([0.2, 0.2, 0.1])[None, :, None]
    P = A @ A.transpose(0, 2, 1) + np.diag([0.1, 0.1, 0.05])[None]
    u = rng.uniform(0.5, 1.5, (F, 2))
    ang = rng.uniform(-np.pi, np.pi, (F, L))
    rad = rng.uniform(500, 3000, (F, L))
    lm = x[:, None, :2] + np.stack([rad * np.cos(ang), rad * np.sin(ang)], -1)
    z = np.zeros((F, L, 2))
    for f in range(F):
        z[f] = ukf.transfer_function(x[f], [ukf._Pos(*p) for p in lm[f]]).reshape(L, 2)
    z += rng.normal(size=(F, L, 2)) * np.array([0.3, 0.02])
    valid = (rng.random((F, L)) < valid_frac).astype(np.int32)
    out = {"x": x, "P": P, "u": u, "z": z, "lm_pos": lm, "lm_valid": valid}
    if line_records:
        # zmode 1: slot = a line record whose origin (ox, oy) is the measured point in the robot frame
        rec = rng.normal(size=(F, L, 8))
        rec[..., 2] = z[..., 0] * np.cos(z[..., 1])
        rec[..., 3] = z[..., 0] * np.sin(z[..., 1])
        out["z"] = rec
    return out


def assert_ukf_close(want, got, what="", rtol=RTOL):
    wx, wP, ws = want
    gx, gP, gs = got
    assert np.array_equal(ws, gs), "%s status differs" % what
    assert np.allclose(gx[:, :2], wx[:, :2], rtol=rtol, atol=POS_FLOOR_MM), "%s position differs: %g" % (
        what, np.abs(gx[:, :2] - wx[:, :2]).max())
    dth = np.abs((gx[:, 2] - wx[:, 2] + np.pi) % (2 * np.pi) - np.pi)
    assert (dth <= rtol * np.maximum(np.abs(wx[:, 2]), 1.0)).all(), "%s heading differs: %g" % (what, dth.max())
    scale = np.abs(wP).reshape(len(wP), -1).max(axis=1)[:, None, None]
    err = np.abs(gP - wP) / scale
    assert (err <= rtol).all(), "%s covariance differs: %g of the largest entry" % (what, err.max())


def oracle_step(d, zmode=0, valid_min=1, flags=7, params=None, use_valid=True):
    return cport.ukf_step(d["x"], d["P"], d["u"], d["z"], d["lm_pos"], d["lm_valid"] if use_valid else None,
                          valid_min=valid_min, zmode=zmode, params=params, flags=flags)


def check_small_batch(eng, F=16, L=8, seed=0):
    """GPU (torch Engine) vs the C oracle on a small batch; used by __graft_entry__.smoke()."""
    import torch
    d = random_filters(F, L, seed=seed)
    want = oracle_step(d)
    dev = eng.device
    st = eng.alloc_ukf(F, L)
    st["x"] = torch.from_numpy(d["x"]).to(dev)
    st["P"] = torch.from_numpy(d["P"]).to(dev)
    eng.ukf_step(st, torch.from_numpy(d["z"]).to(dev), torch.from_numpy(d["lm_pos"]).to(dev),
                 lm_valid=torch.from_numpy(d["lm_valid"]).to(dev), u=torch.from_numpy(d["u"]).to(dev))
    torch.cuda.synchronize()
    assert_ukf_close(want, (st["x"].cpu().numpy(), st["P"].cpu().numpy(), st["status"].cpu().numpy()), "smoke UKF")
