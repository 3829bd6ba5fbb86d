"""GPU parity of the input stage (lslam_express_to_cart / lslam_synth_express, csrc/express.cuh) against the
oracle that is pinned to the reference's own decoder: angles and distances bit for bit, xy to the
last bits of cos/sin (CUDA's and glibc's differ by an ulp or two)."""
import os

import numpy as np
import pytest

from oracle import cport
from tests.test_oracle_express import GOLD, bits, golden_as_scans

pytestmark = pytest.mark.gpu


def gpu_decode(eng, pk, beams):
    import torch
    xy, meas, st = eng.express_to_cart(torch.from_numpy(np.ascontiguousarray(pk)).to(eng.device), beams, want_meas=True)
    torch.cuda.synchronize()
    return xy.cpu().numpy(), meas.cpu().numpy(), st.cpu().numpy()


@pytest.mark.parametrize("P,beams", [(8, 256), (12, 360), (1, 32), (3, 70), (16, 500)])
def test_golden_packets(gpu_engine, P, beams):
    pk, ang, dist = golden_as_scans(P)
    xy, meas, st = gpu_decode(gpu_engine, pk, beams)
    assert not st.any()
    assert np.array_equal(bits(meas[..., 0]), bits(ang[:, :beams]))
    assert np.array_equal(meas[..., 1], dist[:, :beams])
    want, _, _ = cport.express_to_cart(pk, beams)
    np.testing.assert_allclose(xy, want, rtol=1e-13, atol=1e-9)


def test_corrupt_packets(gpu_engine):
    pk, _, _ = golden_as_scans(4)
    pk = pk.copy()
    pk[1, 2, 40] ^= 0x10
    pk[3, 0, 1] = 0x00
    xy, meas, st = gpu_decode(gpu_engine, pk, 128)
    wxy, wmeas, wst = cport.express_to_cart(pk, 128, want_meas=True)
    assert np.array_equal(st, wst) and st.tolist()[:4] == [0, 2, 0, 1]
    assert np.array_equal(np.isnan(xy), np.isnan(wxy))
    assert np.array_equal(bits(meas), bits(wmeas)) or np.array_equal(np.isnan(meas), np.isnan(wmeas))


def test_synthetic_sensor_large_batch(gpu_engine):
    """Device generator -> device decoder == C oracle on the same bytes, at a size that spans many
    blocks; every packet valid; decoded ranges are integers in the sensor's 14-bit range."""
    import torch
    eng = gpu_engine
    B, P, beams = 20000, 12, 360
    pk = eng.synth_express(B, P, seed=5, scan_base=1000)
    xy, meas, st = eng.express_to_cart(pk, beams, want_meas=True)
    torch.cuda.synchronize()
    assert not st.any().item()
    pk_h = pk.cpu().numpy()
    wxy, wmeas, wst = cport.express_to_cart(pk_h, beams, want_meas=True)
    assert np.array_equal(bits(meas.cpu().numpy()), bits(wmeas))
    np.testing.assert_allclose(xy.cpu().numpy(), wxy, rtol=1e-13, atol=1e-9)
    d = wmeas[..., 1]
    assert (d == np.rint(d)).all() and d.min() >= 0 and d.max() < 16384
    # addressed by scan_base: a later call reproduces a sub-range
    again = eng.synth_express(10, P, seed=5, scan_base=1007)
    assert torch.equal(again, pk[7:17])
    # the scans are fit for the path: RANSAC finds the walls
    r = eng.ransac_lines(xy)
    torch.cuda.synchronize()
    assert (r["ninl"] >= 50).float().mean().item() > 0.8
