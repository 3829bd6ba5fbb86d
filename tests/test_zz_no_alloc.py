"""lslam_ukf_aug_step_static - the augmented-state UKF step with its scratch in the library's static device
array (no cudaMalloc anywhere) - against the existing entry point lslam_ukf_aug_step on the same inputs.

Collected last on purpose.  Uses its own handle so that it behaves the
same alone in a fresh process and after the other tests.  The two entry points run the same kernels on the same
batch shape, so the results are compared bit for bit (the suite's floating-point bar for this filter is 1e-5
relative; equality is inside it)."""
import numpy as np
import pytest

from tests import ukf_aug_util as au

gpu = pytest.mark.gpu

STATIC_BYTES = 32 << 20          # LSLAM_STATIC_SCRATCH_BYTES, include/lslam.h
FLAGS = 7                        # PREDICT | UPDATE | REDRAW


def _device_case(eng, F, L, seed):
    import torch
    x, P, u, z = au.aug_case(F, L, seed=seed)
    to = lambda a: torch.from_numpy(a).to(eng.device)
    return to(x), to(P), to(u), to(z)


def _existing(eng, x, P, u, z, L):
    """lslam_ukf_aug_step with a caller-owned workspace holding all F filters (one batch)."""
    import torch
    F = x.shape[0]
    st = eng.alloc_ukf_aug(F, L, workspace_filters=F)
    st["x"], st["P"] = x.clone(), P.clone()
    eng.ukf_aug_step(st, z, u=u, flags=FLAGS)
    torch.cuda.synchronize()
    return st["x"], st["P"], st["status"]


def _static(eng, x, P, u, z, L):
    """The new entry point through the C ABI: raw device addresses, outputs allocated by the caller beforehand."""
    import torch
    F = x.shape[0]
    xs, Ps = x.clone(), P.clone()
    status = torch.full((F,), -1, dtype=torch.int32, device=eng.device)
    stream = torch.cuda.current_stream(eng.device).cuda_stream
    eng.handle.ukf_aug_step_static(xs.data_ptr(), Ps.data_ptr(), u.data_ptr(), z.data_ptr(), F, L, eng.ukf_params(),
                                   FLAGS, status.data_ptr(), stream)
    torch.cuda.synchronize()
    return xs, Ps, status


@gpu
def test_static_scratch_step_matches_workspace_step_twice_in_a_row():
    import torch
    from lidar_slam_b200 import batch
    eng = batch.Engine(0)
    cases = [(5, 12, 11), (3, 20, 12)]      # (F, L, seed): the second call finds the first call's contents, other layout
    data = [_device_case(eng, F, L, seed) for F, L, seed in cases]
    want = [_existing(eng, *d, L=c[1]) for c, d in zip(cases, data)]
    free0 = torch.cuda.mem_get_info(eng.device)[0]
    got = [_static(eng, *d, L=c[1]) for c, d in zip(cases, data)]         # two calls in a row, one shared array
    # the same case once more: the array now holds the other shape's leftovers
    got.append(_static(eng, *data[0], L=cases[0][1]))
    want.append(want[0])
    inputs = [data[0], data[1], data[0]]
    for k, (w, g) in enumerate(zip(want, got)):
        assert not g[2].any().item() and not w[2].any().item(), "call %d: status" % k
        xw, Pw = w[0].cpu().numpy(), w[1].cpu().numpy()
        xg, Pg = g[0].cpu().numpy(), g[1].cpu().numpy()
        assert np.isfinite(xg).all() and np.isfinite(Pg).all()
        assert not np.array_equal(xg, inputs[k][0].cpu().numpy()), "call %d did nothing" % k
        au.assert_aug_close((xw, Pw), (xg, Pg), "static call %d" % k)       # the suite's 1e-5 bar ...
        assert np.array_equal(xw, xg) and np.array_equal(Pw, Pg), "call %d: not bit-identical" % k    # ... and equality
    # device memory the calls took: none beyond what the (torch-allocated) outputs above account for
    held = sum(t.numel() * t.element_size() for g in got for t in g)
    assert free0 - torch.cuda.mem_get_info(eng.device)[0] <= held + (4 << 20), "the static entry point allocated device memory"


@gpu
def test_static_scratch_rejects_what_does_not_fit():
    import torch
    from lidar_slam_b200 import _lib, batch
    eng = batch.Engine(0)
    L = 200
    assert eng.handle.ukf_aug_workspace_bytes(L, 2) <= STATIC_BYTES < eng.handle.ukf_aug_workspace_bytes(L, 3)
    x, P, u, z = _device_case(eng, 3, L, seed=13)
    status = torch.zeros((3,), dtype=torch.int32, device=eng.device)
    x0 = x.clone()
    with pytest.raises(_lib.LslamError) as e:
        eng.handle.ukf_aug_step_static(x.data_ptr(), P.data_ptr(), u.data_ptr(), z.data_ptr(), 3, L, eng.ukf_params(),
                                       FLAGS, status.data_ptr(), 0)
    assert e.value.code == _lib.LSLAM_ERR_CAPACITY and "LSLAM_STATIC_SCRATCH_BYTES" in str(e.value)
    torch.cuda.synchronize()
    assert torch.equal(x, x0)                                            # nothing ran
    # the largest size that fits (2 filters of n = 403) runs, and matches the existing entry point
    x2, P2, u2, z2 = x[:2].contiguous(), P[:2].contiguous(), u[:2].contiguous(), z[:2].contiguous()
    w = _existing(eng, x2, P2, u2, z2, L)
    g = _static(eng, x2, P2, u2, z2, L)
    assert not g[2].any().item()
    assert torch.equal(w[0], g[0]) and torch.equal(w[1], g[1])


def test_static_scratch_entry_point_on_the_emulated_library(emu):
    """CPU leg: the kernel sources compiled for the CPU (tests/emu) - the static-array entry point gives the caller-workspace
    entry point's results bit for bit on two calls in a row, and rejects a size beyond the array."""
    from lidar_slam_b200 import _lib, batch
    h = emu.emu_handle()
    params = batch.Engine.ukf_params()
    for F, L, seed in [(2, 6, 21), (3, 4, 22)]:
        x, P, u, z = au.aug_case(F, L, seed=seed)
        xw, Pw, sw = emu.emu_ukf_aug_step(x, P, u, z)
        xg, Pg = x.copy(), P.copy()
        status = np.full(F, -99, np.int32)
        h.ukf_aug_step_static(emu.ptr(xg), emu.ptr(Pg), emu.ptr(u), emu.ptr(z), F, L, params, FLAGS, emu.ptr(status))
        assert not status.any() and not sw.any()
        assert np.array_equal(xw, xg) and np.array_equal(Pw, Pg)
    F, L = 4, 200
    assert h.ukf_aug_workspace_bytes(L, F) > STATIC_BYTES
    n = 3 + 2 * L
    x, P, z, status = np.zeros((F, n)), np.zeros((F, n, n)), np.zeros((F, L, 2)), np.zeros(F, np.int32)
    with pytest.raises(_lib.LslamError) as e:
        h.ukf_aug_step_static(emu.ptr(x), emu.ptr(P), None, emu.ptr(z), F, L, params, FLAGS, emu.ptr(status))
    assert e.value.code == _lib.LSLAM_ERR_CAPACITY
