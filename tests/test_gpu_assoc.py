"""GPU parity of the association kernel (csrc/assoc.cuh through the C ABI) against
oracle/assoc.py, itself pinned to the real landmarking.py.  Bit-exact: match indices, ids,
projection coefficients and the whole updated landmark table."""
import numpy as np
import pytest

from tests import assoc_util as au

pytestmark = pytest.mark.gpu


def run_gpu(eng, queries, tables, valid=None, **kw):
    import torch
    dev = eng.device
    tt = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in tables.items()}
    q = torch.from_numpy(np.ascontiguousarray(queries)).to(dev)
    v = torch.from_numpy(np.ascontiguousarray(valid)).to(dev) if valid is not None else None
    out = eng.associate(q, tt, q_valid=v, **kw)
    torch.cuda.synchronize()
    return (out["match"].cpu().numpy(), out["match_id"].cpu().numpy(), out["proj_ab"].cpu().numpy(),
            {k: t.cpu().numpy() for k, t in tt.items()})


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("T,M,prefill,Q", [(300, 8, 0, 12), (300, 8, 8, 6), (301, 5, 4, 9), (203, 16, 13, 9), (99, 12, 12, 7),
                                           (64, 40, 37, 10), (40, 70, 66, 5),
                                           (17, 200, 190, 8),
                                           (9, 200, 150, 200)])          # BASELINE.json configs[3]: 200 queries vs 200 stored
def test_first_match_modes(gpu_engine, mode, T, M, prefill, Q):
    queries, valid, tables = au.random_streams(T, Q, M, seed=M + mode, prefill=prefill)
    want = au.oracle_associate(queries, valid, tables, mode=mode)
    got = run_gpu(gpu_engine, queries, tables, valid, mode=mode)
    au.assert_assoc_equal(want, got, "mode %d M %d" % (mode, M))
    assert (want[0] >= 0).any()


@pytest.mark.parametrize("mode", [0, 1])
def test_gate_decided_by_the_last_bit(gpu_engine, mode):
    """As tests/test_emu_assoc.py: gate distances within a few ulps of the tolerance and exactly on it (the kernel compares
    squared sums with assoc_gate_sq's bound instead of taking square roots)."""
    queries, valid, tables = au.gate_boundary_streams(4096, seed=5)
    want = au.oracle_associate(queries, valid, tables, mode=mode)
    got = run_gpu(gpu_engine, queries, tables, valid, mode=mode)
    au.assert_assoc_equal(want, got, "gate mode %d" % mode)
    assert (want[0] >= 0).sum() > 500 and (want[0] < 0).sum() > 500


def _misaligned_dev(a, dev):
    """Device copy of `a` that starts one element past a 16-byte boundary."""
    import torch
    a = np.ascontiguousarray(a)
    flat = torch.zeros(a.size + 4, dtype=torch.from_numpy(a).dtype, device=dev)
    v = flat[1:1 + a.size].view(a.shape)
    v.copy_(torch.from_numpy(a))
    assert v.data_ptr() % 16 == a.itemsize
    return v


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("T,M,prefill,Q,what", [(257, 8, 6, 7, "misaligned"), (130, 12, 9, 5, "misaligned"),
                                                (33, 8, 6, 400, "unstaged"), (33, 8, 6, 400, "unstaged-misaligned")])
def test_load_paths(gpu_engine, mode, T, M, prefill, Q, what):
    """As tests/test_emu_assoc.py::test_load_paths: plain loads for buffers off the 16-byte grid, queries read from global
    memory when they do not fit beside the table - same answers as the asynchronous, staged path."""
    import torch
    eng = gpu_engine
    queries, valid, tables = au.random_streams(T, Q, M, seed=3 * M + mode, prefill=prefill)
    want = au.oracle_associate(queries, valid, tables, mode=mode)
    if "misaligned" in what:
        tt = {k: (_misaligned_dev(v, eng.device) if k in ("rows", "life", "ids") else torch.from_numpy(v.copy()).to(eng.device))
              for k, v in tables.items()}
        q = _misaligned_dev(queries, eng.device)
    else:
        tt = {k: torch.from_numpy(v.copy()).to(eng.device) for k, v in tables.items()}
        q = torch.from_numpy(queries).to(eng.device)
    out = eng.associate(q, tt, q_valid=torch.from_numpy(valid).to(eng.device), mode=mode)
    torch.cuda.synchronize()
    got = (out["match"].cpu().numpy(), out["match_id"].cpu().numpy(), out["proj_ab"].cpu().numpy(),
           {k: t.cpu().numpy() for k, t in tt.items()})
    au.assert_assoc_equal(want, got, "%s mode %d" % (what, mode))


def test_line_record_layout_chained_after_ransac(gpu_engine):
    """layout 1: the queries are the `line` records of lslam_ransac_lines, q_valid its ninl."""
    import torch
    from lidar_slam_b200 import synth
    eng = gpu_engine
    xy = synth.random_room_scans(64, seed=8)
    xy = np.concatenate([xy, xy], axis=1)           # two revolutions per stream: second one re-observes
    r = eng.ransac_lines(torch.from_numpy(xy).to(eng.device), chunk=100)
    tables = eng.alloc_tables(64, 8)
    out = eng.associate(r["line"], tables, q_valid=r["ninl"])
    torch.cuda.synchronize()
    line = r["line"].cpu().numpy()
    ninl = r["ninl"].cpu().numpy()
    zero = {"rows": np.zeros((64, 8, 6)), "life": np.zeros((64, 8), np.int32), "ids": np.zeros((64, 8), np.int32),
            "count": np.zeros(64, np.int32), "next_id": np.zeros(64, np.int32)}
    want = au.oracle_associate(np.ascontiguousarray(line[..., [0, 1, 2, 3, 6, 7]]), ninl, zero)
    got = (out["match"].cpu().numpy(), out["match_id"].cpu().numpy(), out["proj_ab"].cpu().numpy(),
           {k: t.cpu().numpy() for k, t in tables.items()})
    au.assert_assoc_equal(want, got, "chained")
    assert (got[0] >= 0).sum() > 20                 # the second revolution hits landmarks of the first


def test_large_batch_properties(gpu_engine):
    """Full-size property checks (100k streams): determinism; ids of appended landmarks are the
    extraction counter; live count never exceeds capacity; hits point inside the live table."""
    import torch
    eng = gpu_engine
    T, Q, M = 100000, 8, 8
    g = torch.Generator(device="cpu").manual_seed(1)
    sc = torch.tensor([0.5, 500, 500, 500], dtype=torch.float64)
    base = torch.randn((T, 1, 4), generator=g, dtype=torch.float64) * sc
    ab_pos = base + torch.randn((T, Q, 4), generator=g, dtype=torch.float64) * torch.tensor(
        [0.03, 3, 10, 10], dtype=torch.float64)
    end = ab_pos[..., 2:4] + 30 * torch.randn((T, Q, 2), generator=g, dtype=torch.float64)
    q = torch.cat([ab_pos, end], dim=-1).contiguous().to(eng.device)      # tip within the 100 mm gate
    t1, t2 = eng.alloc_tables(T, M), eng.alloc_tables(T, M)
    o1 = eng.associate(q, t1)
    o2 = eng.associate(q, t2)
    torch.cuda.synchronize()
    for k in o1:
        assert torch.equal(o1[k].view(torch.uint8), o2[k].view(torch.uint8)), k
    for k in t1:
        assert torch.equal(t1[k].view(torch.uint8), t2[k].view(torch.uint8)), k
    assert (t1["count"] <= M).all() and (t1["count"] >= 1).all()
    assert (t1["next_id"] == Q).all()
    m = o1["match"]
    assert ((m >= -1) | (m == -3)).all()
    assert (m < t1["count"][:, None]).all()
    assert (m >= 0).float().mean() > 0.3
