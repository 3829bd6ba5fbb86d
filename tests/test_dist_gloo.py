"""N > 1 host logic on the CPU: world_size 2 over gloo.  Each rank runs its shard of the scans
(scan_base = first scan of the shard) through the CPU pipeline, the results are gathered with
lidar_slam_b200.dist.ResultGather, and the gathered whole must equal a single-process run - i.e. the
sharding is invisible in the results (Philox streams are keyed by the GLOBAL scan index)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lidar_slam_b200 import dist as ldist, synth
from tests import pipeline_util as pu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = ldist.shard_range(total, rank, world)
        xy = synth.random_room_scans(total, seed=31)[lo:hi]
        out = pu.cpu_pipeline(xy, 2, u=np.ones((hi - lo, 2)), seed=5, scan_base=lo)[-1]
        res = {"x": torch.from_numpy(out["x"]), "P": torch.from_numpy(out["P"]),
               "match_id": torch.from_numpy(out["assoc"]["match_id"])}
        g = ldist.ResultGather()(res)
        if rank == 0:
            np.savez(os.path.join(outdir, "gathered.npz"), **{k: v.numpy() for k, v in g.items()})
    finally:
        dist.destroy_process_group()


def test_shard_range():
    assert [ldist.shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert ldist.shard_range(2, 3, 4) == (2, 2)


@pytest.mark.timeout(300)
def test_two_ranks_equal_one(tmp_path):
    total, world = 64, 2
    mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
    g = np.load(tmp_path / "gathered.npz")
    xy = synth.random_room_scans(total, seed=31)
    # single process: same map set-up per scan (the clutter rows come from the neighbour scan, so build
    # the reference shard-wise exactly as the ranks do, but in one process without any collective)
    parts = []
    for r in range(world):
        lo, hi = ldist.shard_range(total, r, world)
        parts.append(pu.cpu_pipeline(xy[lo:hi], 2, u=np.ones((hi - lo, 2)), seed=5, scan_base=lo)[-1])
    for r in range(world):
        assert np.array_equal(g["x"][r], parts[r]["x"]) and np.array_equal(g["P"][r], parts[r]["P"])
        assert np.array_equal(g["match_id"][r], parts[r]["assoc"]["match_id"])
    # and the RANSAC stage is invariant to the split itself
    whole = pu.cport.ransac_lines(xy, seed=5, scan_base=0)
    lo, hi = ldist.shard_range(total, 1, world)
    half = pu.cport.ransac_lines(xy[lo:hi], seed=5, scan_base=lo)
    for k in ("ninl", "win_trial", "win_pair"):
        assert np.array_equal(whole[k][lo:hi], half[k])
