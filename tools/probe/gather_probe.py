"""Diagnostic (torchrun): where does the time of step + gather go at N > 1?"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from lidar_slam_b200 import batch, pipeline, dist as ldist  # noqa: E402

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
S = 1 << 20
eng = batch.Engine(local)
pipe = pipeline.ScanPipeline(eng, S, seed=0, scan_base=rank * S)
xy = eng.synth_scans(S, seed=0, scan_base=rank * S)
r0 = eng.ransac_lines(xy, seed=0, scan_base=rank * S, out=pipe.rout)
pipe.install_map_from_lines(r0["line"], r0["ninl"])
g = ldist.ResultGather()


def timed(name, fn, n=3):
    for _ in range(2):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    t_issue = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    if rank == 0:
        print("%-28s events %.2f ms/iter, host issue %.2f ms/iter, wall %.2f ms/iter" %
              (name, e0.elapsed_time(e1) / n, t_issue / n * 1e3, t_all / n * 1e3), flush=True)


res = pipe.step(xy)
timed("step only", lambda: pipe.step(xy))
timed("gather only", lambda: g(res))
timed("step + gather", lambda: g(pipe.step(xy)))
timed("step + gather(x only)", lambda: ldist.ResultGather(("x",))(pipe.step(xy)))
dist.destroy_process_group()
