"""Device-resident step with the scans cut into K slices: line extraction of slice k + 1 on one stream beside association +
filter of slice k on another (ScanPipeline._step_ransac / _step_post).  usage: overlap_probe.py [scans] [K ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lidar_slam_b200 import batch, pipeline  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
Ks = [int(a) for a in sys.argv[2:]] or [1, 2, 4, 8]
eng = batch.Engine(0)
xy = eng.synth_scans(S, seed=1)
pipe = pipeline.ScanPipeline(eng, S)
pipe.install_map_from_lines(*[pipe.step(xy)[k] for k in ("line", "ninl")])
torch.cuda.synchronize()
cur = torch.cuda.current_stream()
s_ran, s_post = torch.cuda.Stream(), torch.cuda.Stream()


def step(K):
    if K == 1:
        pipe.step(xy)
        return
    s_ran.wait_stream(cur); s_post.wait_stream(cur)
    per = (S + K - 1) // K
    for k in range(K):
        a, b = k * per, min(S, (k + 1) * per)
        with torch.cuda.stream(s_ran):
            r = pipe._step_ransac(xy[a:b], a, b)
            ev = torch.cuda.Event(); ev.record()
        with torch.cuda.stream(s_post):
            s_post.wait_event(ev)
            pipe._step_post(r, a, b)
    cur.wait_stream(s_ran); cur.wait_stream(s_post)


for K in Ks:
    for _ in range(3):
        step(K)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step(K)
    e1.record(); torch.cuda.synchronize()
    print("K=%d: %.3f ms per step" % (K, e0.elapsed_time(e1) / 5))
