"""Run lslam_ransac_lines a few times on one configuration (for ncu captures).
usage: ransac_one.py <c2|c3> [B] [key=value ...]   e.g. ransac_one.py c2 262144 kernel=1 tpt=2"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
eng = batch.Engine(0)
NOSSUM = "nossum" in sys.argv
for kv in [a for a in sys.argv[3:] if a != "nossum"]:
    k, v = kv.split("=")
    eng.set_option("ransac." + k, int(v))
if cfg == "c2":
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 262144
    xy = eng.synth_scans(B, seed=1)
    kw = {}
else:
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    xy = eng.synth_scans(B, beams=8192, seed=2)
    kw = {"chunk": 410, "trials": 1024}
out = eng.alloc_ransac_out(B, xy.shape[1], kw.get("chunk", 100), ssum=not NOSSUM)
for _ in range(3):
    eng.ransac_lines(xy, out=out, **kw)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
ms = []
for _ in range(5):
    e0.record(); eng.ransac_lines(xy, out=out, **kw); e1.record(); torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
print(cfg, sys.argv[3:], B, "ms", min(ms), "median", sorted(ms)[2], eng.ransac_counters())
