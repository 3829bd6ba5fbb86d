"""Time lslam_ransac_ols (line_detection.py flavour) on every 50-point window of B synthetic scans.  usage: ols_one.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
eng = batch.Engine(0)
xy = eng.synth_scans(B, seed=6)
ms = []
for r in range(8):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    o = eng.ransac_ols(xy, window=50, thr=3.0)
    e1.record()
    torch.cuda.synchronize()
    if r >= 3:
        ms.append(e0.elapsed_time(e1))
print("ols B=%d windows=%d: min %.3f ms median %.3f ms, mean trials %.2f, %.1f M windows/s" % (
    B, B * 8, min(ms), sorted(ms)[len(ms) // 2], float(o["n_trials"].float().mean().item()), B * 8 / min(ms) * 1e-3))
