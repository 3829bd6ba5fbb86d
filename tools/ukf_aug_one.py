"""Time lslam_ukf_aug_step: F augmented filters with L landmarks in the state (n = 3 + 2L).
usage: ukf_aug_one.py [L] [F] [reps] [pzx: 1 paired (default), 0 literal]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from lidar_slam_b200 import batch  # noqa: E402
from lidar_slam_b200 import synth  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 200
F = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
eng = batch.Engine(0)
if len(sys.argv) > 4:
    eng.set_option("ukf_aug.pzx", int(sys.argv[4]))
dev = eng.device
x, P, u, z = synth.random_aug_filters(min(F, 16), L, seed=0)
rep = (F + len(x) - 1) // len(x)
t = lambda a: torch.from_numpy(np.tile(a, (rep,) + (1,) * (a.ndim - 1))[:F].copy()).to(dev)
st = eng.alloc_ukf_aug(F, L)
x0, P0, ut, zt = t(x), t(P), t(u), t(z)
n = 3 + 2 * L
m, dz = 2 * n + 1, 2 * L
flop = 2.0 * (3 * n ** 3 / 3 + dz ** 3 / 3 + m * n * n + m * dz * dz + m * dz * n + 2 * dz * dz * n + dz * dz * n + dz * n * n)
ms = []
for r in range(reps + 1):
    st["x"].copy_(x0); st["P"].copy_(P0)
    torch.cuda.synchronize()
    l0 = eng.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.ukf_aug_step(st, zt, u=ut)
    e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
    launches = eng.launch_count() - l0
best = min(ms[1:] or ms)
print("ukf_aug L=%d n=%d F=%d: %.2f ms/step (runs %s) -> %.0f filter-steps/s, %.2f TFLOP/s useful fp64 "
      "(%.3f GFLOP/filter), %d launches, workspace %.2f GB, status nonzero %d"
      % (L, n, F, best, ["%.1f" % v for v in ms], F / best * 1e3, flop * F / best / 1e9, flop / 1e9, launches,
         st["workspace"].numel() / 1e9, int((st["status"] != 0).sum().item())))
