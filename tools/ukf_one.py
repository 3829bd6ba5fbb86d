"""Run the pipeline's UKF step a few times on 1M filters (for ncu captures / timing).
usage: ukf_one.py [F] [L]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
L = int(sys.argv[2]) if len(sys.argv) > 2 else 4
eng = batch.Engine(0)
g = torch.Generator(device="cpu").manual_seed(3)
lm = (torch.randn((F, L, 2), generator=g, dtype=torch.float64) * 1500).to(eng.device)
z = torch.stack([lm.norm(dim=-1), torch.atan2(lm[..., 1], lm[..., 0])], -1).contiguous()
valid = torch.ones((F, L), dtype=torch.int32, device=eng.device)
st = eng.alloc_ukf(F, L)
st["u"].fill_(1.0)
for _ in range(3):
    eng.ukf_step(st, z, lm, lm_valid=valid)
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); eng.ukf_step(st, z, lm, lm_valid=valid); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("ukf F=%d L=%d ms %.4f -> %.1f M filter-steps/s, status sum %d" % (F, L, ms, F / ms * 1e-3, int(st["status"].sum())))
