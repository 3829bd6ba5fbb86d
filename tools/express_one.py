"""Time lslam_express_to_cart on B revolutions of 12 packets (360 beams).  usage: express_one.py [B]"""
import sys

import torch

sys.path.insert(0, ".")
from lidar_slam_b200 import batch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
eng = batch.Engine(0)
pk = eng.synth_express(B, 12, seed=0)
xy = torch.empty((B, 360, 2), dtype=torch.float64, device=eng.device)
st = torch.empty(B, dtype=torch.int32, device=eng.device)
for _ in range(3):
    eng.express_to_cart(pk, 360, xy=xy, status=st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    eng.express_to_cart(pk, 360, xy=xy, status=st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
byt = B * (12 * 84 + 360 * 16 + 4)
print("express B=%d ms %.3f -> %.1f M scans/s, %.2f TB/s of algorithmic bytes (%.1f B/point)" %
      (B, ms, B / ms / 1e3, byt / ms / 1e9, byt / B / 360))
