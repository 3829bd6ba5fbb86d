"""Time lslam_ransac_ols on B scans of 360 points in 50-point windows (line_detection.py flavour).
usage: ols_one.py [B] [outlier_fraction]"""
import sys

import torch

sys.path.insert(0, ".")
from lidar_slam_b200 import batch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
eng = batch.Engine(0)
pk = eng.synth_express(B, 12, seed=0)
xy, _, _ = eng.express_to_cart(pk, 360)
for _ in range(3):
    r = eng.ransac_ols(xy)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    r = eng.ransac_ols(xy)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
nt = r["n_trials"].float()
ok = (r["ninl"] >= 2)
print("ols B=%d windows=%d ms %.3f -> %.1f M windows/s; mean trials %.1f, mean inliers %.1f, ok %.3f; "
      "point-trial evals/s %.2f G" % (B, B * 8, ms, B * 8 / ms / 1e3, nt.mean().item(),
                                      r["ninl"][ok].float().mean().item(), ok.float().mean().item(),
                                      (nt.sum().item() * 45) / ms / 1e6))
