"""Time lslam_associate on T tables (M = 8 slots, Q = 4 line queries each) with a known map (hits).
usage: assoc_one.py [T]"""
import sys

import torch

sys.path.insert(0, ".")
from lidar_slam_b200 import batch, pipeline  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
eng = batch.Engine(0)
pk = eng.synth_express(T, 12, seed=0)
xy, _, _ = eng.express_to_cart(pk, 360)
pipe = pipeline.ScanPipeline(eng, T)
r = eng.ransac_lines(xy, out=pipe.rout)
pipe.install_map_from_lines(r["line"], r["ninl"])
for _ in range(3):
    eng.associate(r["line"], pipe.tables, q_valid=r["ninl"], out=pipe.aout)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    a = eng.associate(r["line"], pipe.tables, q_valid=r["ninl"], out=pipe.aout)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
byt = T * (2 * 8 * 56 + 8 + 4 * (64 + 4 + 4 + 4 + 16 + 16))
print("assoc T=%d ms %.3f -> %.1f M tables/s, %.2f TB/s of algorithmic bytes, hit rate %.3f" %
      (T, ms, T / ms / 1e3, byt / ms / 1e9, (a["match"] >= 0).float().mean().item()))
