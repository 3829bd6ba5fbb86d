import sys, ctypes, torch
sys.path.insert(0, '/root/repo')
from lidar_slam_b200 import batch
eng = batch.Engine(0)
for kv in sys.argv[1:]:
    k, v = kv.split("="); eng.set_option("ransac." + k, int(v))
B = 262144
xy = eng.synth_scans(B, seed=1)
out = eng.alloc_ransac_out(B, 360, 100, ssum=False)
for _ in range(2): eng.ransac_lines(xy, out=out)
buf = (ctypes.c_longlong * 8)()
eng.handle.cdll.lslam_debug_counters(eng.handle._h, buf)
eng.ransac_lines(xy, out=out)
eng.handle.cdll.lslam_debug_counters(eng.handle._h, buf)
v = list(buf)
names = {1: "start -> S1 (stage)", 2: "wait S1", 3: "score work", 4: "wait S2", 5: "cand work", 6: "wait S3", 7: "post work"}
tot = sum(v[1:8])
for k in range(1, 8): print("%-26s %6.2f%%  %.0f cycles/warp/tile" % (names[k], 100 * v[k] / tot, v[k] / (B * 4 / 20 * 16)))
