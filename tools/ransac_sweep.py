"""GPU exploration: time lslam_ransac_lines over launch shapes (CUDA events, device-resident
input larger than L2), and probe the fp64 issue rate.  Writes gpurun_out/ransac_sweep.json."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402


def time_call(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    eng = batch.Engine(0)
    res = {"fp64_peak_tflops": [eng.fp64_peak(30.0) for _ in range(3)]}
    print("fp64 probe TFLOP/s:", res["fp64_peak_tflops"], flush=True)
    B = int(os.environ.get("SWEEP_B", 262144))
    xy = eng.synth_scans(B, seed=1)
    out = eng.alloc_ransac_out(B, 360)
    rows = []
    combos = [(0, 0, 0)]
    for tpt in (1, 2, 4):
        for slots in (8, 16, 32, 64):
            for threads in (256, 384, 512):
                combos.append((tpt, slots, threads))
    for tpt, slots, threads in combos:
        eng.set_option("ransac.tpt", tpt)
        eng.set_option("ransac.slots_per_cta", slots)
        eng.set_option("ransac.threads", threads)
        try:
            ms = time_call(lambda: eng.ransac_lines(xy, out=out))
        except Exception as exc:  # capacity errors for some shapes
            print(tpt, slots, threads, "ERR", exc, flush=True)
            continue
        phe = B * 36000.0
        row = {"tpt": tpt, "slots": slots, "threads": threads, "ms": ms, "scans_per_s": B / ms * 1e3,
               "alg_tflops": phe * 8 / ms * 1e-9}
        rows.append(row)
        print(row, flush=True)
    res["c2"] = rows
    for k in ("ransac.tpt", "ransac.slots_per_cta", "ransac.threads"):
        eng.set_option(k, 0)
    # C3: 8192-point scans, chunk 410, 1024 trials
    B3 = 256
    xy3 = eng.synth_scans(B3, beams=8192, seed=2)
    out3 = eng.alloc_ransac_out(B3, 8192, 410)
    rows3 = []
    for tpt, slots, threads in ((0, 0, 0), (4, 1, 256), (4, 2, 512), (2, 1, 512), (2, 1, 256), (2, 2, 512), (2, 4, 512), (1, 1, 512), (1, 2, 512)):
        eng.set_option("ransac.tpt", tpt)
        eng.set_option("ransac.slots_per_cta", slots)
        eng.set_option("ransac.threads", threads)
        ms = time_call(lambda: eng.ransac_lines(xy3, chunk=410, trials=1024, out=out3), iters=3, warm=1)
        phe = B3 * 8192.0 * 1024
        row = {"tpt": tpt, "slots": slots, "threads": threads, "ms": ms, "scans_per_s": B3 / ms * 1e3,
               "alg_tflops": phe * 8 / ms * 1e-9}
        rows3.append(row)
        print("C3", row, flush=True)
    res["c3"] = rows3
    print(eng.ransac_counters())
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/ransac_sweep.json", "w"), indent=1)


if __name__ == "__main__":
    main()
