"""GPU exploration: time lslam_ransac_lines over launch shapes (CUDA events, device-resident
input larger than L2).  Writes gpurun_out/ransac_sweep.json.
usage: ransac_sweep.py [c2|c3|both]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402

KEYS = ("ransac.kernel", "ransac.tpt", "ransac.slots_per_cta", "ransac.scorers", "ransac.posts", "ransac.ring",
        "ransac.threads")


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


def run(eng, combos, fn, phe, tag, rows):
    for combo in combos:
        for k, v in zip(KEYS, combo):
            eng.set_option(k, v)
        try:
            ms = time_call(fn)
        except Exception as exc:
            print(tag, combo, "ERR", exc, flush=True)
            continue
        row = dict(zip([k.split(".")[1] for k in KEYS], combo))
        row.update(ms=ms, alg_tflops=phe * 8 / ms * 1e-9)
        rows.append(row)
        print(tag, row, flush=True)
    for k in KEYS:
        eng.set_option(k, 0)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    eng = batch.Engine(0)
    res = {"fp64_peak_tflops": eng.fp64_peak(30.0)}
    print("fp64 probe TFLOP/s:", res["fp64_peak_tflops"], flush=True)
    if which in ("c2", "both"):
        B = int(os.environ.get("SWEEP_B", 262144))
        xy = eng.synth_scans(B, seed=1)
        out = eng.alloc_ransac_out(B, 360)
        combos = [(1, 0, 0, 0, 0, 0, 0), (2, 0, 0, 0, 0, 0, 0)]
        for tpt in (2, 4):
            for slots in (8, 16, 32):
                for ns, np_ in ((8, 8), (8, 12), (12, 8), (12, 12), (16, 8), (16, 12), (10, 6), (8, 16), (6, 6), (20, 10)):
                    combos.append((2, tpt, slots, ns, np_, 0, 0))
        for ring in (2, 4):
            combos.append((2, 2, 16, 8, 8, ring, 0))
        rows = []
        run(eng, combos, lambda: eng.ransac_lines(xy, out=out), B * 36000.0, "C2", rows)
        res["c2"] = rows
        del xy, out
    if which in ("c3", "both"):
        B3 = 256
        xy3 = eng.synth_scans(B3, beams=8192, seed=2)
        out3 = eng.alloc_ransac_out(B3, 8192, 410)
        combos = [(1, 0, 0, 0, 0, 0, 0), (2, 0, 0, 0, 0, 0, 0)]
        for tpt in (2, 4):
            for slots in (1, 2, 4):
                for ns, np_ in ((8, 8), (12, 4), (16, 4), (16, 8), (20, 4), (24, 4)):
                    combos.append((2, tpt, slots, ns, np_, 0, 0))
        rows3 = []
        run(eng, combos, lambda: eng.ransac_lines(xy3, chunk=410, trials=1024, out=out3), B3 * 8192.0 * 1024, "C3", rows3)
        res["c3"] = rows3
    print(eng.ransac_counters())
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/ransac_sweep.json", "w"), indent=1)


if __name__ == "__main__":
    main()
