"""GPU exploration: lslam_ransac_lines on C2 (360-beam scans, chunk 100, 100 trials) over the tile kernel's launch shapes
against the round-1 phase-structured kernel.  usage: tile_sweep.py [B]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_slam_b200 import batch  # noqa: E402
from tools.ransac_sweep import time_call  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    eng = batch.Engine(0)
    peak = eng.fp64_peak(30.0)
    xy = eng.synth_scans(B, seed=1)
    phe = B * 360 * 100
    rows = []
    for ssum in (False, True):
        out = eng.alloc_ransac_out(B, 360, 100, ssum=ssum)
        combos = [(1, 0, 0, 0)] + [(3, tpt, thr, g) for tpt in (2, 4, 6, 8) for thr in (128, 256, 384, 512) for g in (0,)]
        for kernel, tpt, threads, slots in combos:
            for k, v in (("ransac.kernel", kernel), ("ransac.tpt", tpt), ("ransac.threads", threads), ("ransac.slots_per_cta", slots)):
                eng.set_option(k, v)
            try:
                ms = time_call(lambda: eng.ransac_lines(xy, out=out))
            except Exception as exc:
                print("ERR", kernel, tpt, threads, exc, flush=True)
                continue
            row = dict(kernel=kernel, tpt=tpt, threads=threads, slots=slots, ssum=ssum, ms=ms, alg_tflops=phe * 8 / ms * 1e-9,
                       frac=phe * 8 / ms * 1e-9 / peak)
            rows.append(row)
            print(row, flush=True)
    print("counters", eng.ransac_counters())
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"fp64_peak_tflops": peak, "B": B, "rows": rows}, open("gpurun_out/tile_sweep.json", "w"), indent=1)


if __name__ == "__main__":
    main()
