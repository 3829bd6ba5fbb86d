"""Host <-> device copy ceiling of this box: N ranks (torchrun), each with pinned host buffers, run the e2e leg's
copies - H2D 1.06 GB + D2H 0.41 GB per step, concurrently on two streams - and NOTHING else (no kernels).  The
aggregate GB/s over all ranks is what bounds `e2e` at N GPUs.

    python tools/probe/host_copy_probe.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/probe/host_copy_probe.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lidar_slam_b200 import dist as ldist  # noqa: E402


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        ldist.bind_to_gpu_numa(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    S = 1 << 20
    h2d_b, d2h_b = S * 1008, S * 388                                        # bench.py e2e: packets in, results out
    hin = torch.empty(h2d_b, dtype=torch.uint8).pin_memory()
    hout = torch.empty(d2h_b, dtype=torch.uint8).pin_memory()
    din = torch.empty(h2d_b, dtype=torch.uint8, device="cuda")
    dout = torch.empty(d2h_b, dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "both"):
        def step():
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s_in):
                    din.copy_(hin, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s_out):
                    hout.copy_(dout, non_blocking=True)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            step()
        s_in.synchronize(); s_out.synchronize()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        nbytes = (h2d_b if mode != "d2h" else 0) + (d2h_b if mode != "h2d" else 0)
        res[mode] = {"ms_per_step": ms, "aggregate_gbs": world * nbytes / (ms * 1e-3) * 1e-9,
                     "scans_per_s_ceiling": world * S / (ms * 1e-3) if mode == "both" else None}
    if rank == 0:
        print(json.dumps({"ranks": world, "h2d_bytes": h2d_b, "d2h_bytes": d2h_b, "result": res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
