"""Diagnostic: all_gather bandwidth and transport between the GPUs of one box (run under torchrun)."""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if rank == 0:
    print("peer access 0->1:", torch.cuda.can_device_access_peer(0, 1) if torch.cuda.device_count() > 1 else None,
          "devices visible:", torch.cuda.device_count())
for mb in (1, 16, 128, 512):
    n = mb * (1 << 20) // 8
    src = torch.full((n,), float(rank), dtype=torch.float64, device="cuda")
    dst = torch.empty((world, n), dtype=torch.float64, device="cuda")
    for _ in range(3):
        dist.all_gather_into_tensor(dst, src)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        dist.all_gather_into_tensor(dst, src)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    if rank == 0:
        print("all_gather %4d MB per rank: %.3f ms -> %.1f GB/s inbound per GPU" % (mb, ms, mb * (world - 1) / 1024 / (ms * 1e-3)))
    assert float(dst[(rank + 1) % world, 0]) == float((rank + 1) % world)
dist.destroy_process_group()
