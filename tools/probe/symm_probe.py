"""Diagnostic (torchrun): does torch symmetric memory (peer-mapped buffers) work on this box?"""
import os

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = symm_mem.empty((world, 1 << 20), dtype=torch.float64, device=torch.device("cuda", local))
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast", hex(hdl.multicast_ptr) if hdl.has_multicast_support else None,
      "backend", symm_mem.get_backend(torch.device("cuda", local)), flush=True)
t.fill_(-1.0)
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (world, 1 << 20), torch.float64)
peer[rank].fill_(float(rank))          # P2P store into the peer's buffer
hdl.barrier(channel=0)
torch.cuda.synchronize()
print(rank, "row written by peer:", float(t[(rank + 1) % world, 5]), "expected", float((rank + 1) % world), flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    hdl.barrier(channel=0)
e1.record(); torch.cuda.synchronize()
print(rank, "barrier %.1f us" % (e0.elapsed_time(e1) / 20 * 1e3), flush=True)
dist.destroy_process_group()
