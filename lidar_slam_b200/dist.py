"""Multi-GPU plumbing: the path shards by scan / stream index with no data-path exchange
(SURVEY.md 8(e)); the only collective is the gather of fixed-size per-scan results."""
import torch
import torch.distributed as dist


def shard_range(total, rank, world):
    """Contiguous block partition: rank r owns [r*ceil(total/world), min(total, (r+1)*ceil(total/world)))."""
    per = (total + world - 1) // world
    lo = min(total, rank * per)
    return lo, min(total, lo + per)


class ResultGather:
    """all_gather of the per-scan results of a step (NCCL over NVLink on GPUs, gloo in the CPU
    tests) into persistent [world, ...] buffers; asynchronous on the current stream for NCCL."""

    def __init__(self, keys=("x", "P", "match_id")):
        self.keys = keys
        self.buf = None

    def __call__(self, res):
        world = dist.get_world_size() if dist.is_initialized() else 1
        if world == 1:
            return {k: res[k].unsqueeze(0) for k in self.keys}
        if self.buf is None:
            self.buf = {k: torch.empty((world,) + tuple(res[k].shape), dtype=res[k].dtype, device=res[k].device)
                        for k in self.keys}
        nccl = dist.get_backend() == "nccl"
        for k in self.keys:
            if nccl:
                dist.all_gather_into_tensor(self.buf[k], res[k].contiguous())
            else:                                   # gloo (CPU tests): list form into views of the same buffer
                dist.all_gather(list(self.buf[k].unbind(0)), res[k].contiguous())
        return self.buf
