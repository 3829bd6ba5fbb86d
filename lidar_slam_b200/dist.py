"""Multi-GPU plumbing: the path shards by scan / stream index with no data-path exchange
(SURVEY.md 8(e)); the only collective is the gather of fixed-size per-scan results.

Two implementations of that gather:
  * ``PeerGather`` - the product path on GPUs: the ranks' gather buffers live in symmetric memory
    (peer-mapped over NVLink, with an NVSwitch multicast mapping when the fabric offers one) and the UKF
    kernel itself stores every filter's result into all of them (lslam_ukf_step_gather): no separate
    collective, only a cross-rank barrier on the stream before the records are read;
  * ``ResultGather`` - a plain all_gather (NCCL on GPUs; gloo in the CPU tests of the host logic),
    kept as the baseline the fused path is measured against (bench.py --gather nccl)."""
import torch
import torch.distributed as dist


def bind_to_gpu_numa(device_index):
    """Pin this process to the CPUs NVML reports as local to the GPU, so that pinned host buffers
    (first touch) and the copy-issuing thread sit on the GPU's own socket.  Returns the CPU set used, or
    None when NVML or the affinity call is unavailable (the process then stays where it is)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


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


class PeerGather:
    """Gather buffers in symmetric memory + the descriptor lslam_ukf_step_gather needs.

    state [2, world, F, 12] float64 (x[3], P[9] of every filter of every rank) and ids
    [2, world, F, id_slots] int32: TWO copies, written alternately by step parity.  After ``barrier()`` has passed
    on the stream that ran step k, ``result()`` is step k's records, identical on every rank.  One barrier per step is
    enough: a peer that runs ahead writes step k + 1 into the OTHER copy, and cannot start step k + 2 (which reuses
    step k's copy) before it has passed barrier k + 1 - which needs this rank's arrival there.  So the records of
    step k stay valid until this rank itself queues ``barrier()`` for step k + 1; read (or copy) them before that.
    ``multicast=None`` uses the multicast mapping when there is one."""

    def __init__(self, filters_per_rank, id_slots, device, group=None, multicast=None):
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.F, self.id_slots = int(filters_per_rank), int(id_slots)
        self.state = symm_mem.empty((2, self.world, self.F, 12), dtype=torch.float64, device=device)
        self.ids = symm_mem.empty((2, self.world, self.F, max(self.id_slots, 1)), dtype=torch.int32, device=device)
        self._parity, self._done = 0, 0          # copy the running step writes / copy of the last completed step
        self.h_state = symm_mem.rendezvous(self.state, group)
        self.h_ids = symm_mem.rendezvous(self.ids, group)
        has_mc = bool(self.h_state.has_multicast_support and self.h_state.multicast_ptr and self.h_ids.multicast_ptr)
        if multicast and not has_mc:
            raise RuntimeError("symmetric memory has no multicast mapping on this system")
        self.multicast = has_mc if multicast is None else bool(multicast)

    def descriptor(self, ids, s0=0):
        """For a call that runs the local filters [s0, s0 + n): ``ids`` = its [n, id_slots] int32 tensor."""
        d = {"world": self.world, "first_record": (self._parity * self.world + self.rank) * self.F + s0,
             "ids": ids if self.id_slots else None,
             "id_slots": self.id_slots}
        if self.multicast:
            d["mc_state"], d["mc_ids"] = int(self.h_state.multicast_ptr), int(self.h_ids.multicast_ptr)
        else:
            d["peer_state"], d["peer_ids"] = list(self.h_state.buffer_ptrs), list(self.h_ids.buffer_ptrs)
        return d

    def barrier(self):
        """End of a step: cross-rank barrier on the current stream (every rank's stores of the step have landed), then
        the next step writes the other copy."""
        self.h_state.barrier(channel=0)
        self._done = self._parity
        self._parity ^= 1

    def result(self):
        """Records of the last completed step: state [world, F, 12], ids [world, F, id_slots] (views)."""
        return {"state": self.state[self._done], "ids": self.ids[self._done]}
