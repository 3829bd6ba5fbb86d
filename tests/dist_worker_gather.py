"""Worker of tests/test_gpu_dist.py (launched by torchrun, one process per GPU): the fused UKF + all-gather
over peer memory must leave, on EVERY rank, exactly what an NCCL all_gather of the per-rank results gives."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lidar_slam_b200 import batch, pipeline, dist as ldist  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = batch.Engine(local)
    S = 50000
    pk = eng.synth_express(S, 12, seed=4, scan_base=rank * S)
    xy, _, _ = eng.express_to_cart(pk, 360)
    for multicast in (False, True, None):
        ref = pipeline.ScanPipeline(eng, S, seed=4, scan_base=rank * S)
        fused = pipeline.ScanPipeline(eng, S, seed=4, scan_base=rank * S)
        r0 = eng.ransac_lines(xy, seed=4, scan_base=rank * S)
        for p in (ref, fused):
            p.install_map_from_lines(r0["line"], r0["ninl"])
            p.ukf["u"].fill_(1.0)
        try:
            pg = fused.enable_peer_gather(multicast=multicast)
        except RuntimeError as e:
            if multicast:                      # no multicast mapping on this fabric: nothing to test for this mode
                if rank == 0:
                    print("multicast unavailable:", e, flush=True)
                continue
            raise
        nccl = ldist.ResultGather()
        for step in range(3):
            res = ref.step(xy)
            want = nccl(res)
            if step == 1:                      # a step issued in slices fills the same records
                for a, b in ((0, 12345), (12345, S)):
                    fused.step(xy[a:b], s0=a, s1=b)
            else:
                fused.step(xy)
            pg.barrier()
            torch.cuda.synchronize()
            got = pg.result()
            x = got["state"][..., :3]
            P = got["state"][..., 3:].reshape(world, S, 3, 3)
            assert torch.equal(x.contiguous().view(torch.uint8), want["x"].view(torch.uint8)), (multicast, step, "x")
            assert torch.equal(P.contiguous().view(torch.uint8), want["P"].view(torch.uint8)), (multicast, step, "P")
            assert torch.equal(got["ids"], want["match_id"]), (multicast, step, "ids")
            # no second barrier: a peer that is already in the next step writes the OTHER copy (PeerGather docstring)
        if rank == 0:
            print("peer gather ok: multicast=%s used=%s world=%d" % (multicast, pg.multicast, world), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
