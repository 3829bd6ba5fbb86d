#!/usr/bin/env python
"""bench.py - the headline benchmark (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--scans S] [--impl ours|reference]

Workload (BASELINE.json configs[4], per GPU): S = 1,048,576 synthetic 360-beam scans per
step, generated on the device, each through RANSAC line extraction (4 chunks x 100 trials) ->
landmark association (4 queries against the stream's 8-slot table) -> UKF predict/update.
One process per GPU; scans shard by index (scan_base = rank * S), no data-path collective;
the per-rank results are gathered over NCCL at the end of each step (N > 1).

Prints ONE JSON line (rank 0).  `value` = scans/s, whole job, inputs resident in HBM;
`e2e` = the same through the host-buffer API (pinned host scans in, host results out, copies
inside the timed region); `roofline` = the RANSAC kernel against the measured fp64 issue rate
of this GPU and its HBM traffic against MEASURED_PEAKS.json; `cpu_baseline` = the C oracle
(all host threads) on a bounded sample, on rank 0 at N = 1.

--impl reference times the CPU restatement of the reference path (oracle/, all host threads)
on a bounded sample of the same workload (the reference itself is pure Python on top of
skimage/filterpy, neither installable here: DESIGN.md "Reference arm").
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BEAMS, CHUNK, TRIALS, THR, TABLE_M = 360, 100, 100, 20.0, 8
PACKETS = 12                                       # express packets per revolution (32 measures each)
E2E_SLICES = 16                                    # slices of the overlapped host path, one step in flight
E2E_SLICES_PIPELINED = 8                           # ... two steps in flight (22.5 ms per step against 23.3 at 16 and 22.4 at 4)
PHE_PER_SCAN = BEAMS * TRIALS                      # point-hypothesis evaluations per scan
FLOP_PER_PHE = 8                                   # algorithmic fp64 flop per PHE
BYTES_PER_SCAN = BEAMS * 16 + 4 * 96               # 5,760 in + 4 x 96 out = 6,144 B
# DRAM bytes per scan of the RANSAC kernel: profiles/ransac_traffic.json, written from the latest ncu --set full
# capture of the CURRENT kernel (tools/ncu_traffic.py); None when the file is absent


def ncu_traffic():
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ransac_traffic.json")))
        return (float(t["dram_bytes_read"]) + float(t["dram_bytes_write"])) / float(t["scans"]), t
    except Exception:
        return None, None


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--scans", type=int, default=1 << 20, help="scans per GPU per step")
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--cpu-sample", type=int, default=0, help="scans in the CPU sample (0 = auto)")
    p.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-aug", action="store_true", help="skip the configs[3] leg (augmented-state UKF)")
    p.add_argument("--aug-filters", type=int, default=10000, help="configs[3]: 10,000 filters")
    p.add_argument("--no-legs", action="store_true", help="skip the configs[1] / configs[2] RANSAC-only legs and the OLS leg")
    p.add_argument("--no-points", action="store_true", help="skip the e2e leg that also returns the projected inlier points")
    p.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                   help="weak: --scans per GPU (default, the headline); strong: --scans in total, split over the GPUs")
    p.add_argument("--slices", type=int, default=0, help="slices of the overlapped host path (0 = default)")
    p.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                   help="N > 1 result gather: fused into the UKF kernel over peer memory, or an NCCL all_gather")
    return p.parse_args()


# ---- clocks under load ----------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.rows, self.thread = index, None, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for ts, line in self.rows:
            if t0 is not None and not (t0 <= ts <= t1 + 0.1):
                continue
            f = [v.strip() for v in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(self.NAMES, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- CPU legs (the only place bench.py touches oracle/) -----------------------------------------
def host_threads():
    """All the host threads this process may use - set explicitly, because torchrun exports OMP_NUM_THREADS=1 to its
    workers and the reference arm would otherwise be a one-thread run at N > 1."""
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        from oracle import cport
        cport.set_threads(n)
        n = cport.max_threads()
    except Exception:
        pass
    return n


def numpy_rates(n1=24, npool=None):
    """The same job in Python/NumPy (oracle/numpy_pipeline.py: per-trial NumPy calls like skimage / filterpy make), on
    one core and on a multiprocessing.Pool of all cores.  Returns dict."""
    import multiprocessing as mp
    import numpy as np
    from lidar_slam_b200 import synth
    from oracle import numpy_pipeline as npp
    cores = len(os.sched_getaffinity(0))
    xy = synth.random_room_scans(n1, seed=11)
    n, dt = npp.run_scans((xy, 0, 0))
    out = {"numpy_1core": {"value": n / dt, "unit": "scans/s", "cores": 1, "kind": "port (Python/NumPy restatement)",
                           "sample": "%d scans, %.1f s" % (n, dt)}}
    npool = npool or 12 * cores
    xy = synth.random_room_scans(npool, seed=12)
    parts = [(xy[i::cores], i, 0) for i in range(cores)]
    try:
        with mp.get_context("fork").Pool(cores) as pool:
            t0 = time.perf_counter()
            done = pool.map(npp.run_scans, parts)
            wall = time.perf_counter() - t0
        # the pool's wall time includes each worker's (untimed-in-the-other-arms) map set-up: use the slowest worker's own clock
        slow = max(d for _, d in done)
        out["numpy_pool"] = {"value": sum(k for k, _ in done) / slow, "unit": "scans/s", "cores": cores,
                             "kind": "port (Python/NumPy restatement, multiprocessing.Pool)",
                             "sample": "%d scans, slowest worker %.1f s (pool wall %.1f s)" % (npool, slow, wall)}
    except Exception as e:
        out["numpy_pool"] = {"unavailable": str(e)[:100]}
    return out


def sklearn_ols_rate(nwin=300):
    """R7 as the reference runs it (line_detection.py:57-58): sklearn RANSACRegressor(residual_threshold=3).fit per
    50-point window, timed as-is."""
    import numpy as np
    try:
        from sklearn.linear_model import RANSACRegressor
    except Exception as e:
        return {"unavailable": str(e)[:100]}
    from lidar_slam_b200 import synth
    xy = synth.random_room_scans((nwin * 50 + 359) // 360, seed=13).reshape(-1, 2)[:nwin * 50].reshape(nwin, 50, 2)
    t0 = time.perf_counter()
    ok = 0
    for w in xy:
        try:
            RANSACRegressor(residual_threshold=3, random_state=0).fit(w[:, :1], w[:, 1])
            ok += 1
        except ValueError:
            pass
    dt = time.perf_counter() - t0
    return {"value": nwin / dt, "unit": "windows/s", "cores": 1, "kind": "reference (scikit-learn as-is)",
            "sample": "%d windows of 50 points, %.1f s, %d fitted" % (nwin, dt, ok)}


def cpu_pipeline_rate(nscans, seed=0):
    """Scans/s of the CPU restatement of the same job (oracle/: express-packet decode + polar->cartesian,
    RANSAC, association and UKF in C, OpenMP over scans, all host threads) on `nscans` synthetic
    revolutions given as raw sensor packets, with the same known-map set-up as the GPU arm.
    Returns (scans_per_s, threads, seconds)."""
    import numpy as np
    from lidar_slam_b200 import synth
    from oracle import cport
    from tests import pipeline_util
    pk = synth.random_room_express(nscans, PACKETS, seed=seed)
    threads = host_threads()
    xy, _, _ = cport.express_to_cart(pk, BEAMS)
    r0 = cport.ransac_lines(xy)                                    # untimed set-up: the known map
    tables = pipeline_util.map_from_lines(r0["line"], r0["ninl"], TABLE_M)
    x = np.zeros((nscans, 3))
    P = np.tile(np.diag([0.1, 0.1, 0.05]), (nscans, 1, 1))
    u = np.ones((nscans, 2))
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        xy, _, _ = cport.express_to_cart(pk, BEAMS)
        _, _, x, P, _ = cport.pipeline_step(xy, tables, x, P, u)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return nscans / best, threads, best


def run_reference(args, rank, world):
    if rank != 0:
        return
    n = args.cpu_sample or 16384
    rates = []
    for _ in range(args.warmup):
        cpu_pipeline_rate(min(n, 2048))
    t_all = time.perf_counter()
    threads = 1
    for _ in range(args.steps):
        r, threads, _ = cpu_pipeline_rate(n)
        rates.append(r)
    dt = time.perf_counter() - t_all
    value = statistics.median(rates)
    sample = "%d scans per step (bounded sample of the %d-scan step), C oracle with OpenMP" % (n, args.scans)
    line = {
        "impl": "reference", "metric": "scans/s (RANSAC+assoc+UKF)", "value": value, "unit": "scans/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": "scans/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def aug_leg(eng, F, steps, fp64_peak, L=200):
    """BASELINE.json configs[3]: F independent filters with 3-DOF pose + L landmarks in the state (n = 3 + 2L),
    one predict + update per step through lslam_ukf_aug_step, state and measurements resident in HBM (the
    per-step working set, 13 MB of scratch per filter, is far larger than L2).  filter-steps/s; the flop
    figure is SURVEY.md 8(d)'s literal-filterpy dense count (1.34 GFLOP per step at L = 200, scaled by n^3)."""
    import numpy as np
    import torch
    from lidar_slam_b200 import synth
    nb = min(F, 16)
    x, P, u, z = synth.random_aug_filters(nb, L, seed=0)
    rep = (F + nb - 1) // nb
    dev = eng.device
    t = lambda a: torch.from_numpy(np.tile(a, (rep,) + (1,) * (a.ndim - 1))[:F].copy()).to(dev)
    st = eng.alloc_ukf_aug(F, L)
    x0, P0, ut, zt = t(x), t(P), t(u), t(z)
    ms, launches = [], 0
    for r in range(steps + 2):
        st["x"].copy_(x0); st["P"].copy_(P0)
        torch.cuda.synchronize()
        l0 = eng.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.ukf_aug_step(st, zt, u=ut)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        launches = eng.launch_count() - l0
    bad = int((st["status"] != 0).sum().item())
    t_ms = sum(ms[2:]) / len(ms[2:])
    # configs[3]'s association: L queries (the observed landmarks) against the filter's table of L stored landmarks,
    # first-match mode; the stored rows are the queries of the previous revolution, slightly moved, so most queries hit
    g = torch.Generator(device="cpu").manual_seed(3)
    ab = torch.randn((F, L, 2), generator=g, dtype=torch.float64) * torch.tensor([0.5, 1500.0], dtype=torch.float64)
    px = torch.randn((F, L, 1), generator=g, dtype=torch.float64) * 2000.0
    ex = px + 80.0 * torch.randn((F, L, 1), generator=g, dtype=torch.float64)
    rows = torch.cat([ab, px, ab[..., :1] * px + ab[..., 1:], ex, ab[..., :1] * ex + ab[..., 1:]], dim=-1).contiguous().to(dev)
    q = rows.clone()
    q[..., 1] += 2.0                                                          # b moves by 2 mm: inside every gate
    a_ms = []
    for r in range(3):
        tb = eng.alloc_tables(F, L)
        tb["rows"].copy_(rows); tb["life"].fill_(40); tb["count"].fill_(L); tb["next_id"].fill_(L)
        tb["ids"].copy_(torch.arange(L, dtype=torch.int32, device=dev).expand(F, L))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ao = eng.associate(q, tb, mode=1)
        e1.record()
        torch.cuda.synchronize()
        a_ms.append(e0.elapsed_time(e1))
    assoc_hits = float((ao["match"] >= 0).float().mean().item())
    n = 3 + 2 * L
    flop = 1.34e9 * (n / 403.0) ** 3                                          # literal filterpy count
    m, dz = 2 * n + 1, 2 * L
    # what the kernels execute: 3 Cholesky, Pbar and S on the upper triangle, the paired cross-covariance D^T T over the
    # non-zero rows of T, the forward solve W = U^-T [Pzx | r], and P = Pbar - W^T W on the upper triangle
    flop_exec = (2 * n ** 3 / 3 + dz ** 3 / 3) + m * n * n + m * dz * dz + 2 * dz * n * (n / 2 + 32) + dz * dz * (n + 1) + dz * n * n
    tf = flop_exec * F / (t_ms * 1e-3) * 1e-12
    tf_lit = flop * F / (t_ms * 1e-3) * 1e-12
    del st, x0, P0
    torch.cuda.empty_cache()
    return {"workload": "BASELINE.json configs[3]: %d filters, pose + %d landmarks in the state (n = %d), predict + update" % (F, L, n),
            "value": F / (t_ms * 1e-3), "unit": "filter-steps/s", "ms_per_step": t_ms, "gpu_launches": launches,
            "status_nonzero": bad,
            "associate": {"queries": L, "stored": L, "mode": "first match", "ms": min(a_ms), "hit_rate": assoc_hits},
            "roofline": {"bound": "fp64 tensor pipe (DMMA = DFMA rate)", "achieved": tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": tf / fp64_peak if fp64_peak > 0 else None,
                         "flop_per_filter_step": flop_exec,
                         "flop_source": "executed: W-form update, triangle-only products, paired cross-covariance (DESIGN.md 4.6)",
                         "literal_equivalent": {"tflops": tf_lit, "flop_per_filter_step": flop,
                                                "source": "SURVEY.md 8(d) literal filterpy count; its ceiling at the fp64 peak is %.1f k filter-steps/s"
                                                          % (fp64_peak * 1e12 / flop / 1e3 if fp64_peak > 0 else 0.0)}}}


def l2_flush_buffer(torch, dev):
    return torch.empty(256 << 20, dtype=torch.uint8, device=dev)          # > the 126 MB L2


def ransac_leg(eng, name, B, beams, chunk, trials, steps, fp64_peak, flush):
    """BASELINE.json configs[1] / configs[2]: RANSAC line extraction only (lslam_ransac_lines with its default kernel
    choice and knobs), scans resident in HBM.  These batches are smaller than L2, so L2 is flushed (a 256 MB memset)
    before every timed launch; each launch is timed on its own with CUDA events."""
    import torch
    xy = eng.synth_scans(B, beams=beams, seed=5)
    out = eng.alloc_ransac_out(B, beams, chunk, ssum=False)
    C = (beams + chunk - 1) // chunk
    ms = []
    for r in range(steps + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.ransac_lines(xy, chunk=chunk, trials=trials, thr=THR, seed=0, out=out)
        e1.record()
        torch.cuda.synchronize()
        if r >= 3:
            ms.append(e0.elapsed_time(e1))
    t = statistics.median(ms)
    phe = B * beams * trials
    tf = phe * FLOP_PER_PHE / (t * 1e-3) * 1e-12
    ok = int((out["ninl"] >= 2).sum().item())
    return {"workload": "%s: %d scans x %d beams, chunk %d (%d lines per scan), %d hypotheses per line, RANSAC + refit only"
                        % (name, B, beams, chunk, C, trials),
            "kernel_ms": t, "scans_per_s": B / (t * 1e-3), "hyps_per_s": B * C * trials / (t * 1e-3), "lines_found": ok,
            "l2": "flushed before every launch (input %.1f MB < 126 MB L2)" % (B * beams * 16 / 1e6),
            "roofline": {"bound": "fp64-issue", "achieved": tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": tf / fp64_peak if fp64_peak > 0 else None,
                         "algorithmic_flop": phe * FLOP_PER_PHE,
                         "hbm_gbs": (B * beams * 16 + B * C * 96) / (t * 1e-3) * 1e-9}}


def ols_leg(eng, steps, flush, B=65536):
    """R7 / N4 (line_detection.py:51-73): every 50-point window of B scans through lslam_ransac_ols."""
    import torch
    xy = eng.synth_scans(B, seed=6)
    ms = []
    for r in range(steps + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        o = eng.ransac_ols(xy, window=50, thr=3.0)
        e1.record()
        torch.cuda.synchronize()
        if r >= 2:
            ms.append(e0.elapsed_time(e1))
    t = statistics.median(ms)
    nwin = B * 8
    return {"workload": "line_detection.py flavour: %d windows of <= 50 points, RANSACRegressor(residual_threshold=3) semantics" % nwin,
            "kernel_ms": t, "windows_per_s": nwin / (t * 1e-3), "mean_trials": float(o["n_trials"].float().mean().item()),
            "hbm_gbs": (B * 360 * 16 + nwin * 84) / (t * 1e-3) * 1e-9}


def workload_config(args, world):
    return {"workload": "BASELINE.json configs[4]: 1M synthetic 360-beam scans end-to-end "
                        "(RANSAC 4 chunks x 100 trials + association M=8 + UKF n=3) per GPU",
            "scans_per_gpu_per_step": args.scans if args.scaling == "weak" else args.scans // world, "beams": BEAMS, "chunk": CHUNK, "trials": TRIALS,
            "threshold_mm": THR, "table_capacity": TABLE_M,
            "input": "synthetic sensor: %d express packets (84 B, 32 measures) per revolution, first %d measures "
                     "form the scan; `value` starts from the decoded points resident in HBM, `e2e` and the CPU "
                     "legs start from the packet bytes in host memory" % (PACKETS, BEAMS), "sharding": "scans by index, %d rank(s)" % world,
            "l2": "inputs (%.2f GB per step) exceed the 126 MB L2" % (args.scans * BEAMS * 16 / 1e9)}


# ---- our arm ---------------------------------------------------------------------------------------
def main():
    args = parse()
    global E2E_SLICES, E2E_SLICES_PIPELINED
    if args.slices > 0:
        E2E_SLICES = E2E_SLICES_PIPELINED = args.slices
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from lidar_slam_b200 import batch, pipeline

    torch.cuda.set_device(local)
    from lidar_slam_b200 import dist as ldist
    numa_cpus = ldist.bind_to_gpu_numa(local) if world > 1 else None       # host buffers on the GPU's socket
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = batch.Engine(local)
    S = args.scans if args.scaling == "weak" else args.scans // world
    pipe = pipeline.ScanPipeline(eng, S, beams=BEAMS, chunk=CHUNK, trials=TRIALS, thr=THR,
                                 table_capacity=TABLE_M, seed=0, scan_base=rank * S)
    # untimed set-up: the synthetic sensor's bytes, decoded once; xy stays resident in HBM
    pk = eng.synth_express(S, PACKETS, seed=0, scan_base=rank * S)
    xy, _, pk_status = eng.express_to_cart(pk, BEAMS)
    assert not bool(pk_status.any().item()), "synthetic packets rejected"
    r0 = eng.ransac_lines(xy, chunk=CHUNK, trials=TRIALS, thr=THR, seed=0, scan_base=rank * S, out=pipe.rout)
    pipe.install_map_from_lines(r0["line"], r0["ninl"])                     # untimed set-up: the known map
    pipe.ukf["u"].fill_(1.0)
    gatherer = ldist.ResultGather()
    gather_mode = "none (1 rank)"
    if world > 1:
        gather_mode = "nccl all_gather"
        if args.gather == "peer":
            # the product path: every filter's (x, P, match_id) is stored into all ranks' symmetric-memory
            # buffers by the UKF kernel itself; the step ends with a cross-rank barrier on the stream
            try:
                pg = pipe.enable_peer_gather()
                gather_mode = "fused in the UKF kernel over %s" % (
                    "NVSwitch multicast (multimem stores)" if pg.multicast else "NVLink peer stores")
            except Exception as e:                                          # transport set-up, not compute
                gather_mode = "nccl all_gather (peer-memory rendezvous failed: %s)" % (str(e).splitlines()[0][:120])
                pipe.peer_gather = None

    def gather(res):
        """N > 1: all ranks' per-scan UKF/landmark results, inside the timed region."""
        if world > 1:
            if pipe.peer_gather is not None:
                pipe.peer_gather.barrier()
            else:
                gatherer(res)

    def one_step():
        res = pipe.step(xy)
        gather(res)
        return res

    fp64_peak = eng.fp64_peak(40.0) if rank == 0 else 0.0                   # TFLOP/s, dependent DFMA chains
    for _ in range(max(args.warmup, 3)):
        one_step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                                                     # nvidia-smi needs a moment to start
        time.sleep(0.3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()                                                      # all ranks enter the timed region together
    torch.cuda.synchronize()
    n0 = eng.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0 = time.perf_counter()
    ev[0].record()
    for k in range(args.steps):
        # the dominant kernel (lslam_ransac_lines) is bracketed by its own events on the same stream
        res = pipe.step(xy, ransac_events=kev[k])
        gather(res)
    ev[1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    launches = eng.launch_count() - n0
    ms = ev[0].elapsed_time(ev[1])
    k_ms = statistics.mean(a_.elapsed_time(b_) for a_, b_ in kev)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    faults = eng.ransac_counters()["faults"]
    hit_rate = float((pipe.aout["match"] >= 0).float().mean().item())
    assert int(pipe.ukf["status"].abs().sum().item()) == 0, "UKF status flags set"

    # ---- end to end through the host-buffer API --------------------------------------------
    def e2e_leg(host_in, pipelined=True, slices=None):
        """K steps through ScanPipeline.step_host, host buffers in and out.  pipelined: a step is queued before the host
        has read the previous step's results (two pinned result sets alternate), so step k + 1's uploads overlap step k's
        kernels - what a stream of revolutions allows; every step still uploads its own input and the host reads every
        step's results inside the timed region.  pipelined=False: the host waits for each step before queueing the next."""
        slices = slices or (E2E_SLICES_PIPELINED if pipelined else E2E_SLICES)
        for _ in range(3):                                                   # untimed: both pinned result sets exist afterwards
            pipe.step_host(host_in, slices=slices, wait=not pipelined)
        pipe.drain()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        n0e = eng.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if pipelined:
            pending = None
            for _ in range(args.steps):
                out_h = pipe.step_host(host_in, slices=slices, wait=False)
                if pending is not None:
                    pending["done"].synchronize()                                # the previous step's results are on the host
                    _ = float(pending["ninl"][0, 0])
                pending = out_h
            pending["done"].synchronize()
            _ = float(pending["ninl"][0, 0])
            pipe.drain()
        else:
            for _ in range(args.steps):
                out_h = pipe.step_host(host_in, slices=slices)
                torch.cuda.current_stream().synchronize()                        # results are on the host
                _ = float(out_h["ninl"][0, 0])
        e1.record()
        torch.cuda.synchronize()
        ems = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ems], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ems = float(t.item())
        h2d, d2h = pipe.io_bytes(host_in)
        return {"value": world * S * args.steps / (ems * 1e-3), "unit": "scans/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ems / args.steps,
                "slices": slices, "gpu_launches": eng.launch_count() - n0e,
                "steps_in_flight": 2 if pipelined else 1}

    e2e = e2e_xy = None
    if not args.no_e2e:
        # headline: the sensor's own bytes in (what Lidar._read_response returns, lidar.py:161-167),
        # decoded on the device, per-scan results out
        pk_h = pipe.alloc_host_io("packets", PACKETS)
        pk_h.copy_(pk)                                                       # untimed: fill the host buffer
        torch.cuda.synchronize()
        # two ways to drive the same API, both timed: two steps in flight (step k + 1 is queued, its uploads running, before the
        # host has read step k's results) and one step in flight (the host waits for every step before it queues the next; the
        # e2e figure of round 1).  Either way every step uploads its own input and the host reads every step's results inside
        # the timed region.  `e2e` is the faster of the two on this box (at 1-2 GPUs the pipelined one; with 8 ranks saturating
        # the host's copy bandwidth the other), the slower one is kept beside it.
        e2e_p = e2e_leg(pk_h, pipelined=True)
        e2e_s = e2e_leg(pk_h, pipelined=False)
        if world > 1:                                                        # every rank must pick the same leg
            t = torch.tensor([e2e_p["value"], e2e_s["value"]], dtype=torch.float64, device="cuda")
            dist.broadcast(t, src=0)
            e2e_p["value"], e2e_s["value"] = float(t[0]), float(t[1])
        e2e, other = (e2e_p, e2e_s) if e2e_p["value"] >= e2e_s["value"] else (e2e_s, e2e_p)
        e2e["input"] = "RPLidar express packets, %d x 84 B per scan, decoded on the device" % PACKETS
        e2e["mode"] = "%d step(s) in flight" % e2e["steps_in_flight"]
        e2e["other_mode"] = {"steps_in_flight": other["steps_in_flight"], "value": other["value"], "ms_per_step": other["ms_per_step"],
                             "slices": other["slices"]}
        del pk_h
        # the same with cartesian float64 scans in (what scanning() queues, functions.py:59-66)
        xy_h = pipe.alloc_host_io("xy")
        xy_h.copy_(xy)
        torch.cuda.synchronize()
        e2e_xy = e2e_leg(xy_h)
        e2e_xy["input"] = "cartesian float64 points, 5760 B per scan"
        del xy_h

    # ---- the same end to end INCLUDING what landmark_extraction itself returns: the projected inlier points
    #      (ransac_functions.py:28,46-55) compacted on the device (lslam_project_inliers) and copied to the host
    e2e_points = None
    if not args.no_e2e and not args.no_points and world == 1:
        pk_h = pipe.alloc_host_io("packets", PACKETS)
        pk_h.copy_(pk)
        torch.cuda.synchronize()
        for _ in range(2):
            pipe.step_host(pk_h, slices=E2E_SLICES, points=True)
        torch.cuda.synchronize()
        n0e = eng.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nsteps = max(2, args.steps // 2)
        e0.record()
        rows = 0
        for _ in range(nsteps):
            out_h = pipe.step_host(pk_h, slices=E2E_SLICES, points=True)
            torch.cuda.current_stream().synchronize()
            rows = sum(int(v.shape[0]) for v in out_h["points"])
            _ = float(out_h["points"][0][0, 0])
        e1.record()
        torch.cuda.synchronize()
        ems = e0.elapsed_time(e1)
        h2d, d2h = pipe.io_bytes(pk_h)
        e2e_points = {"value": S * nsteps / (ems * 1e-3), "unit": "scans/s", "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": d2h + rows * 16 + 8 * E2E_SLICES, "ms_per_step": ems / nsteps, "slices": E2E_SLICES,
                      "gpu_launches": eng.launch_count() - n0e, "points_per_step": rows,
                      "input": "express packets in; per-scan results AND every chunk's projected inlier points out"}
        del pk_h

    # ---- N > 1: the fused gather's records against an NCCL all_gather of the same step (once, untimed)
    gather_check = None
    if world > 1 and pipe.peer_gather is not None:
        res = pipe.step(xy)
        pipe.peer_gather.barrier()
        want = ldist.ResultGather()(res)
        torch.cuda.synchronize()
        got = pipe.peer_gather.result()
        same = (torch.equal(got["state"][..., :3].contiguous().view(torch.uint8), want["x"].view(torch.uint8)) and
                torch.equal(got["state"][..., 3:].contiguous().view(torch.uint8), want["P"].reshape(world, S, 9).view(torch.uint8)) and
                torch.equal(got["ids"], want["match_id"]))
        t = torch.tensor([1.0 if same else 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        gather_check = "bit-identical to nccl all_gather on every rank" if t.item() == 1.0 else "MISMATCH"

    # ---- N > 1, weak run: the same job as STRONG scaling beside it (configs[4] as written: 1M scans in total)
    strong = None
    if world > 1 and args.scaling == "weak":
        S2 = args.scans // world
        pipe2 = pipeline.ScanPipeline(eng, S2, beams=BEAMS, chunk=CHUNK, trials=TRIALS, thr=THR, table_capacity=TABLE_M,
                                      seed=0, scan_base=rank * S2)
        xy2 = xy[:S2]
        r2 = eng.ransac_lines(xy2, chunk=CHUNK, trials=TRIALS, thr=THR, seed=0, scan_base=rank * S2, out=pipe2.rout)
        pipe2.install_map_from_lines(r2["line"], r2["ninl"])
        pipe2.ukf["u"].fill_(1.0)
        g2 = None
        if pipe.peer_gather is not None:
            try:
                g2 = pipe2.enable_peer_gather()
            except Exception:
                g2 = None
        gat2 = ldist.ResultGather()

        def step2():
            r = pipe2.step(xy2)
            if g2 is not None:
                g2.barrier()
            else:
                gat2(r)
        for _ in range(5):
            step2()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nst = 4 * args.steps
        e0.record()
        for _ in range(nst):
            step2()
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        strong = {"scaling": "strong", "scans_total_per_step": S2 * world, "scans_per_gpu_per_step": S2, "steps": nst,
                  "ms_per_step": float(t.item()) / nst, "value": S2 * world * nst / (float(t.item()) * 1e-3), "unit": "scans/s",
                  "gather": "fused peer stores" if g2 is not None else "nccl all_gather",
                  "limit": "3 launches + the cross-rank barrier per step; the gather's tail does not shrink with the shard"}
        del pipe2

    # ---- BASELINE.json configs[1], configs[2] (RANSAC only) and the OLS flavour, beside the headline
    legs = None
    if world == 1 and not args.no_legs:
        flush = l2_flush_buffer(torch, eng.device)
        legs = {"ransac_c2": ransac_leg(eng, "BASELINE.json configs[1]", 4096, 360, 100, 100, args.steps, fp64_peak, flush),
                "ransac_c3": ransac_leg(eng, "BASELINE.json configs[2]", 256, 8192, 410, 1024, args.steps, fp64_peak, flush),
                "ransac_ols": ols_leg(eng, args.steps, flush)}
        del flush

    # ---- BASELINE.json configs[3]: augmented-state UKF (pose + 200 landmarks, n = 403), reported beside the headline
    aug = None
    if world == 1 and not args.no_aug:
        del pipe
        torch.cuda.empty_cache()
        aug = aug_leg(eng, args.aug_filters, max(2, min(args.steps, 3)), fp64_peak)

    if rank == 0:
        value = world * S * args.steps / (ms * 1e-3)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        ach_tflops = S * PHE_PER_SCAN * FLOP_PER_PHE / (k_ms * 1e-3) * 1e-12
        per_scan, tsrc = ncu_traffic()
        roofline = {
            "kernel": "ransac_tile_kernel (lslam_ransac_lines, default kernel choice)", "bound": "fp64-issue",
            "achieved": ach_tflops, "peak": fp64_peak,
            "unit": "TFLOP/s", "frac": ach_tflops / fp64_peak if fp64_peak > 0 else None,
            "flop_basis": "algorithmic: SURVEY.md 8(d) 8 fp64 flop per point-hypothesis evaluation x 36,000 per scan; the kernel "
                          "EXECUTES fewer fp64 operations than that (distances screened in fp32 with an exact re-check band, "
                          "residual sums from chunk moments - DESIGN.md 4.1), so this is a rate of useful work, not a pipe utilisation",
            "traffic": S * per_scan if (per_scan is not None and S % 4 == 0 and BEAMS == 360) else None,
            "traffic_source": ("ncu dram__bytes_read+write of one launch of the current kernel at %d scans, scaled per scan (%s)"
                               % (tsrc["scans"], tsrc["file"])) if tsrc else "no capture recorded (profiles/ransac_traffic.json absent)",
            "algorithmic_bytes": S * BYTES_PER_SCAN,
            "peak_source": "measured in this run: lslam_fp64_peak (dependent DFMA chains, 2 flop each)",
            "kernel_ms": k_ms, "hyps_per_s": S * 4 * TRIALS / (k_ms * 1e-3),
            "hbm": {"achieved": S * BYTES_PER_SCAN / (k_ms * 1e-3) * 1e-9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": S * BYTES_PER_SCAN / (k_ms * 1e-3) * 1e-9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }
        if tsrc and "issue_active_pct" in tsrc:
            # north_star's other yardstick ("HBM or issue roofline"): issue slots in use, from the same ncu capture - a
            # profiler figure of the current kernel at the captured size, not measured in this run
            roofline["issue"] = {"achieved": tsrc["issue_active_pct"], "peak": 100.0, "unit": "% of issue slots active",
                                 "frac": tsrc["issue_active_pct"] / 100.0, "warps_active_pct": tsrc.get("warps_active_pct"),
                                 "source": "ncu smsp__issue_active.avg.pct_of_peak_sustained_active (%s)" % tsrc["file"]}
        cpu = None
        if world == 1 and not args.no_cpu:
            n = args.cpu_sample or 16384
            rate, threads, secs = cpu_pipeline_rate(n)
            cpu = {"value": rate, "unit": "scans/s", "cores": threads, "kind": "port",
                   "sample": "%d of the %d scans of one step from packet bytes, C oracle (OpenMP), %.1f s" % (n, S, secs)}
            cpu.update(numpy_rates())
            cpu["sklearn_ols"] = sklearn_ols_rate()
        line = {
            "metric": "scans/s (RANSAC+assoc+UKF)", "value": value, "unit": "scans/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, world),
            "stages": ["ransac", "associate", "ukf"], "hit_rate": hit_rate, "gather": gather_mode,
            "ransac_hyps_per_s": world * S * 4 * TRIALS * args.steps / (ms * 1e-3),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_xy": e2e_xy, "gpu_launches": launches,
            "clocks": clocks, "parity_faults": faults, "ukf_aug": aug, "e2e_points": e2e_points,
            "gather_check": gather_check, "strong_scaling": strong,
        }
        if legs:
            line.update(legs)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
