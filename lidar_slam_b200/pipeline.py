"""Batched scan pipeline: RANSAC line extraction -> landmark association -> UKF predict/update.

One `ScanPipeline` owns, for S independent streams (one scan per stream per step), the
device-resident state the reference keeps in Python objects of one process: the landmark list
of ransac_functions.check_ransac (ransac_functions.py:63-78) as a table per stream, and the
filter of systemClass.System (systemClass.py:14-29) as (x, P) per stream.  A step is what the
reference does for one revolution of one sensor - landmark_extraction per chunk
(ransac_functions.py:15-59), then (README.md:4, never wired up in the reference) the UKF
predict/update with the matched landmarks - for all S streams at once, three kernel launches.

All computation is liblslam.so (include/lslam.h); torch holds the buffers and the stream.
"""
import torch

from . import batch


class ScanPipeline:
    _OV_STREAMS = ("in", "dec", "ran", "run", "out")     # the overlapped host path's streams (step_host with slices > 1)

    def __init__(self, eng, streams, beams=360, chunk=batch.CHUNK, trials=batch.MAX_TRIALS,
                 thr=batch.THRESHOLD, table_capacity=8, seed=0, scan_base=0):
        self.eng = eng
        self.S, self.N, self.chunk, self.trials, self.thr = int(streams), int(beams), int(chunk), int(trials), thr
        self.C = (self.N + self.chunk - 1) // self.chunk
        self.M = int(table_capacity)
        self.seed, self.scan_base = seed, scan_base
        self.rout = eng.alloc_ransac_out(self.S, self.N, self.chunk, ssum=False)
        self.tables = eng.alloc_tables(self.S, self.M)
        self.aout = eng.alloc_assoc_out(self.S, self.C)
        self.ukf = eng.alloc_ukf(self.S, self.C)
        self.xy_dev = None
        self.pk_dev = None
        self.status_dev = None
        self._host_out = None
        self._flip = 0                # pinned result set of the last pipelined step
        self._view_cache = {}
        self._ov = None
        self.peer_gather = None
        self._pts = None              # projected inlier points (step_host(points=True)): device + pinned host buffers

    def enable_peer_gather(self, group=None, multicast=None):
        """N > 1: publish every step's (x, P, match_id) into all ranks' gather buffers from inside the UKF
        kernel (lidar_slam_b200/dist.PeerGather).  Call ``self.peer_gather.barrier()`` after the step
        before reading ``self.peer_gather.result()``; the records stay valid until the NEXT step's barrier() is
        queued on this rank (the buffers are double-buffered by step parity, one barrier per step is enough)."""
        from . import dist as ldist
        self.peer_gather = ldist.PeerGather(self.S, self.C, self.eng.device, group=group, multicast=multicast)
        return self.peer_gather

    # ---- device-resident step: inputs already in HBM ------------------------------------
    @staticmethod
    def _slice(d, s0, s1):
        return {k: (v[s0:s1] if isinstance(v, torch.Tensor) and v.dim() > 0 and k != "params" else v)
                for k, v in d.items()}

    def _views(self, s0, s1):
        """Views of the pipeline's state and result buffers for streams [s0, s1) (cached)."""
        key = (s0, s1)
        v = self._view_cache.get(key)
        if v is None:
            v = {"rout": self._slice(self.rout, s0, s1), "tables": self._slice(self.tables, s0, s1),
                 "aout": self._slice(self.aout, s0, s1), "ukf": self._slice(self.ukf, s0, s1)}
            self._view_cache[key] = v
        return v

    def step(self, xy, u=None, ransac_events=None, s0=0, s1=None):
        """xy [s1-s0, N, 2] float64 on the device: the scans of streams [s0, s1) (default: all).
        Asynchronous on the current stream.  Returns the dict of result tensors (views of buffers
        owned by the pipeline, overwritten next step).  ``ransac_events`` = (start, end)
        torch.cuda.Event pair recorded around the RANSAC launch."""
        s1 = self.S if s1 is None else s1
        if ransac_events is not None:
            ransac_events[0].record()
        r = self._step_ransac(xy, s0, s1)
        if ransac_events is not None:
            ransac_events[1].record()
        return self._step_post(r, s0, s1, u)

    def _step_ransac(self, xy, s0, s1):
        """First kernel of a step (line extraction) for streams [s0, s1), on the current stream."""
        v = self._views(s0, s1)
        return self.eng.ransac_lines(xy, chunk=self.chunk, trials=self.trials, thr=self.thr, seed=self.seed,
                                     scan_base=self.scan_base + s0, out=v["rout"])

    def _step_post(self, r, s0, s1, u=None):
        """Association and filter kernels of a step for streams [s0, s1), on the current stream; r = _step_ransac's result."""
        eng = self.eng
        v = self._views(s0, s1)
        a = eng.associate(r["line"], v["tables"], q_valid=r["ninl"], out=v["aout"])
        # measurement slot c = chunk c's line: range/bearing of its inlier centroid against the matched
        # stored landmark's position; slots whose line missed (match < 0) are not used
        gd = self.peer_gather.descriptor(a["match_id"], s0) if self.peer_gather is not None else None
        eng.ukf_step(v["ukf"], r["line"], a["hit_pos"], lm_valid=a["match"], valid_min=0, u=u, gather=gd)
        return {"line": r["line"], "ninl": r["ninl"], "match": a["match"], "match_id": a["match_id"],
                "x": v["ukf"]["x"], "P": v["ukf"]["P"], "inl_mask": r["inl_mask"], "proj_ab": a["proj_ab"]}

    def project(self, xy, res, s0=0, s1=None):
        """The projected inlier points of streams [s0, s1) (what landmark_extraction returns per chunk,
        ransac_functions.py:28,46-55) compacted in (stream, chunk, data) order into the pipeline's device buffer.
        Returns (points view [<= (s1-s0)*N, 2], offsets [(s1-s0)*C + 1] int64); offsets[-1] = rows written."""
        s1 = self.S if s1 is None else s1
        if self._pts is None:
            dev = self.eng.device
            self._pts = {"dev": torch.empty((self.S * self.N, 2), dtype=torch.float64, device=dev),
                         "off": torch.zeros((self.S * self.C + 1,), dtype=torch.int64, device=dev),
                         "host": torch.empty((self.S * self.N, 2), dtype=torch.float64).pin_memory(),
                         "total": torch.zeros((self.S,), dtype=torch.int64).pin_memory()}
        n = (s1 - s0) * self.C
        off = self._pts["off"][s0 * self.C:s0 * self.C + n + 1]
        off[0] = 0
        torch.cumsum(res["ninl"].reshape(-1).clamp(min=0), 0, out=off[1:])
        out = self._pts["dev"][s0 * self.N:s1 * self.N]
        self.eng.handle.project_inliers(xy.data_ptr(), None, s1 - s0, self.N, self.chunk, res["inl_mask"].data_ptr(),
                                        res["ninl"].data_ptr(), res["proj_ab"].data_ptr(), 2, off.data_ptr(),
                                        out.data_ptr(), self.eng._stream())
        return out, off

    # ---- a known map (benchmark / test set-up; data movement only) -----------------------------
    def install_map_from_lines(self, line, ninl, life=batch.Engine.LIFE):
        """Fill every stream's table from line records [S, C, 8] of a first pass over its scan:
        landmark c = line c with its tip placed on its centroid, so that a re-observation of the same
        wall passes the 100 mm tip/centroid gate of Landmark.is_equal (landmarking.py:66-77; two
        observations of one wall from the same pose do NOT pass it - tip and centroid are half a
        chunk apart); the remaining slots hold the next stream's lines as clutter.  Lines that were
        not extracted (ninl <= 0) become NaN rows, which never match."""
        C, M, t = self.C, self.M, self.tables
        own = line[..., [0, 1, 2, 3, 2, 3]].clone()
        own[ninl <= 0] = float("nan")
        k = min(C, M)
        t["rows"].zero_()
        t["rows"][:, :k] = own[:, :k]
        k2 = min(2 * C, M)
        if k2 > k:
            t["rows"][:, k:k2] = torch.roll(own, 1, dims=0)[:, :k2 - k]
        t["life"].zero_()
        t["life"][:, :k2] = life
        t["ids"].zero_()
        t["ids"][:, :k2] = torch.arange(k2, dtype=torch.int32, device=line.device)
        t["count"].fill_(k2)
        t["next_id"].fill_(k2)

    # ---- end-to-end steps: host buffers in, host buffers out ------------------------------
    def result_keys(self):
        return ("line", "ninl", "match_id", "x", "P")

    def alloc_host_io(self, kind="xy", packets_per_scan=12):
        """Pinned host input buffer for step_host: cartesian scans ("xy": [S, N, 2] float64, what
        scanning() puts on the queue, functions.py:59-66) or raw sensor bytes ("packets":
        [S, P, 84] uint8, what Lidar._read_response returns, lidar.py:161-167)."""
        if kind == "xy":
            return torch.empty((self.S, self.N, 2), dtype=torch.float64).pin_memory()
        if kind == "packets":
            if 32 * packets_per_scan < self.N:
                raise ValueError("packets_per_scan * 32 must cover the scan's beams")
            return torch.empty((self.S, packets_per_scan, 84), dtype=torch.uint8).pin_memory()
        raise ValueError("kind must be 'xy' or 'packets'")

    def _ensure_host_out(self, which=0):
        """Pinned result buffers; set 1 only exists once a caller pipelines steps (step_host(wait=False))."""
        if self._host_out is None:
            self._host_out = [None, None]
        if self._host_out[which] is None:
            shapes = {"line": ((self.S, self.C, 8), torch.float64), "ninl": ((self.S, self.C), torch.int32),
                      "match_id": ((self.S, self.C), torch.int32), "x": ((self.S, 3), torch.float64),
                      "P": ((self.S, 3, 3), torch.float64), "status": ((self.S,), torch.int32)}
            self._host_out[which] = {k: torch.empty(sh, dtype=dt).pin_memory() for k, (sh, dt) in shapes.items()}
        return self._host_out[which]

    def step_host(self, host_in, slices=1, points=False, wait=True):
        """One step from pinned host memory to pinned host memory.

        wait=False (slices > 1, points=False) PIPELINES successive steps: the call returns as soon as the step is queued,
        without making the current stream wait for it, so the next call's uploads run while this step's kernels do (the
        events between the three streams keep every buffer's reuse in order).  The results then carry an extra entry
        "done", a CUDA event recorded after the step's last device->host copy: synchronise it before reading them.  Two
        sets of pinned result buffers alternate, so a step's results stay valid until the call after the next; call
        drain() before timing or before touching the pipeline's state from the current stream.

        host_in: [S, N, 2] float64 scans, or [S, P, 84] uint8 express packets (decoded on the device:
        lslam_express_to_cart).  With slices == 1 everything is queued on the current stream.  With
        slices > 1 the streams are cut into that many contiguous slices and the host->device copies,
        the kernels and the device->host copies of successive slices run on three CUDA streams, so
        the PCIe transfers in both directions overlap the kernels (slice k's kernels wait for slice
        k's input only).  The current stream waits for all three at the end; the caller synchronises
        it before reading the returned pinned tensors (result_keys(), plus "status" per scan for
        packets).  points=True also returns what landmark_extraction itself returns - the projected inlier points
        (ransac_functions.py:28,46-55) - as out["points"] = list of pinned views [rows_k, 2], one per slice, in
        (stream, chunk, data) order: the row count of a slice is only known on the device, so the host waits for
        each slice's count (8 bytes) before it queues that slice's copy; the kernels of later slices keep running."""
        packets = host_in.dtype == torch.uint8
        dev = self.eng.device
        if packets:
            if self.pk_dev is None or self.pk_dev.shape != host_in.shape:
                self.pk_dev = torch.empty(host_in.shape, dtype=torch.uint8, device=dev)
                self.status_dev = torch.zeros((self.S,), dtype=torch.int32, device=dev)
            dev_in = self.pk_dev
        if self.xy_dev is None:
            self.xy_dev = torch.empty((self.S, self.N, 2), dtype=torch.float64, device=dev)
        if not packets:
            dev_in = self.xy_dev
        K = max(1, min(int(slices), self.S))
        if not wait and (K == 1 or points):
            raise ValueError("step_host(wait=False) needs slices > 1 and points=False")
        if not wait:
            self._flip ^= 1
        out = self._ensure_host_out(0 if wait else self._flip)
        keys = self.result_keys()
        bounds = [(self.S * k // K, self.S * (k + 1) // K) for k in range(K)]
        if K == 1:
            dev_in.copy_(host_in, non_blocking=True)
            if packets:
                self.eng.express_to_cart(dev_in, self.N, xy=self.xy_dev, status=self.status_dev)
            res = self.step(self.xy_dev)
            for k in keys:
                out[k].copy_(res[k], non_blocking=True)
            if packets:
                out["status"].copy_(self.status_dev, non_blocking=True)
            if points:
                pts, off = self.project(self.xy_dev, res)
                self._pts["total"][:1].copy_(off[-1:], non_blocking=True)
                torch.cuda.current_stream(dev).synchronize()
                rows = int(self._pts["total"][0])
                self._pts["host"][:rows].copy_(pts[:rows], non_blocking=True)
                out = dict(out, points=[self._pts["host"][:rows]])
            return out
        if self._ov is None or self._ov["K"] != K:
            self._ov = {"K": K, "in": torch.cuda.Stream(dev), "run": torch.cuda.Stream(dev), "dec": torch.cuda.Stream(dev),
                        "ran": torch.cuda.Stream(dev), "ev_ran": [torch.cuda.Event() for _ in range(K)],
                        "out": torch.cuda.Stream(dev), "ev_in": [torch.cuda.Event() for _ in range(K)],
                        "ev_dec": [torch.cuda.Event() for _ in range(K)],
                        "ev_run": [torch.cuda.Event() for _ in range(K)],
                        "ev_out": [torch.cuda.Event() for _ in range(K)], "primed": False}
        ov = self._ov
        cur = torch.cuda.current_stream(dev)
        for st in self._OV_STREAMS:
            ov[st].wait_stream(cur)
        for k, (a, b) in enumerate(bounds):
            with torch.cuda.stream(ov["in"]):
                if ov["primed"]:
                    ov["in"].wait_event(ov["ev_run"][k])      # the previous step's kernels have read this slice
                dev_in[a:b].copy_(host_in[a:b], non_blocking=True)
                ov["ev_in"][k].record()
        for k, (a, b) in enumerate(bounds):
            if packets:
                # the packet decode (store-bound) on a stream of its own: slice k + 1 is decoded while slice k's association
                # and filter kernels run (the RANSAC kernel fills every SM, so it only overlaps that kernel's tail)
                with torch.cuda.stream(ov["dec"]):
                    ov["dec"].wait_event(ov["ev_in"][k])
                    if ov["primed"]:
                        ov["dec"].wait_event(ov["ev_run"][k])  # the previous step's kernels have read this slice's points
                        ov["dec"].wait_event(ov["ev_out"][k])  # ... and its status words have left
                    self.eng.express_to_cart(dev_in[a:b], self.N, xy=self.xy_dev[a:b], status=self.status_dev[a:b])
                    ov["ev_dec"][k].record()
            with torch.cuda.stream(ov["ran"]):
                ov["ran"].wait_event(ov["ev_dec" if packets else "ev_in"][k])
                if ov["primed"]:
                    ov["ran"].wait_event(ov["ev_out"][k])     # the previous step's results have left
                rr = self._step_ransac(self.xy_dev[a:b], a, b)
                ov["ev_ran"][k].record()
            # association + filter on the next stream: they leave most of an SM's warp slots empty, and slice k + 1's line
            # extraction fills in beside them
            with torch.cuda.stream(ov["run"]):
                ov["run"].wait_event(ov["ev_ran"][k])
                res = self._step_post(rr, a, b)
                if points:
                    pts_k, off_k = self.project(self.xy_dev[a:b], res, a, b)
                    self._pts["total"][k:k + 1].copy_(off_k[-1:], non_blocking=True)
                ov["ev_run"][k].record()
            with torch.cuda.stream(ov["out"]):
                ov["out"].wait_event(ov["ev_run"][k])
                for key in keys:
                    out[key][a:b].copy_(res[key], non_blocking=True)
                if packets:
                    out["status"][a:b].copy_(self.status_dev[a:b], non_blocking=True)
                ov["ev_out"][k].record()
        if points:
            views = []
            for k, (a, b) in enumerate(bounds):
                ov["ev_run"][k].synchronize()                  # slice k's row count has landed
                rows = int(self._pts["total"][k])
                with torch.cuda.stream(ov["out"]):
                    self._pts["host"][a * self.N:a * self.N + rows].copy_(self._pts["dev"][a * self.N:a * self.N + rows],
                                                                          non_blocking=True)
                    ov["ev_out"][k].record()
                views.append(self._pts["host"][a * self.N:a * self.N + rows])
            out = dict(out, points=views)
        ov["primed"] = True
        if not wait:
            done = torch.cuda.Event()
            done.record(ov["out"])
            return dict(out, done=done)
        for st in self._OV_STREAMS:
            cur.wait_stream(ov[st])
        return out

    def drain(self):
        """After pipelined steps (step_host(wait=False)): the current stream waits for everything they queued."""
        if self._ov is not None:
            cur = torch.cuda.current_stream(self.eng.device)
            for st in self._OV_STREAMS:
                cur.wait_stream(self._ov[st])

    def io_bytes(self, host_in=None):
        h2d = self.S * self.N * 16 if host_in is None else host_in.numel() * host_in.element_size()
        d2h = 0
        per = {"line": self.C * 64, "ninl": self.C * 4, "match_id": self.C * 4, "x": 24, "P": 72}
        for k in self.result_keys():
            d2h += self.S * per[k]
        if host_in is not None and host_in.dtype == torch.uint8:
            d2h += self.S * 4
        return h2d, d2h

    def launches_per_step(self, host_in=None, slices=1):
        """Kernel launches of this library inside one step (lslam_launch_count difference)."""
        per = 3 + (1 if host_in is not None and host_in.dtype == torch.uint8 else 0)
        return per * max(1, min(int(slices), self.S))
