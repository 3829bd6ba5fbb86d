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
    def __init__(self, eng, streams, beams=360, chunk=batch.CHUNK, trials=batch.MAX_TRIALS,
                 thr=batch.THRESHOLD, table_capacity=8, seed=0, scan_base=0):
        self.eng = eng
        self.S, self.N, self.chunk, self.trials, self.thr = int(streams), int(beams), int(chunk), int(trials), thr
        self.C = (self.N + self.chunk - 1) // self.chunk
        self.M = int(table_capacity)
        self.seed, self.scan_base = seed, scan_base
        self.rout = eng.alloc_ransac_out(self.S, self.N, self.chunk)
        self.tables = eng.alloc_tables(self.S, self.M)
        self.aout = eng.alloc_assoc_out(self.S, self.C)
        self.ukf = eng.alloc_ukf(self.S, self.C)
        self.xy_dev = None
        self._host_out = None

    # ---- device-resident step: inputs already in HBM ------------------------------------
    def step(self, xy, u=None, ransac_events=None):
        """xy [S, N, 2] float64 on the device.  Asynchronous on the current stream.  Returns the
        dict of result tensors (views of buffers owned by the pipeline, overwritten next step).
        ``ransac_events`` = (start, end) torch.cuda.Event pair recorded around the RANSAC launch."""
        eng = self.eng
        if ransac_events is not None:
            ransac_events[0].record()
        r = eng.ransac_lines(xy, chunk=self.chunk, trials=self.trials, thr=self.thr, seed=self.seed,
                             scan_base=self.scan_base, out=self.rout)
        if ransac_events is not None:
            ransac_events[1].record()
        a = eng.associate(r["line"], self.tables, q_valid=r["ninl"], out=self.aout)
        # measurement slot c = chunk c's line: range/bearing of its inlier centroid against the matched
        # stored landmark's position; slots whose line missed (match < 0) are not used
        eng.ukf_step(self.ukf, r["line"], a["hit_pos"], lm_valid=a["match"], valid_min=0, u=u)
        return {"line": r["line"], "ninl": r["ninl"], "match": a["match"], "match_id": a["match_id"],
                "x": self.ukf["x"], "P": self.ukf["P"]}

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

    # ---- end-to-end step: host buffers in, host buffers out -----------------------------
    def result_keys(self):
        return ("line", "ninl", "match_id", "x", "P")

    def alloc_host_io(self):
        """Pinned host buffers for step_host: the scans in, the per-step results out."""
        xy_h = torch.empty((self.S, self.N, 2), dtype=torch.float64).pin_memory()
        return xy_h

    def step_host(self, xy_host):
        """xy_host [S, N, 2] float64 in pinned host memory -> dict of pinned host result tensors.
        Host->device copy, the three kernels and the device->host copies are queued on the current
        stream; the caller synchronises."""
        if self.xy_dev is None:
            self.xy_dev = torch.empty((self.S, self.N, 2), dtype=torch.float64, device=self.eng.device)
        self.xy_dev.copy_(xy_host, non_blocking=True)
        res = self.step(self.xy_dev)
        if self._host_out is None:
            self._host_out = {k: torch.empty(res[k].shape, dtype=res[k].dtype).pin_memory()
                              for k in self.result_keys()}
        for k in self.result_keys():
            self._host_out[k].copy_(res[k], non_blocking=True)
        return self._host_out

    def io_bytes(self):
        h2d = self.S * self.N * 16
        d2h = 0
        per = {"line": self.C * 64, "ninl": self.C * 4, "match_id": self.C * 4, "x": 24, "P": 72}
        for k in self.result_keys():
            d2h += self.S * per[k]
        return h2d, d2h
