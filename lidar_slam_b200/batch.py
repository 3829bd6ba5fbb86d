"""Batched device API: torch tensors in, torch tensors out, kernels through the C ABI.

PyTorch is plumbing only here (device memory, streams): every computation is a call into
liblslam.so (include/lslam.h) with ``tensor.data_ptr()`` and the current CUDA stream.  There
is no CPU or eager fallback: without a CUDA device or without the built library the
constructor raises.
"""
import torch

from . import _lib

# defaults = the reference's module constants (ransac_functions.py:9-11, functions.py:14)
THRESHOLD = 20
MAX_TRIALS = 100
CHUNK = 100


def _dev_index(device):
    if device is None:
        return torch.cuda.current_device()
    d = torch.device(device)
    if d.type != "cuda":
        raise ValueError("lidar_slam_b200 runs on CUDA devices only, got %r" % (device,))
    return d.index if d.index is not None else torch.cuda.current_device()


class Engine:
    """One liblslam handle bound to one CUDA device."""

    def __init__(self, device=None, library_path=None):
        if not torch.cuda.is_available():
            raise RuntimeError("lidar_slam_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.index = _dev_index(device)
        self.device = torch.device("cuda", self.index)
        self.handle = _lib.Handle(self.index, path=library_path)

    # ---- helpers -------------------------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _check(self, t, dtype, name):
        if not isinstance(t, torch.Tensor) or t.device != self.device:
            raise ValueError("%s must be a tensor on %s" % (name, self.device))
        if t.dtype != dtype or not t.is_contiguous():
            raise ValueError("%s must be contiguous %s" % (name, dtype))
        return t

    def launch_count(self):
        return self.handle.launch_count()

    def set_option(self, key, value):
        self.handle.set_option(key, value)

    # ---- entry points --------------------------------------------------------------------
    def polar_to_cart(self, dist_mm, angle_deg):
        """functions.py:59-60 for device tensors of any (equal) shape -> xy[..., 2]."""
        dist_mm = self._check(dist_mm, torch.float64, "dist_mm")
        angle_deg = self._check(angle_deg, torch.float64, "angle_deg")
        if dist_mm.shape != angle_deg.shape:
            raise ValueError("dist_mm and angle_deg must have the same shape")
        xy = torch.empty(dist_mm.shape + (2,), dtype=torch.float64, device=self.device)
        self.handle.polar_to_cart(dist_mm.data_ptr(), angle_deg.data_ptr(), xy.data_ptr(),
                                  dist_mm.numel(), self._stream())
        return xy

    def express_to_cart(self, packets, beams, xy=None, want_meas=False, status=None):
        """Raw RPLidar express packets [B, P, 84] uint8 on this device -> xy [B, beams, 2] (lidar.py:66-91,
        179-187 + functions.py:59-60 on the device).  Returns (xy, meas or None, status [B] int32)."""
        packets = self._check(packets, torch.uint8, "packets")
        if packets.dim() != 3 or packets.shape[2] != 84:
            raise ValueError("packets must be [B, P, 84]")
        B, P = int(packets.shape[0]), int(packets.shape[1])
        if xy is None:
            xy = torch.empty((B, beams, 2), dtype=torch.float64, device=self.device)
        meas = torch.empty((B, beams, 2), dtype=torch.float64, device=self.device) if want_meas else None
        if status is None:
            status = torch.empty(B, dtype=torch.int32, device=self.device)
        self.handle.express_to_cart(packets.data_ptr(), B, P, int(beams), xy.data_ptr(),
                                    meas.data_ptr() if want_meas else None, status.data_ptr(), self._stream())
        return xy, meas, status

    def synth_express(self, B, packets_per_scan=12, seed=0, scan_base=0, noise_mm=5.0, dropout=0.01, out=None):
        """Synthetic sensor bytes [B, P, 84] uint8 (device twin of the random-room scene)."""
        if out is None:
            out = torch.empty((B, packets_per_scan, 84), dtype=torch.uint8, device=self.device)
        self.handle.synth_express(out.data_ptr(), B, packets_per_scan, seed, scan_base, noise_mm, dropout,
                                  self._stream())
        return out

    def alloc_ransac_out(self, B, N, chunk=CHUNK, ssum=True):
        """Output buffers of ransac_lines.  ``ssum=False`` leaves out the winner's residual sum (a
        diagnostic the reference never exposes); the kernel then skips the literal residual pass for
        chunks with an unambiguous winner."""
        C = (N + chunk - 1) // chunk
        W = (chunk + 31) // 32
        dev = self.device
        out = {
            "line": torch.empty((B, C, 8), dtype=torch.float64, device=dev),
            "ninl": torch.empty((B, C), dtype=torch.int32, device=dev),
            "win_trial": torch.empty((B, C), dtype=torch.int32, device=dev),
            "win_pair": torch.empty((B, C, 2), dtype=torch.int32, device=dev),
            "ssum": torch.empty((B, C), dtype=torch.float64, device=dev),
            "inl_mask": torch.empty((B, C, W), dtype=torch.int32, device=dev),
        }
        if not ssum:
            del out["ssum"]
        return out

    def ransac_lines(self, xy, npts=None, chunk=CHUNK, trials=MAX_TRIALS, thr=THRESHOLD,
                     seed=0, scan_base=0, out=None, pairs=None):
        """Batched landmark extraction (ransac_functions.py:20-31 per chunk).

        xy [B, N, 2] float64 on this device; npts [B] int32 or None; pairs None (Philox sampler) or
        [B, C, trials, 2] int32 explicit sample pairs (what upstream's RandomState drew).  Returns the
        dict of lslam_ransac_lines outputs (include/lslam.h); ``inl_mask`` words are int32 bit
        patterns.  Asynchronous on the current stream.
        """
        xy = self._check(xy, torch.float64, "xy")
        if xy.dim() != 3 or xy.shape[2] != 2:
            raise ValueError("xy must be [B, N, 2]")
        B, N = int(xy.shape[0]), int(xy.shape[1])
        if npts is not None:
            npts = self._check(npts, torch.int32, "npts")
            if npts.numel() != B:
                raise ValueError("npts must have B elements")
        if out is None:
            out = self.alloc_ransac_out(B, N, chunk)
        if pairs is not None:
            pairs = self._check(pairs, torch.int32, "pairs")
            if tuple(pairs.shape) != (B, (N + chunk - 1) // chunk, int(trials), 2):
                raise ValueError("pairs must be [B, C, trials, 2]")
        self.handle.ransac_lines(
            xy.data_ptr(), npts.data_ptr() if npts is not None else None, B, N, int(chunk),
            int(trials), float(thr), seed, scan_base, out["line"].data_ptr(), out["ninl"].data_ptr(),
            out["win_trial"].data_ptr(), out["win_pair"].data_ptr(),
            out["ssum"].data_ptr() if "ssum" in out else None, out["inl_mask"].data_ptr(), self._stream(),
            pairs=pairs.data_ptr() if pairs is not None else None)
        return out

    def project_inliers(self, xy, r, ab=None, npts=None, chunk=CHUNK, compact=True, out=None):
        """The projected inlier points landmark_extraction returns (ransac_functions.py:28,46-55), on the device:
        for every chunk slot with a model, (x, a' x + b') of its inliers in data order.  ``r`` = the dict of
        ransac_lines; ``ab`` = ``proj_ab`` of associate ([.., 2], the reference's choice of coefficients) or None for
        the slots' own lines.  compact=True returns (points [total, 2], offsets [B*C + 1] int64); compact=False
        returns points [B, C, chunk, 2] (rows >= ninl of a slot are NaN)."""
        xy = self._check(xy, torch.float64, "xy")
        B, N = int(xy.shape[0]), int(xy.shape[1])
        C = (N + chunk - 1) // chunk
        ninl = self._check(r["ninl"], torch.int32, "ninl")
        mask = self._check(r["inl_mask"], torch.int32, "inl_mask")
        if ab is None:
            ab, stride = self._check(r["line"], torch.float64, "line"), 8
        else:
            ab, stride = self._check(ab, torch.float64, "ab"), 2
            if ab.numel() != B * C * 2:
                raise ValueError("ab must hold [B*C, 2] coefficients")
        if npts is not None:
            npts = self._check(npts, torch.int32, "npts")
        offsets = None
        if compact:
            offsets = torch.zeros(B * C + 1, dtype=torch.int64, device=self.device)
            torch.cumsum(ninl.reshape(-1).clamp(min=0), 0, out=offsets[1:])
            if out is None:
                out = torch.empty((int(offsets[-1].item()), 2), dtype=torch.float64, device=self.device)
        elif out is None:
            out = torch.full((B, C, chunk, 2), float("nan"), dtype=torch.float64, device=self.device)
        self.handle.project_inliers(xy.data_ptr(), npts.data_ptr() if npts is not None else None, B, N, int(chunk),
                                    mask.data_ptr(), ninl.data_ptr(), ab.data_ptr(), stride,
                                    offsets.data_ptr() if offsets is not None else None, out.data_ptr(), self._stream())
        return (out, offsets) if compact else out

    def ransac_ols(self, xy, npts=None, window=50, trials=100, thr=3.0, stop_probability=0.99, seed=0,
                   scan_base=0, pairs=None):
        """Batched OLS-flavour RANSAC (line_detection.py:51-73: RANSACRegressor(residual_threshold=3) per
        50-point window).  xy [B, N, 2] float64 on this device; pairs None (Philox sampler) or
        [B, C, trials, 2] int32 explicit sample pairs.  Returns the dict of lslam_ransac_ols outputs."""
        xy = self._check(xy, torch.float64, "xy")
        if xy.dim() != 3 or xy.shape[2] != 2:
            raise ValueError("xy must be [B, N, 2]")
        B, N = int(xy.shape[0]), int(xy.shape[1])
        C = (N + window - 1) // window
        W = (window + 31) // 32
        if npts is not None:
            npts = self._check(npts, torch.int32, "npts")
            if npts.numel() != B:
                raise ValueError("npts must have B elements")
        if pairs is not None:
            pairs = self._check(pairs, torch.int32, "pairs")
            if tuple(pairs.shape) != (B, C, trials, 2):
                raise ValueError("pairs must be [B, C, trials, 2]")
        dev = self.device
        out = {"fit": torch.empty((B, C, 8), dtype=torch.float64, device=dev),
               "ninl": torch.empty((B, C), dtype=torch.int32, device=dev),
               "n_trials": torch.empty((B, C), dtype=torch.int32, device=dev),
               "best_trial": torch.empty((B, C), dtype=torch.int32, device=dev),
               "inl_mask": torch.empty((B, C, W), dtype=torch.int32, device=dev)}
        self.handle.ransac_ols(xy.data_ptr(), npts.data_ptr() if npts is not None else None, B, N, int(window),
                               int(trials), thr, stop_probability, seed, scan_base,
                               pairs.data_ptr() if pairs is not None else None, out["fit"].data_ptr(),
                               out["ninl"].data_ptr(), out["n_trials"].data_ptr(), out["best_trial"].data_ptr(),
                               out["inl_mask"].data_ptr(), self._stream())
        return out

    def ransac_counters(self):
        return self.handle.ransac_counters(self._stream())

    # landmarking.py:3-6
    LIFE = 40
    TOLERANCE_A = 0.1
    TOLERANCE_B = 10
    TOLERANCE = 100

    def alloc_tables(self, T, M):
        """Empty landmark tables for T independent streams, capacity M each (lslam_associate)."""
        dev = self.device
        return {
            "rows": torch.zeros((T, M, 6), dtype=torch.float64, device=dev),
            "life": torch.zeros((T, M), dtype=torch.int32, device=dev),
            "ids": torch.zeros((T, M), dtype=torch.int32, device=dev),
            "count": torch.zeros((T,), dtype=torch.int32, device=dev),
            "next_id": torch.zeros((T,), dtype=torch.int32, device=dev),
        }

    def alloc_assoc_out(self, T, Q):
        dev = self.device
        return {
            "match": torch.empty((T, Q), dtype=torch.int32, device=dev),
            "match_id": torch.empty((T, Q), dtype=torch.int32, device=dev),
            "proj_ab": torch.empty((T, Q, 2), dtype=torch.float64, device=dev),
            "hit_pos": torch.empty((T, Q, 2), dtype=torch.float64, device=dev),
        }

    def associate(self, query, tables, q_valid=None, mode=0, append=True, tolA=TOLERANCE_A,
                  tolB=TOLERANCE_B, tol=TOLERANCE, life_init=LIFE, out=None):
        """Batched data association (ransac_functions.py:34-54 + :74-78, landmarking.py:48-77).

        query [T, Q, 6] (a, b, pos, end) or [T, Q, 8] (``line`` records of ransac_lines);
        tables = dict from alloc_tables, updated IN PLACE; q_valid [T, Q] int32 or None (e.g.
        ``ninl`` of ransac_lines).  Returns dict match / match_id / proj_ab (include/lslam.h).
        """
        query = self._check(query, torch.float64, "query")
        if query.dim() != 3 or query.shape[2] not in (6, 8):
            raise ValueError("query must be [T, Q, 6] or [T, Q, 8]")
        T, Q, qs = (int(v) for v in query.shape)
        rows = self._check(tables["rows"], torch.float64, "tables['rows']")
        if rows.dim() != 3 or rows.shape[0] != T or rows.shape[2] != 6:
            raise ValueError("tables['rows'] must be [T, M, 6]")
        M = int(rows.shape[1])
        life = self._check(tables["life"], torch.int32, "tables['life']")
        ids = self._check(tables["ids"], torch.int32, "tables['ids']")
        count = self._check(tables["count"], torch.int32, "tables['count']")
        next_id = self._check(tables["next_id"], torch.int32, "tables['next_id']")
        if life.shape != (T, M) or ids.shape != (T, M) or count.numel() != T or next_id.numel() != T:
            raise ValueError("table field shapes do not match [T, M]")
        if q_valid is not None:
            q_valid = self._check(q_valid, torch.int32, "q_valid")
            if q_valid.numel() != T * Q:
                raise ValueError("q_valid must be [T, Q]")
        if out is None:
            out = self.alloc_assoc_out(T, Q)
        self.handle.associate(
            query.data_ptr(), qs, 1 if qs == 8 else 0, q_valid.data_ptr() if q_valid is not None else None,
            T, Q, M, rows.data_ptr(), life.data_ptr(), ids.data_ptr(), count.data_ptr(), next_id.data_ptr(),
            tolA, tolB, tol, life_init, mode, append, out["match"].data_ptr(), out["match_id"].data_ptr(),
            out["proj_ab"].data_ptr(), out["hit_pos"].data_ptr() if "hit_pos" in out else None,
            stream=self._stream())
        return out

    # ---- UKF (systemClass.py:7-29, UKFMethods.py:6-8) ----------------------------------------
    UKF_PREDICT, UKF_UPDATE, UKF_REDRAW = 1, 2, 4

    @staticmethod
    def ukf_params(dt=0.005, wheel_r=50.0, base_l=200.0, alpha=1e-4, beta=2.0, kappa=0.0,
                   var_dist=0.5 ** 2, var_angle=0.3 ** 2, Q=None):
        """The 17 doubles lslam_ukf_step takes; defaults = systemClass.py:8-10,20,28-29 (DT, VAR_DIST,
        VAR_ANGLE, MerweScaledSigmaPoints(alpha=1e-4, beta=2, kappa=0), Q = 0.001 I) and
        UKFMethods.py:6-7 (R = 50, L = 200)."""
        q = [0.001, 0, 0, 0, 0.001, 0, 0, 0, 0.001] if Q is None else [float(v) for row in Q for v in row]
        return [dt, wheel_r, base_l, alpha, beta, kappa, var_dist, var_angle] + q

    def alloc_ukf(self, F, L=8):
        """State of F filters as systemClass.config_ukf sets it (x = 0, P = diag(.1, .1, .05))."""
        dev = self.device
        P0 = torch.diag(torch.tensor([0.1, 0.1, 0.05], dtype=torch.float64, device=dev))
        return {
            "x": torch.zeros((F, 3), dtype=torch.float64, device=dev),
            "P": P0.expand(F, 3, 3).contiguous(),
            "u": torch.zeros((F, 2), dtype=torch.float64, device=dev),
            "status": torch.zeros((F,), dtype=torch.int32, device=dev),
            "params": self.ukf_params(),
            "L": L,
        }

    def ukf_step(self, ukf, z, lm_pos, lm_valid=None, valid_min=1, u=None, zmode=None, flags=None,
                 sigmas_f=None, gather=None):
        """predict(u) + update(z, landmarks) of every filter in ``ukf`` (dict of alloc_ukf), in place.

        z [F, L, 2] (range, bearing) or [F, L, 8] line records of ransac_lines (zmode 1: the
        measurement is the range/bearing of the line's inlier centroid); lm_pos [F, L, 2];
        lm_valid [F, L] int32 (slot used iff >= valid_min) or None."""
        x = self._check(ukf["x"], torch.float64, "ukf['x']")
        P = self._check(ukf["P"], torch.float64, "ukf['P']")
        F = int(x.shape[0])
        if x.shape != (F, 3) or P.shape != (F, 3, 3):
            raise ValueError("ukf['x'] must be [F, 3] and ukf['P'] [F, 3, 3]")
        u = ukf.get("u") if u is None else u
        if u is not None:
            u = self._check(u, torch.float64, "u")
            if u.shape != (F, 2):
                raise ValueError("u must be [F, 2]")
        if flags is None:
            flags = self.UKF_PREDICT | self.UKF_UPDATE | self.UKF_REDRAW
        L, zs = 0, 2
        if z is not None:
            z = self._check(z, torch.float64, "z")
            if z.dim() != 3 or z.shape[0] != F:
                raise ValueError("z must be [F, L, zstride]")
            L, zs = int(z.shape[1]), int(z.shape[2])
            if zmode is None:
                zmode = 1 if zs == 8 else 0
            lm_pos = self._check(lm_pos, torch.float64, "lm_pos")
            if lm_pos.shape != (F, L, 2):
                raise ValueError("lm_pos must be [F, L, 2]")
            if lm_valid is not None:
                lm_valid = self._check(lm_valid, torch.int32, "lm_valid")
                if lm_valid.shape != (F, L):
                    raise ValueError("lm_valid must be [F, L]")
        st = ukf.get("status")
        common = (x.data_ptr(), P.data_ptr(), u.data_ptr() if u is not None else None,
                  z.data_ptr() if z is not None else None, zs, zmode or 0,
                  lm_pos.data_ptr() if z is not None else None, lm_valid.data_ptr() if lm_valid is not None else None,
                  valid_min, F, L, ukf["params"], flags, st.data_ptr() if st is not None else None)
        if gather is None:
            self.handle.ukf_step(*common, sigmas_f.data_ptr() if sigmas_f is not None else None, self._stream())
        else:
            # fused all-gather over peer memory: gather = dict(world, first_record, ids, id_slots, and either
            # mc_state/mc_ids (multicast addresses) or peer_state/peer_ids (lists of device addresses));
            # built by lidar_slam_b200/dist.PeerGather.descriptor
            ids = gather.get("ids")
            if ids is not None:
                ids = self._check(ids, torch.int32, "gather ids")
            self.handle.ukf_step_gather(*common, ids.data_ptr() if ids is not None else None,
                                        int(gather.get("id_slots", 0)), int(gather["world"]), int(gather["first_record"]),
                                        gather.get("peer_state"), gather.get("mc_state"), gather.get("peer_ids"),
                                        gather.get("mc_ids"), self._stream())
        return ukf

    # ---- augmented-state filters (pose + L landmarks in the state; BASELINE.json configs[3]) ----
    def alloc_ukf_aug(self, F, L, workspace_filters=None, workspace_bytes=None):
        """F filters of dimension n = 3 + 2L: x = 0, P = diag(.1, .1, .05, 25, 25, ...) (pose prior of
        systemClass.config_ukf, 5 mm landmark sigma), and the scratch lslam_ukf_aug_step needs.
        The scratch holds ``workspace_filters`` filters at a time (default: all F, capped at ~1/4 of the
        device's memory); more filters are processed in batches."""
        dev = self.device
        n = 3 + 2 * L
        per = self.handle.ukf_aug_workspace_bytes(L, 1)
        if workspace_bytes is None:
            k = F if workspace_filters is None else int(workspace_filters)
            cap = torch.cuda.get_device_properties(dev).total_memory // 4 if dev.type == "cuda" else per * k
            k = max(1, min(k, cap // per))
            workspace_bytes = self.handle.ukf_aug_workspace_bytes(L, k)
        d = torch.tensor([0.1, 0.1, 0.05] + [25.0] * (2 * L), dtype=torch.float64, device=dev)
        return {
            "x": torch.zeros((F, n), dtype=torch.float64, device=dev),
            "P": torch.diag(d).expand(F, n, n).contiguous(),
            "u": torch.zeros((F, 2), dtype=torch.float64, device=dev),
            "status": torch.zeros((F,), dtype=torch.int32, device=dev),
            "params": self.ukf_params(),
            "workspace": torch.empty((workspace_bytes,), dtype=torch.uint8, device=dev),
            "L": L,
        }

    def ukf_aug_step(self, ukf, z, u=None, flags=None):
        """predict(u) + update(z) of every augmented filter in ``ukf`` (dict of alloc_ukf_aug), in place.
        z [F, L, 2]: (range, bearing) of every landmark of the state."""
        L = int(ukf["L"])
        n = 3 + 2 * L
        x = self._check(ukf["x"], torch.float64, "ukf['x']")
        P = self._check(ukf["P"], torch.float64, "ukf['P']")
        F = int(x.shape[0])
        if x.shape != (F, n) or P.shape != (F, n, n):
            raise ValueError("ukf['x'] must be [F, 3 + 2L] and ukf['P'] [F, 3 + 2L, 3 + 2L]")
        u = ukf.get("u") if u is None else u
        if u is not None:
            u = self._check(u, torch.float64, "u")
            if u.shape != (F, 2):
                raise ValueError("u must be [F, 2]")
        if flags is None:
            flags = self.UKF_PREDICT | self.UKF_UPDATE | self.UKF_REDRAW
        if z is not None:
            z = self._check(z, torch.float64, "z")
            if z.shape != (F, L, 2):
                raise ValueError("z must be [F, L, 2]")
        st = self._check(ukf["status"], torch.int32, "ukf['status']")
        ws = ukf["workspace"]
        self.handle.ukf_aug_step(x.data_ptr(), P.data_ptr(), u.data_ptr() if u is not None else None,
                                 z.data_ptr() if z is not None else None, F, L, ukf["params"], flags, st.data_ptr(),
                                 ws.data_ptr(), ws.numel(), self._stream())
        return ukf

    def synth_scans(self, B, beams=360, seed=0, scan_base=0, noise_mm=5.0, dropout=0.02, out=None):
        """Random-room scans generated on the device (csrc/aux_kernels.cuh) -> xy[B, beams, 2]."""
        if out is None:
            out = torch.empty((B, beams, 2), dtype=torch.float64, device=self.device)
        self.handle.synth_scans(out.data_ptr(), B, beams, seed, scan_base, noise_mm, dropout,
                                self._stream())
        return out

    def fp64_peak(self, ms_target=50.0):
        """Measured dependent-chain DFMA rate of this device, TFLOP/s (roofline denominator)."""
        return self.handle.fp64_peak(ms_target)


_ENGINES = {}


def engine(device=None):
    """Process-wide Engine per device, created on first use (after fork, SLAM.py:20)."""
    idx = _dev_index(device)
    if idx not in _ENGINES:
        _ENGINES[idx] = Engine(idx)
    return _ENGINES[idx]
