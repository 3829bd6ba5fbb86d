"""ctypes binding of the C ABI (include/lslam.h) - the only door to the CUDA kernels.

The library is built in-tree by ``lidar_slam_b200/_build.py`` (nvcc, sm_100a) as
``lidar_slam_b200/csrc/liblslam.so``.  There is NO fallback: if the library is missing or a
call fails, this module raises.  All pointer arguments are raw device addresses (ints), as
the ABI takes them; tensor plumbing lives in ``lidar_slam_b200/batch.py``.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_PATH = os.path.join(_HERE, "csrc", "liblslam.so")

LSLAM_OK = 0
LSLAM_ERR_BAD_ARG = -1
LSLAM_ERR_TOO_FEW_POINTS = -2
LSLAM_ERR_NO_MODEL = -3
LSLAM_ERR_CUDA = -4
LSLAM_ERR_NONFINITE = -5
LSLAM_ERR_CAPACITY = -6
ABI_VERSION = 6

_vp = ctypes.c_void_p
_i32 = ctypes.c_int32
_i64 = ctypes.c_int64
_u64 = ctypes.c_uint64
_f64 = ctypes.c_double

# name -> (restype, argtypes); mirrors include/lslam.h one to one
SIGNATURES = {
    "lslam_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(_vp)]),
    "lslam_destroy": (ctypes.c_int, [_vp]),
    "lslam_last_error": (ctypes.c_char_p, [_vp]),
    "lslam_abi_version": (ctypes.c_int, []),
    "lslam_set_option": (ctypes.c_int, [_vp, ctypes.c_char_p, _i64]),
    "lslam_launch_count": (_i64, [_vp]),
    "lslam_polar_to_cart": (ctypes.c_int, [_vp, _vp, _vp, _vp, _i64, _vp]),
    "lslam_ransac_lines": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _f64, _u64, _u64, _vp,
                                          _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "lslam_project_inliers": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "lslam_ransac_counters": (ctypes.c_int, [_vp, ctypes.POINTER(_i64), _vp]),
    "lslam_associate": (ctypes.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp,
                                       _f64, _f64, _f64, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "lslam_ukf_step": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32,
                                      ctypes.POINTER(_f64), _i32, _vp, _vp, _vp]),
    "lslam_ukf_step_gather": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32,
                                             ctypes.POINTER(_f64), _i32, _vp, _vp, _i32, _i32, _i64,
                                             ctypes.POINTER(_vp), _vp, ctypes.POINTER(_vp), _vp, _vp]),
    "lslam_ukf_predict": (ctypes.c_int, [_vp, _vp, _vp, _vp, _i32, ctypes.POINTER(_f64), _i32, _vp, _vp, _vp]),
    "lslam_ukf_update": (ctypes.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32,
                                        ctypes.POINTER(_f64), _vp, _vp, _vp]),
    "lslam_ukf_aug_workspace_bytes": (ctypes.c_int, [_i32, _i32, ctypes.POINTER(_i64)]),
    "lslam_ukf_aug_step": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, ctypes.POINTER(_f64), _i32, _vp, _vp,
                                          _i64, _vp]),
    "lslam_ukf_aug_step_static": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, ctypes.POINTER(_f64), _i32, _vp, _vp]),
    "lslam_ransac_ols": (ctypes.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _f64, _f64, _u64, _u64,
                                        _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "lslam_express_to_cart": (ctypes.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp, _vp]),
    "lslam_synth_express": (ctypes.c_int, [_vp, _vp, _i64, _i32, _u64, _u64, _f64, _f64, _vp]),
    "lslam_synth_scans": (ctypes.c_int, [_vp, _vp, _i32, _i32, _u64, _u64, _f64, _f64, _vp]),
    "lslam_fp64_peak": (ctypes.c_int, [_vp, _f64, ctypes.POINTER(_f64)]),
}


class LslamError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("liblslam error %d: %s" % (code, text))
        self.code = code


def load(path=None):
    """dlopen the C-ABI library and attach the signatures.  Raises if it is absent."""
    path = path or os.environ.get("LSLAM_LIBRARY") or DEFAULT_PATH
    if not os.path.exists(path):
        raise ImportError(
            "%s not found: build it with `python -m lidar_slam_b200._build` "
            "(nvcc, sm_100a). There is no CPU fallback." % path)
    cdll = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(cdll, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if cdll.lslam_abi_version() != ABI_VERSION:
        raise ImportError("%s has ABI version %d, binding expects %d"
                          % (path, cdll.lslam_abi_version(), ABI_VERSION))
    return cdll


class Handle:
    """One lslam handle (one device).  Methods take raw addresses and mirror the ABI."""

    def __init__(self, device=0, path=None, cdll=None):
        self.cdll = cdll if cdll is not None else load(path)
        self.device = device
        h = _vp()
        rc = self.cdll.lslam_create(device, ctypes.byref(h))
        if rc != LSLAM_OK or not h.value:
            raise LslamError(rc, "lslam_create(device=%d) failed (no usable CUDA device?)" % device)
        self._h = h

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.cdll.lslam_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != LSLAM_OK:
            raise LslamError(rc, self.cdll.lslam_last_error(self._h).decode("utf-8", "replace"))

    # -- plumbing --
    def set_option(self, key, value):
        self._check(self.cdll.lslam_set_option(self._h, key.encode(), int(value)))

    def launch_count(self):
        return int(self.cdll.lslam_launch_count(self._h))

    # -- entry points --
    def polar_to_cart(self, dist, ang_deg, xy, npts, stream=0):
        self._check(self.cdll.lslam_polar_to_cart(self._h, dist, ang_deg, xy, npts, stream))

    def ransac_ols(self, xy, npts, B, N, window, trials, thr, stop_probability, seed, scan_base, pairs,
                   fit, ninl, ntrials, best_trial, inl_mask, stream=0):
        self._check(self.cdll.lslam_ransac_ols(
            self._h, xy, npts, B, N, window, trials, float(thr), float(stop_probability),
            int(seed) & (2**64 - 1), int(scan_base) & (2**64 - 1), pairs, fit, ninl, ntrials, best_trial,
            inl_mask, stream))

    def express_to_cart(self, packets, nscans, packets_per_scan, beams, xy, meas=None, status=None, stream=0):
        self._check(self.cdll.lslam_express_to_cart(self._h, packets, nscans, packets_per_scan, beams, xy, meas,
                                                    status, stream))

    def synth_express(self, packets, nscans, packets_per_scan, seed, scan_base, noise_mm, dropout, stream=0):
        self._check(self.cdll.lslam_synth_express(self._h, packets, nscans, packets_per_scan,
                                                  int(seed) & (2**64 - 1), int(scan_base) & (2**64 - 1),
                                                  float(noise_mm), float(dropout), stream))

    def ransac_lines(self, xy, npts, B, N, chunk, trials, thr, seed, scan_base,
                     line, ninl, win_trial, win_pair, ssum, inl_mask, stream=0, pairs=None):
        self._check(self.cdll.lslam_ransac_lines(
            self._h, xy, npts, B, N, chunk, trials, float(thr), int(seed) & (2**64 - 1),
            int(scan_base) & (2**64 - 1), pairs, line, ninl, win_trial, win_pair, ssum, inl_mask, stream))

    def project_inliers(self, xy, npts, B, N, chunk, inl_mask, ninl, ab, ab_stride, offsets, points, stream=0):
        self._check(self.cdll.lslam_project_inliers(self._h, xy, npts, B, N, chunk, inl_mask, ninl, ab, ab_stride, offsets,
                                                    points, stream))

    def ransac_counters(self, stream=0):
        out = (_i64 * 3)()
        self._check(self.cdll.lslam_ransac_counters(self._h, out, stream))
        return {"literal_rescored": int(out[0]), "extra_candidates": int(out[1]), "faults": int(out[2])}

    def associate(self, query, qstride, layout, q_valid, T, Q, M, table, life, ids, count, next_id,
                  tolA, tolB, tol, life_init, mode, append, match, match_id, proj_ab, hit_pos=None, stream=0):
        self._check(self.cdll.lslam_associate(
            self._h, query, qstride, layout, q_valid, T, Q, M, table, life, ids, count, next_id,
            float(tolA), float(tolB), float(tol), int(life_init), int(mode), int(bool(append)),
            match, match_id, proj_ab, hit_pos, stream))

    @staticmethod
    def _params(params):
        if len(params) != 17:
            raise ValueError("UKF params: 17 doubles (dt, wheel_r, base_l, alpha, beta, kappa, var_dist, "
                             "var_angle, Q[9])")
        return (_f64 * 17)(*[float(v) for v in params])

    def ukf_step(self, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params, flags,
                 status=None, sigmas_f=None, stream=0):
        self._check(self.cdll.lslam_ukf_step(self._h, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min,
                                             F, L, self._params(params), flags, status, sigmas_f, stream))

    def ukf_step_gather(self, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params, flags,
                        status, ids, id_slots, world, first_record, peer_state=None, mc_state=None, peer_ids=None,
                        mc_ids=None, stream=0):
        """peer_state / peer_ids: sequences of `world` device addresses (or None with a multicast address)."""
        def arr(ptrs):
            if ptrs is None:
                return None
            return (_vp * world)(*[int(p) for p in ptrs])
        self._check(self.cdll.lslam_ukf_step_gather(
            self._h, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, self._params(params), flags,
            status, ids, id_slots, world, first_record, arr(peer_state), mc_state, arr(peer_ids), mc_ids, stream))

    def ukf_predict(self, x, P, u, F, params, redraw=True, status=None, sigmas_f=None, stream=0):
        self._check(self.cdll.lslam_ukf_predict(self._h, x, P, u, F, self._params(params), int(bool(redraw)),
                                                status, sigmas_f, stream))

    def ukf_update(self, x, P, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params,
                   status=None, sigmas_f=None, stream=0):
        self._check(self.cdll.lslam_ukf_update(self._h, x, P, z, zstride, zmode, lm_pos, lm_valid, valid_min,
                                               F, L, self._params(params), status, sigmas_f, stream))

    def ukf_aug_workspace_bytes(self, L, filters):
        out = _i64(0)
        if self.cdll.lslam_ukf_aug_workspace_bytes(int(L), int(filters), ctypes.byref(out)) != LSLAM_OK:
            raise LslamError(-1, "lslam_ukf_aug_workspace_bytes: need 1 <= L <= 1024, filters >= 1")
        return int(out.value)

    def ukf_aug_step(self, x, P, u, z, F, L, params, flags, status, workspace, workspace_bytes, stream=0):
        self._check(self.cdll.lslam_ukf_aug_step(self._h, x, P, u, z, F, L, self._params(params), flags, status,
                                                 workspace, int(workspace_bytes), stream))

    def ukf_aug_step_static(self, x, P, u, z, F, L, params, flags, status, stream=0):
        """lslam_ukf_aug_step with the scratch in the library's static device array (no allocation)."""
        self._check(self.cdll.lslam_ukf_aug_step_static(self._h, x, P, u, z, F, L, self._params(params), flags, status,
                                                        stream))

    def synth_scans(self, xy, B, beams, seed, scan_base, noise_mm, dropout, stream=0):
        self._check(self.cdll.lslam_synth_scans(self._h, xy, B, beams, int(seed) & (2**64 - 1),
                                                int(scan_base) & (2**64 - 1), float(noise_mm),
                                                float(dropout), stream))

    def fp64_peak(self, ms_target=50.0):
        out = _f64(0.0)
        self._check(self.cdll.lslam_fp64_peak(self._h, float(ms_target), ctypes.byref(out)))
        return float(out.value)
