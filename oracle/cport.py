"""ctypes loader for the C restatement of the oracle (oracle/*.c -> oracle/liboracle.so).

TEST INFRASTRUCTURE (see oracle/__init__.py).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int32_p = ctypes.POINTER(ctypes.c_int32)
c_uint32_p = ctypes.POINTER(ctypes.c_uint32)


def build(force=False):
    """Compile oracle/liboracle.so with the committed Makefile (gcc only)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c")]
    if force or not os.path.exists(so) or any(
            os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.oracle_max_threads.restype = ctypes.c_int
    return _LIB


def _p(a, ty):
    return a.ctypes.data_as(ty) if a is not None else None


def ransac_lines(xy, npts=None, chunk=100, trials=100, thr=20.0, seed=0, scan_base=0,
                 nthreads=0, pairs=None):
    """Batched C oracle; same argument meaning and output layout as lslam_ransac_lines
    (pairs: None = Philox sampler, or [B, C, trials, 2] int32 explicit sample pairs)."""
    xy = np.ascontiguousarray(xy, dtype=np.float64)
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    W = (chunk + 31) // 32
    if npts is not None:
        npts = np.ascontiguousarray(npts, dtype=np.int32)
    out = {
        "line": np.empty((B, C, 8), np.float64),
        "ninl": np.empty((B, C), np.int32),
        "win_trial": np.empty((B, C), np.int32),
        "win_pair": np.empty((B, C, 2), np.int32),
        "ssum": np.empty((B, C), np.float64),
        "inl_mask": np.empty((B, C, W), np.uint32),
    }
    if pairs is not None:
        pairs = np.ascontiguousarray(pairs, dtype=np.int32)
        assert pairs.shape == (B, C, trials, 2), pairs.shape
    rc = lib().oracle_ransac_lines_pairs(
        _p(xy, c_double_p), _p(npts, c_int32_p), ctypes.c_int32(B), ctypes.c_int32(N),
        ctypes.c_int32(chunk), ctypes.c_int32(trials), ctypes.c_double(thr),
        ctypes.c_uint64(seed), ctypes.c_uint64(scan_base), _p(pairs, c_int32_p),
        _p(out["line"], c_double_p), _p(out["ninl"], c_int32_p),
        _p(out["win_trial"], c_int32_p), _p(out["win_pair"], c_int32_p),
        _p(out["ssum"], c_double_p), _p(out["inl_mask"], c_uint32_p),
        ctypes.c_int32(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_ransac_lines failed: %d" % rc)
    return out


def max_threads():
    return int(lib().oracle_max_threads())


def set_threads(n):
    """OpenMP threads of every later call (a launcher may have exported OMP_NUM_THREADS=1)."""
    lib().oracle_set_threads(ctypes.c_int(int(n)))


# ---- UKF (oracle/ukf.c) ----------------------------------------------------------------------
UKF_PREDICT, UKF_UPDATE, UKF_REDRAW = 1, 2, 4


def ukf_params(dt=0.005, wheel_r=50.0, base_l=200.0, alpha=1e-4, beta=2.0, kappa=0.0,
               var_dist=0.25, var_angle=0.09, Q=None):
    """The 17 doubles oracle_ukf_step / lslam_ukf_step take (defaults: systemClass.py:7-29,
    UKFMethods.py:6-7)."""
    Q = np.eye(3) * 0.001 if Q is None else np.asarray(Q, dtype=np.float64)
    return np.concatenate([[dt, wheel_r, base_l, alpha, beta, kappa, var_dist, var_angle],
                           Q.reshape(9)]).astype(np.float64)


def ukf_step(x, P, u, z, lm_pos, lm_valid=None, valid_min=1, zmode=0, params=None,
             flags=UKF_PREDICT | UKF_UPDATE | UKF_REDRAW, sigmas_f=None, nthreads=0):
    """Batched pinned-order UKF step.  x [F,3], P [F,3,3], u [F,2], z [F,L,zs], lm_pos [F,L,2],
    lm_valid [F,L] int32 or None.  Returns (x_new, P_new, status[F]); inputs are not modified
    (sigmas_f [F,7,3], when given, is read/written in place)."""
    x = np.array(x, dtype=np.float64, order="C")
    P = np.array(P, dtype=np.float64, order="C")
    F = x.shape[0]
    u = None if u is None else np.ascontiguousarray(u, dtype=np.float64)
    if z is not None:
        z = np.ascontiguousarray(z, dtype=np.float64)
        L, zs = z.shape[1], z.shape[2]
        lm_pos = np.ascontiguousarray(lm_pos, dtype=np.float64)
    else:
        L, zs = 0, 2
    if lm_valid is not None:
        lm_valid = np.ascontiguousarray(lm_valid, dtype=np.int32)
    params = ukf_params() if params is None else np.ascontiguousarray(params, dtype=np.float64)
    status = np.zeros(F, np.int32)
    fn = lib().oracle_ukf_step
    fn.restype = ctypes.c_int
    rc = fn(_p(x, c_double_p), _p(P, c_double_p), _p(u, c_double_p), _p(z, c_double_p),
            ctypes.c_int32(zs), ctypes.c_int32(zmode), _p(lm_pos, c_double_p) if z is not None else None,
            _p(lm_valid, c_int32_p), ctypes.c_int32(valid_min), ctypes.c_int32(F), ctypes.c_int32(L),
            _p(params, c_double_p), ctypes.c_int32(flags), _p(status, c_int32_p),
            _p(sigmas_f, c_double_p), ctypes.c_int32(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_ukf_step failed: %d" % rc)
    return x, P, status


# ---- association (oracle/assoc.c) and the whole pipeline -------------------------------------------
def associate(query, tables, q_valid=None, mode=0, append=True, tolA=0.1, tolB=10.0, tol=100.0,
              life_init=40, nthreads=0):
    """Batched C association; same meaning as lslam_associate.  query [T,Q,6|8]; tables = dict of
    rows [T,M,6], life, ids [T,M], count, next_id [T], updated IN PLACE.  Returns dict."""
    query = np.ascontiguousarray(query, dtype=np.float64)
    T, Q, qs = query.shape
    M = tables["rows"].shape[1]
    for k, ty in (("rows", np.float64), ("life", np.int32), ("ids", np.int32), ("count", np.int32), ("next_id", np.int32)):
        assert tables[k].dtype == ty and tables[k].flags.c_contiguous, k
    if q_valid is not None:
        q_valid = np.ascontiguousarray(q_valid, dtype=np.int32)
    out = {"match": np.empty((T, Q), np.int32), "match_id": np.empty((T, Q), np.int32),
           "proj_ab": np.empty((T, Q, 2)), "hit_pos": np.empty((T, Q, 2))}
    fn = lib().oracle_associate
    fn.restype = ctypes.c_int
    rc = fn(_p(query, c_double_p), ctypes.c_int32(qs), ctypes.c_int32(1 if qs == 8 else 0), _p(q_valid, c_int32_p),
            ctypes.c_int32(T), ctypes.c_int32(Q), ctypes.c_int32(M), _p(tables["rows"], c_double_p),
            _p(tables["life"], c_int32_p), _p(tables["ids"], c_int32_p), _p(tables["count"], c_int32_p),
            _p(tables["next_id"], c_int32_p), ctypes.c_double(tolA), ctypes.c_double(tolB), ctypes.c_double(tol),
            ctypes.c_int32(life_init), ctypes.c_int32(mode), ctypes.c_int32(int(bool(append))),
            _p(out["match"], c_int32_p), _p(out["match_id"], c_int32_p), _p(out["proj_ab"], c_double_p),
            _p(out["hit_pos"], c_double_p), ctypes.c_int32(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_associate failed: %d" % rc)
    return out


def empty_tables(T, M):
    return {"rows": np.zeros((T, M, 6)), "life": np.zeros((T, M), np.int32), "ids": np.zeros((T, M), np.int32),
            "count": np.zeros(T, np.int32), "next_id": np.zeros(T, np.int32)}


def ransac_ols(xy, npts=None, window=50, trials=100, thr=3.0, probability=0.99, seed=0, scan_base=0, pairs=None,
               nthreads=0):
    """Batched C oracle of the OLS flavour; argument meaning and outputs of lslam_ransac_ols."""
    xy = np.ascontiguousarray(xy, dtype=np.float64)
    B, N, _ = xy.shape
    C = (N + window - 1) // window
    W = (window + 31) // 32
    if npts is not None:
        npts = np.ascontiguousarray(npts, dtype=np.int32)
    if pairs is not None:
        pairs = np.ascontiguousarray(pairs, dtype=np.int32)
        assert pairs.shape == (B, C, trials, 2)
    out = {"fit": np.empty((B, C, 8), np.float64), "ninl": np.empty((B, C), np.int32),
           "n_trials": np.empty((B, C), np.int32), "best_trial": np.empty((B, C), np.int32),
           "inl_mask": np.empty((B, C, W), np.uint32)}
    rc = lib().oracle_ransac_ols(
        _p(xy, c_double_p), _p(npts, c_int32_p), ctypes.c_int32(B), ctypes.c_int32(N), ctypes.c_int32(window),
        ctypes.c_int32(trials), ctypes.c_double(thr), ctypes.c_double(probability), ctypes.c_uint64(seed),
        ctypes.c_uint64(scan_base), _p(pairs, c_int32_p), _p(out["fit"], c_double_p), _p(out["ninl"], c_int32_p),
        _p(out["n_trials"], c_int32_p), _p(out["best_trial"], c_int32_p), _p(out["inl_mask"], c_uint32_p),
        ctypes.c_int32(nthreads))
    if rc != 0:
        raise ValueError("oracle_ransac_ols: bad arguments")
    return out


def express_to_cart(packets, beams, want_meas=False, nthreads=0):
    """oracle_express_to_cart: packets [B, P, 84] uint8 -> (xy [B, beams, 2], meas or None, status [B])."""
    packets = np.ascontiguousarray(packets, dtype=np.uint8)
    B, P, _ = packets.shape
    xy = np.empty((B, beams, 2), np.float64)
    meas = np.empty((B, beams, 2), np.float64) if want_meas else None
    status = np.empty(B, np.int32)
    lib().oracle_express_to_cart(packets.ctypes.data_as(ctypes.c_void_p), ctypes.c_longlong(B), ctypes.c_int(P),
                                 ctypes.c_int(beams), _p(xy, c_double_p), _p(meas, c_double_p),
                                 _p(status, c_int32_p), ctypes.c_int32(nthreads))
    return xy, meas, status


def pipeline_step(xy, tables, x, P, u=None, chunk=100, trials=100, thr=20.0, seed=0, scan_base=0,
                  params=None, nthreads=0):
    """One step of lidar_slam_b200.pipeline.ScanPipeline on the CPU (all three stages in C, OpenMP
    over scans): RANSAC -> association (tables updated in place) -> UKF predict/update with the
    matched lines as measurements.  Returns (ransac_out, assoc_out, x_new, P_new, status)."""
    r = ransac_lines(xy, chunk=chunk, trials=trials, thr=thr, seed=seed, scan_base=scan_base, nthreads=nthreads)
    a = associate(r["line"], tables, q_valid=r["ninl"], nthreads=nthreads)
    xn, Pn, st = ukf_step(x, P, u, r["line"], a["hit_pos"], a["match"], valid_min=0, zmode=1, params=params,
                          nthreads=nthreads)
    return r, a, xn, Pn, st
