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
                 nthreads=0):
    """Batched C oracle; same argument meaning and output layout as lslam_ransac_lines."""
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
    rc = lib().oracle_ransac_lines(
        _p(xy, c_double_p), _p(npts, c_int32_p), ctypes.c_int32(B), ctypes.c_int32(N),
        ctypes.c_int32(chunk), ctypes.c_int32(trials), ctypes.c_double(thr),
        ctypes.c_uint64(seed), ctypes.c_uint64(scan_base),
        _p(out["line"], c_double_p), _p(out["ninl"], c_int32_p),
        _p(out["win_trial"], c_int32_p), _p(out["win_pair"], c_int32_p),
        _p(out["ssum"], c_double_p), _p(out["inl_mask"], c_uint32_p),
        ctypes.c_int32(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_ransac_lines failed: %d" % rc)
    return out


def max_threads():
    return int(lib().oracle_max_threads())
