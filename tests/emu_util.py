"""Test helper: build and drive tests/emu/liblslam_emu.so, the kernel sources of
lidar_slam_b200/csrc compiled for the CPU with tests/emu/cuda_emu.h (test infrastructure:
lets the CPU-only suite execute the real kernel code; never used by the product)."""
import ctypes
import os
import subprocess

import numpy as np

from lidar_slam_b200 import _lib

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
EMU_SO = os.path.join(_HERE, "emu", "liblslam_emu.so")
CSRC = os.path.join(_ROOT, "lidar_slam_b200", "csrc")


def build_emu(force=False):
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))]
    srcs += [os.path.join(_HERE, "emu", "cuda_emu.h"), os.path.join(_ROOT, "include", "lslam.h")]
    def stale():
        return force or not os.path.exists(EMU_SO) or any(os.path.getmtime(s) > os.path.getmtime(EMU_SO) for s in srcs)
    if stale():
        # one builder at a time (pytest-xdist workers start together), and the library appears under its name only when
        # it is complete: compile to a temporary name, then rename
        import fcntl
        with open(EMU_SO + ".lock", "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            if stale():
                gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
                tmp = "%s.%d.tmp" % (EMU_SO, os.getpid())
                subprocess.check_call([
                    gxx, "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared",
                    "-Wall", "-Wno-unknown-pragmas", "-include", os.path.join(_HERE, "emu", "cuda_emu.h"),
                    "-x", "c++", os.path.join(CSRC, "lslam.cu"), "-o", tmp, "-lpthread"])
                os.replace(tmp, EMU_SO)
    return EMU_SO


_HANDLE = None
WANT_SSUM = True      # default of emu_ransac_lines(want_ssum=None); the RANSAC tests run both ways


def emu_handle():
    global _HANDLE
    if _HANDLE is None:
        _HANDLE = _lib.Handle(0, path=build_emu())
    return _HANDLE


def ptr(a):
    return None if a is None else a.ctypes.data


def emu_ransac_lines(xy, npts=None, chunk=100, trials=100, thr=20.0, seed=0, scan_base=0, handle=None,
                     want_ssum=None, pairs=None):
    """lslam_ransac_lines on the emulated library with NumPy buffers standing in for device
    memory.  Same outputs as oracle.cport.ransac_lines."""
    h = handle or emu_handle()
    xy = np.ascontiguousarray(xy, dtype=np.float64)
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    W = (chunk + 31) // 32
    if npts is not None:
        npts = np.ascontiguousarray(npts, dtype=np.int32)
    out = {
        "line": np.full((B, C, 8), -7.0, np.float64),
        "ninl": np.full((B, C), -77, np.int32),
        "win_trial": np.full((B, C), -77, np.int32),
        "win_pair": np.full((B, C, 2), -77, np.int32),
        "ssum": np.full((B, C), -7.0, np.float64),
        "inl_mask": np.full((B, C, W), 0xdeadbeef, np.uint32),
    }
    if not (WANT_SSUM if want_ssum is None else want_ssum):
        del out["ssum"]
    h.ransac_lines(ptr(xy), ptr(npts), B, N, chunk, trials, thr, seed, scan_base,
                   ptr(out["line"]), ptr(out["ninl"]), ptr(out["win_trial"]), ptr(out["win_pair"]),
                   ptr(out.get("ssum")), ptr(out["inl_mask"]),
                   pairs=ptr(None if pairs is None else np.ascontiguousarray(pairs, dtype=np.int32)))
    return out


def emu_project_inliers(xy, r, ab=None, npts=None, chunk=100, compact=True, handle=None):
    """lslam_project_inliers on the emulated library; r = the dict of emu_ransac_lines."""
    h = handle or emu_handle()
    xy = np.ascontiguousarray(xy, dtype=np.float64)
    B, N, _ = xy.shape
    C = (N + chunk - 1) // chunk
    ninl = np.ascontiguousarray(r["ninl"], dtype=np.int32)
    mask = np.ascontiguousarray(r["inl_mask"]).view(np.uint32)
    if ab is None:
        ab, stride = np.ascontiguousarray(r["line"], dtype=np.float64), 8
    else:
        ab, stride = np.ascontiguousarray(ab, dtype=np.float64), 2
    if npts is not None:
        npts = np.ascontiguousarray(npts, dtype=np.int32)
    offsets = None
    if compact:
        offsets = np.zeros(B * C + 1, np.int64)
        offsets[1:] = np.cumsum(np.maximum(ninl.reshape(-1), 0))
        out = np.full((int(offsets[-1]) + 4, 2), -7.0)            # 4 guard rows
    else:
        out = np.full((B, C, chunk, 2), np.nan)
    h.project_inliers(ptr(xy), ptr(npts), B, N, chunk, ptr(mask), ptr(ninl), ptr(ab), stride, ptr(offsets), ptr(out))
    if compact:
        assert (out[-4:] == -7.0).all(), "lslam_project_inliers wrote past the last row"
        return out[:-4], offsets
    return out


def emu_associate(queries, tables, q_valid=None, mode=0, append=True, layout=0, handle=None,
                  tolA=0.1, tolB=10, tol=100, life_init=40, hit_pos=None):
    """lslam_associate on the emulated library.  queries [T,Q,6|8]; tables = dict with rows
    [T,M,6], life [T,M], ids [T,M], count [T], next_id [T] (updated in place)."""
    h = handle or emu_handle()
    queries = np.ascontiguousarray(queries, dtype=np.float64)
    T, Q, qs = queries.shape
    M = tables["rows"].shape[1]
    match = np.full((T, Q), -99, np.int32)
    match_id = np.full((T, Q), -99, np.int32)
    proj = np.full((T, Q, 2), -7.0)
    if q_valid is not None:
        q_valid = np.ascontiguousarray(q_valid, dtype=np.int32)
    h.associate(ptr(queries), qs, layout, ptr(q_valid), T, Q, M, ptr(tables["rows"]), ptr(tables["life"]),
                ptr(tables["ids"]), ptr(tables["count"]), ptr(tables["next_id"]), tolA, tolB, tol,
                life_init, mode, append, ptr(match), ptr(match_id), ptr(proj), ptr(hit_pos))
    return match, match_id, proj


def emu_ukf_step(x, P, u, z, lm_pos, lm_valid=None, valid_min=1, zmode=0, params=None, flags=7,
                 sigmas_f=None, handle=None):
    """lslam_ukf_step on the emulated library (NumPy buffers stand in for device memory).
    Returns (x_new, P_new, status); inputs are copied."""
    from lidar_slam_b200 import batch
    h = handle or emu_handle()
    x = np.array(x, dtype=np.float64, order="C")
    P = np.array(P, dtype=np.float64, order="C")
    F = x.shape[0]
    u = None if u is None else np.ascontiguousarray(u, dtype=np.float64)
    L, zs = 0, 2
    if z is not None:
        z = np.ascontiguousarray(z, dtype=np.float64)
        L, zs = z.shape[1], z.shape[2]
        lm_pos = np.ascontiguousarray(lm_pos, dtype=np.float64)
    if lm_valid is not None:
        lm_valid = np.ascontiguousarray(lm_valid, dtype=np.int32)
    status = np.full(F, -99, np.int32)
    params = batch.Engine.ukf_params() if params is None else list(params)
    h.ukf_step(ptr(x), ptr(P), ptr(u), ptr(z), zs, zmode, ptr(lm_pos) if z is not None else None,
               ptr(lm_valid), valid_min, F, L, params, flags, ptr(status), ptr(sigmas_f))
    return x, P, status


def emu_ukf_aug_step(x, P, u, z, params=None, flags=7, workspace_filters=None, handle=None, ws_fill=0):
    """lslam_ukf_aug_step on the emulated library.  x [F, n], P [F, n, n], z [F, L, 2]; returns copies
    (x_new, P_new, status).  workspace_filters < F exercises the batching."""
    from lidar_slam_b200 import batch
    h = handle or emu_handle()
    x = np.array(x, dtype=np.float64, order="C")
    P = np.array(P, dtype=np.float64, order="C")
    F, n = x.shape
    L = (n - 3) // 2
    u = None if u is None else np.ascontiguousarray(u, dtype=np.float64)
    z = None if z is None else np.ascontiguousarray(z, dtype=np.float64)
    status = np.full(F, -99, np.int32)
    params = batch.Engine.ukf_params() if params is None else list(params)
    nbytes = h.ukf_aug_workspace_bytes(L, workspace_filters or F)
    # guard bands around the scratch (and ws_fill = 0xFF: a workspace full of NaNs) catch out-of-bounds writes and
    # reads of scratch the step did not write first
    guard = 4096
    big = np.full(nbytes + 2 * guard, ws_fill, np.uint8)
    big[:guard] = 0xA5; big[guard + nbytes:] = 0xA5
    ws = big[guard:guard + nbytes]
    h.ukf_aug_step(ptr(x), ptr(P), ptr(u), ptr(z), F, L, params, flags, ptr(status), ptr(ws), nbytes)
    assert (big[:guard] == 0xA5).all() and (big[guard + nbytes:] == 0xA5).all(), "lslam_ukf_aug_step wrote outside its workspace"
    return x, P, status


def emu_ukf_step_gather(x, P, u, z, lm_pos, lm_valid, ids, first_record, peer_state, peer_ids, multicast=False,
                        valid_min=0, zmode=1, params=None, flags=7, handle=None):
    """lslam_ukf_step_gather on the emulated library: peer_state / peer_ids are lists of NumPy arrays
    standing in for the ranks' gather buffers (multicast=True: a single array reached through the
    'multicast' address).  Returns (x_new, P_new, status)."""
    from lidar_slam_b200 import batch
    h = handle or emu_handle()
    x = np.array(x, dtype=np.float64, order="C")
    P = np.array(P, dtype=np.float64, order="C")
    F = x.shape[0]
    u = None if u is None else np.ascontiguousarray(u, dtype=np.float64)
    z = np.ascontiguousarray(z, dtype=np.float64)
    L, zs = z.shape[1], z.shape[2]
    lm_pos = np.ascontiguousarray(lm_pos, dtype=np.float64)
    lm_valid = np.ascontiguousarray(lm_valid, dtype=np.int32)
    ids = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32)
    status = np.full(F, -99, np.int32)
    params = batch.Engine.ukf_params() if params is None else list(params)
    world = len(peer_state)
    kw = ({"mc_state": ptr(peer_state[0]), "mc_ids": ptr(peer_ids[0]) if peer_ids else None} if multicast else
          {"peer_state": [ptr(b) for b in peer_state], "peer_ids": [ptr(b) for b in peer_ids] if peer_ids else None})
    h.ukf_step_gather(ptr(x), ptr(P), ptr(u), ptr(z), zs, zmode, ptr(lm_pos), ptr(lm_valid), valid_min, F, L, params,
                      flags, ptr(status), ptr(ids), 0 if ids is None else ids.shape[1], world, first_record, **kw)
    return x, P, status


def emu_express_to_cart(packets, beams, want_meas=True, handle=None):
    """lslam_express_to_cart on the emulated library."""
    h = handle or emu_handle()
    packets = np.ascontiguousarray(packets, dtype=np.uint8)
    B, P, _ = packets.shape
    xy = np.full((B, beams, 2), -7.0)
    meas = np.full((B, beams, 2), -7.0) if want_meas else None
    status = np.full(B, -99, np.int32)
    h.express_to_cart(ptr(packets), B, P, beams, ptr(xy), ptr(meas), ptr(status))
    return xy, meas, status


def emu_synth_express(B, P=12, seed=0, scan_base=0, noise_mm=5.0, dropout=0.01, handle=None):
    h = handle or emu_handle()
    packets = np.zeros((B, P, 84), np.uint8)
    h.synth_express(ptr(packets), B, P, seed, scan_base, noise_mm, dropout)
    return packets


def emu_ransac_ols(xy, npts=None, window=50, trials=100, thr=3.0, probability=0.99, seed=0, scan_base=0, pairs=None,
                   handle=None):
    """lslam_ransac_ols on the emulated library; same outputs as oracle.cport.ransac_ols."""
    h = handle or emu_handle()
    xy = np.ascontiguousarray(xy, dtype=np.float64)
    B, N, _ = xy.shape
    C = (N + window - 1) // window
    W = (window + 31) // 32
    if npts is not None:
        npts = np.ascontiguousarray(npts, dtype=np.int32)
    if pairs is not None:
        pairs = np.ascontiguousarray(pairs, dtype=np.int32)
    out = {"fit": np.full((B, C, 8), -7.0), "ninl": np.full((B, C), -77, np.int32),
           "n_trials": np.full((B, C), -77, np.int32), "best_trial": np.full((B, C), -77, np.int32),
           "inl_mask": np.full((B, C, W), 0xdeadbeef, np.uint32)}
    h.ransac_ols(ptr(xy), ptr(npts), B, N, window, trials, thr, probability, seed, scan_base, ptr(pairs),
                 ptr(out["fit"]), ptr(out["ninl"]), ptr(out["n_trials"]), ptr(out["best_trial"]), ptr(out["inl_mask"]))
    return out
