"""Shared set-up for the pipeline tests: the CPU twin of ScanPipeline.install_map_from_lines and the
multi-step CPU pipeline (oracle/cport.pipeline_step)."""
import numpy as np

from oracle import cport


def map_from_lines(line, ninl, M, life=40):
    S, C, _ = line.shape
    own = line[..., [0, 1, 2, 3, 2, 3]].copy()
    own[ninl <= 0] = np.nan
    t = cport.empty_tables(S, M)
    k = min(C, M)
    t["rows"][:, :k] = own[:, :k]
    k2 = min(2 * C, M)
    if k2 > k:
        t["rows"][:, k:k2] = np.roll(own, 1, axis=0)[:, :k2 - k]
    t["life"][:, :k2] = life
    t["ids"][:, :k2] = np.arange(k2, dtype=np.int32)
    t["count"][:] = k2
    t["next_id"][:] = k2
    return t


def cpu_pipeline(xy, steps, M=8, u=None, seed=0, scan_base=0, chunk=100, trials=100, with_map=True):
    """`steps` pipeline steps over the same scans (as bench.py does).  Returns per-step results."""
    S = xy.shape[0]
    r0 = cport.ransac_lines(xy, chunk=chunk, trials=trials, seed=seed, scan_base=scan_base)
    tables = map_from_lines(r0["line"], r0["ninl"], M) if with_map else cport.empty_tables(S, M)
    x = np.zeros((S, 3))
    P = np.tile(np.diag([0.1, 0.1, 0.05]), (S, 1, 1))
    out = []
    for _ in range(steps):
        r, a, x, P, st = cport.pipeline_step(xy, tables, x, P, u, chunk=chunk, trials=trials, seed=seed,
                                             scan_base=scan_base)
        out.append({"ransac": r, "assoc": a, "x": x.copy(), "P": P.copy(), "status": st,
                    "tables": {k: v.copy() for k, v in tables.items()}})
    return out
