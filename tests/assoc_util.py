"""Shared generators / checkers for the association tests (oracle vs kernel)."""
import numpy as np

from oracle import assoc


def random_streams(T, Q, M, seed=0, prefill=0, short_lives=True):
    """T independent streams: landmark tables (optionally pre-filled) and Q queries each, drawn
    around a few base lines per stream so hits, misses, expiries and removals all occur."""
    rng = np.random.default_rng(seed)
    tables = {"rows": np.zeros((T, M, 6)), "life": np.zeros((T, M), np.int32), "ids": np.zeros((T, M), np.int32),
              "count": np.zeros(T, np.int32), "next_id": np.zeros(T, np.int32)}
    queries = np.zeros((T, Q, 6))
    valid = np.ones((T, Q), np.int32)
    for t in range(T):
        nb = int(rng.integers(2, 7))
        base = [(rng.uniform(-1, 1), rng.uniform(-1500, 1500), rng.uniform(-500, 500)) for _ in range(nb)]

        def draw():
            a0, b0, x0 = base[rng.integers(0, nb)]
            a = a0 + rng.normal(0, 0.04)
            b = b0 + rng.normal(0, 5.0)
            px = x0 + rng.uniform(-60, 60)
            ex = px + rng.uniform(-150, 150)
            return [a, b, px, a * px + b, ex, a * ex + b]
        n0 = min(prefill, M)
        for m in range(n0):
            tables["rows"][t, m] = draw()
            tables["life"][t, m] = rng.integers(1, 5) if short_lives else assoc.LIFE
            tables["ids"][t, m] = m
        tables["count"][t] = n0
        tables["next_id"][t] = n0
        for q in range(Q):
            queries[t, q] = draw()
        valid[t] = rng.random(Q) > 0.1
    return queries, valid, tables


def oracle_associate(queries, valid, tables, mode=0, append=True):
    """oracle/assoc.py over every stream; returns outputs and the updated tables (copies)."""
    T, Q, _ = queries.shape
    M = tables["rows"].shape[1]
    out = {k: v.copy() for k, v in tables.items()}
    match = np.zeros((T, Q), np.int32)
    mid = np.zeros((T, Q), np.int32)
    proj = np.zeros((T, Q, 2))
    for t in range(T):
        tb = assoc.LandmarkTable(M)
        tb.rows[:] = tables["rows"][t]
        tb.life[:] = tables["life"][t]
        tb.ids[:] = tables["ids"][t]
        tb.count = int(tables["count"][t])
        tb.next_id = int(tables["next_id"][t])
        match[t], mid[t], proj[t] = assoc.associate_stream(tb, queries[t], valid[t] if valid is not None else None,
                                                           append=append, reference_exact=(mode == 0))
        out["rows"][t], out["life"][t], out["ids"][t] = tb.rows, tb.life, tb.ids
        out["count"][t], out["next_id"][t] = tb.count, tb.next_id
    return match, mid, proj, out


def assert_assoc_equal(want, got, what=""):
    wm, wi, wp, wt = want
    gm, gi, gp, gt = got
    assert np.array_equal(wm, gm), "%s match indices differ" % what
    assert np.array_equal(wi, gi), "%s match ids differ" % what
    assert np.array_equal(np.isnan(wp), np.isnan(gp)) and np.array_equal(np.nan_to_num(wp), np.nan_to_num(gp)), \
        "%s projection coefficients differ" % what
    for k in ("count", "next_id", "life", "ids", "rows"):
        assert np.array_equal(wt[k], gt[k]), "%s table field %s differs" % (what, k)
