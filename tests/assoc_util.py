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


def gate_boundary_streams(T=512, seed=0, tol=100.0):
    """T one-entry tables and one query each whose only open gate, ||end_s - pos_q||, sits within a few ulps of the
    tolerance on either side (the other gate is far away): the decision hangs on the last bit of the rounded norm."""
    rng = np.random.default_rng(seed)
    tables = {"rows": np.zeros((T, 1, 6)), "life": np.full((T, 1), 9, np.int32), "ids": np.zeros((T, 1), np.int32),
              "count": np.ones(T, np.int32), "next_id": np.ones(T, np.int32)}
    queries = np.zeros((T, 1, 6))
    th = rng.uniform(0, 2 * np.pi, T)
    v = tol * np.stack([np.cos(th), np.sin(th)], -1)
    for k in range(T):                                      # nudge the vector by a few ulps in or out
        for _ in range(int(rng.integers(0, 4))):
            v[k] = np.nextafter(v[k], v[k] * (2.0 if rng.random() < 0.5 else 0.5))
    pos_q = rng.uniform(-2000, 2000, (T, 2))
    tables["rows"][:, 0, 0] = 0.5
    tables["rows"][:, 0, 1] = 10.0
    tables["rows"][:, 0, 2:4] = pos_q + 5000.0              # pos_s: far from the query's end
    tables["rows"][:, 0, 4:6] = pos_q + v                   # end_s
    queries[:, 0, 0] = 0.5
    queries[:, 0, 1] = 10.0
    queries[:, 0, 2:4] = pos_q
    queries[:, 0, 4:6] = pos_q - 5000.0
    v[::7] = [60.0, 80.0]                                   # exactly on the gate (3-4-5): || . || == tol, a hit
    tables["rows"][::7, 0, 4:6] = 0.0 + v[::7]
    queries[::7, 0, 2:4] = 0.0
    tables["rows"][::7, 0, 2:4] = 5000.0
    queries[::7, 0, 4:6] = -5000.0
    return queries, None, tables


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


def misaligned_copy(a):
    """A C-contiguous copy of `a` whose first byte sits 8 (float64) or 4 (int32) bytes past a 16-byte boundary: what the
    association kernel's 16-byte asynchronous copies cannot take (it must fall back to plain loads)."""
    a = np.ascontiguousarray(a)
    raw = np.zeros(a.nbytes + 64, np.uint8)
    off = (-raw.ctypes.data) % 16 + a.itemsize
    v = raw[off:off + a.nbytes].view(a.dtype).reshape(a.shape)
    v[...] = a
    assert v.ctypes.data % 16 == a.itemsize and v.flags.c_contiguous
    return v
