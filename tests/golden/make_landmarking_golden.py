"""Generates tests/golden/landmarking_golden.npz from the REAL reference module
/root/reference/landmarking.py (imports with NumPy only).  It pins the oracle's Landmark
restatement (oracle/assoc.py) against the reference itself:
  - is_equal on 2000 random (stored, new) pairs drawn around every gate of the test
  - the life bookkeeping: decrease_life return values / life after k calls, reset_life
  - a replay of the association loop of ransac_functions.py:34-54 (restated here line by line,
    since that file cannot be imported) driven with REAL Landmark objects, including the
    remove-while-iterating skip
Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_landmarking_golden.py
"""
import importlib.util
import os

import numpy as np

spec = importlib.util.spec_from_file_location("ref_landmarking", "/root/reference/landmarking.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

rng = np.random.default_rng(2024)
N = 2000
stored = np.empty((N, 6))
new = np.empty((N, 6))
equal = np.empty(N, bool)
for k in range(N):
    a = rng.uniform(-3, 3)
    b = rng.uniform(-2000, 2000)
    pos = rng.uniform(-3000, 3000, 2)
    end = pos + rng.uniform(-800, 800, 2)
    s = ref.Landmark(a, b, k, pos[0], pos[1], end[0], end[1])
    # perturb around the gates: TOLERANCE_A = .1, TOLERANCE_B = 10, TOLERANCE = 100
    da = rng.choice([0.0, 0.05, 0.0999999, 0.1, 0.1000001, 0.3]) * rng.choice([-1, 1])
    db = rng.choice([0.0, 5.0, 9.9999, 10.0, 10.0001, 40.0]) * rng.choice([-1, 1])
    mode = rng.integers(0, 4)
    ang = rng.uniform(0, 2 * np.pi)
    rad = rng.choice([0.0, 50.0, 99.999, 100.0, 100.001, 400.0])
    off = rad * np.array([np.cos(ang), np.sin(ang)])
    npos = rng.uniform(-3000, 3000, 2)
    nend = rng.uniform(-3000, 3000, 2)
    if mode == 0:
        npos = end + off          # end_s close to pos_n
    elif mode == 1:
        nend = pos + off          # pos_s close to end_n
    elif mode == 2:
        npos = end + off
        nend = pos - off
    n = ref.Landmark(a + da, b + db, -1, npos[0], npos[1], nend[0], nend[1])
    stored[k] = [s.a, s.b, s.pos[0], s.pos[1], s.end[0], s.end[1]]
    new[k] = [n.a, n.b, n.pos[0], n.pos[1], n.end[0], n.end[1]]
    equal[k] = bool(s.is_equal(n))

# life bookkeeping
lm = ref.Landmark(0, 0, 0, 0, 0, 1, 1)
life_trace, ret_trace = [], []
for k in range(ref.LIFE + 3):
    r = lm.decrease_life()
    ret_trace.append(-1 if r is None else int(bool(r)))
    life_trace.append(lm.get_life())
lm.reset_life()
life_after_reset = lm.get_life()


def reference_loop(landmarks, fitted):
    """ransac_functions.py:34-47 verbatim in structure (prints dropped)."""
    i = 0
    equal = False
    if len(landmarks) > 0:
        while i < len(landmarks) and not equal:
            equal = landmarks[i].is_equal(fitted)
            if not equal:
                deleteLandmark = landmarks[i].decrease_life()
                if deleteLandmark:
                    landmarks.remove(landmarks[i])
            i += 1
        if equal:
            landmarks[i - 1].reset_life()
            return i - 1
    return -1


# association replay: 30 streams of 60 extractions, low initial lives so removals happen
streams = []
for s in range(30):
    landmarks = []
    number = 0
    base = [(rng.uniform(-1, 1), rng.uniform(-1500, 1500), rng.uniform(-500, 500)) for _ in range(6)]
    qs, res, counts, lives = [], [], [], []
    for step in range(60):
        a0, b0, x0 = base[rng.integers(0, len(base))]
        a = a0 + rng.normal(0, 0.04)
        b = b0 + rng.normal(0, 5.0)
        px = x0 + rng.uniform(-60, 60)
        py = a * px + b
        ex = px + rng.uniform(-150, 150)
        ey = a * ex + b
        f = ref.Landmark(a, b, number, px, py, ex, ey)
        idx = reference_loop(landmarks, f)
        if idx < 0:
            landmarks.append(f)
            f.life = int(rng.integers(1, 6))      # short lives: exercise expiry and the skip bug
        number += 1
        qs.append([a, b, px, py, ex, ey])
        res.append(idx)
        counts.append(len(landmarks))
        lives.append([l.life for l in landmarks] + [-1] * (32 - len(landmarks)))
        # remember the life assigned on append so the replay can do the same
        qs[-1].append(f.life if idx < 0 else -1)
    streams.append((np.array(qs), np.array(res), np.array(counts), np.array(lives)))

path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "landmarking_golden.npz")
np.savez_compressed(
    path, stored=stored, new=new, equal=equal,
    consts=np.array([ref.LIFE, ref.TOLERANCE_A, ref.TOLERANCE_B, ref.TOLERANCE]),
    life_trace=np.array(life_trace), ret_trace=np.array(ret_trace), life_after_reset=life_after_reset,
    stream_q=np.stack([s[0] for s in streams]), stream_match=np.stack([s[1] for s in streams]),
    stream_count=np.stack([s[2] for s in streams]), stream_lives=np.stack([s[3] for s in streams]))
print("wrote", path, "equal fraction", equal.mean(), "hits", np.mean(np.stack([s[1] for s in streams]) >= 0))
