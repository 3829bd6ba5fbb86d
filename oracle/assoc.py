"""Landmark record and data association (oracle side).

TEST INFRASTRUCTURE (see oracle/__init__.py).

PINNED against the reference itself: /root/reference/landmarking.py imports with NumPy only,
so tests/golden/landmarking_golden.npz holds outputs of the REAL ``Landmark.is_equal`` /
``decrease_life`` / ``reset_life`` (generator: tests/golden/make_landmarking_golden.py) and
tests/test_oracle_assoc.py checks this restatement against them.  The association loop
itself lives in ransac_functions.py (which cannot be imported: skimage, PyQt5) and is
restated from its lines.

Arithmetic order (the definition the CUDA kernel is bit-exact against for match indices):
    is_equal(stored, new):   |a_s - a_n| <= TOLERANCE_A  and  |b_s - b_n| <= TOLERANCE_B
                             and ( ||end_s - pos_n|| <= TOLERANCE  or  ||pos_s - end_n|| <= TOLERANCE )
    ||v|| = sqrt(fma(v_y, v_y, v_x*v_x))
np.linalg.norm of a 2-vector is sqrt(dot(v, v)); in the build container (where the golden
vectors were taken from the real module) NumPy's BLAS dot evaluates it as fma(v_y, v_y, v_x*v_x)
- checked on 50,000 random vectors and by the golden vector that sits exactly on the gate - so
that is the pinned definition.  (On a CPU without FMA the reference itself would round
differently; the difference can only flip a decision that sits within 1 ulp of TOLERANCE.)
"""
import ctypes
import ctypes.util

import numpy as np

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.fma.restype = ctypes.c_double
_libm.fma.argtypes = [ctypes.c_double] * 3


def norm2(vx, vy):
    """||(vx, vy)|| with the pinned rounding: sqrt(fma(vy, vy, vx*vx))."""
    vx = float(vx)
    vy = float(vy)
    return float(np.sqrt(_libm.fma(vy, vy, vx * vx)))

# /root/reference/landmarking.py:3-6
LIFE = 40
TOLERANCE_A = 0.1
TOLERANCE_B = 10
TOLERANCE = 100


class Landmark:
    """landmarking.py:9-77 (same fields, same method names)."""
    spec = "line"

    def __init__(self, a, b, ID, x, y, tipX, tipY):
        self.a = a
        self.b = b
        self.id = ID
        self.life = LIFE
        self.pos = np.array([x, y])
        self.end = np.array([tipX, tipY])
        self.timesObserved = 0

    def get_id(self):
        return self.id

    def get_a(self):
        return self.a

    def get_b(self):
        return self.b

    def get_pos(self):
        return self.pos

    def get_end(self):
        return self.end

    def get_life(self):
        return self.life

    def observed(self):
        self.timesObserved += 1

    def decrease_life(self):                       # :48-52  True iff life reaches 0, else None
        if self.life > 0:
            self.life -= 1
        if self.life == 0:
            return True

    def reset_life(self):                          # :54-55
        self.life = LIFE

    @staticmethod
    def _norm2(v):
        return norm2(v[0], v[1])

    def distance_end_origin(self, landmark):       # :57-59
        return self._norm2(self.end - landmark.get_pos())

    def distance_origin_end(self, landmark):       # :62-64
        return self._norm2(self.pos - landmark.get_end())

    def is_equal(self, landmark):                  # :66-77
        dist_a = abs(self.a - landmark.get_a())
        dist_b = abs(self.b - landmark.get_b())
        d_eo = self.distance_end_origin(landmark)
        d_oe = self.distance_origin_end(landmark)
        if dist_a <= TOLERANCE_A and dist_b <= TOLERANCE_B:
            return bool(d_eo <= TOLERANCE or d_oe <= TOLERANCE)
        return False


def associate(landmarks, fitted, reference_exact=True):
    """The association loop of landmark_extraction (ransac_functions.py:34-54), on a Python
    list that is mutated exactly like the reference's: first match in list order; every
    landmark tested before the hit loses one life; a landmark whose life reaches 0 is removed
    and - the reference's skip bug - the element that slides into its slot is neither tested
    nor aged (``i`` still increments).  ``reference_exact=False`` visits every element.

    Returns (matched_landmark_or_None, index_of_match_in_the_list_after_removals_or_-1).
    On a hit the matched landmark's life is reset (its a, b, pos, end are NOT updated).
    The caller appends ``fitted`` on a miss (ransac_functions.py:75-76).
    """
    i = 0
    equal = False
    while i < len(landmarks) and not equal:
        equal = landmarks[i].is_equal(fitted)
        if not equal:
            if landmarks[i].decrease_life():
                landmarks.remove(landmarks[i])
                if not reference_exact:
                    i -= 1
        i += 1
    if equal:
        landmarks[i - 1].reset_life()
        return landmarks[i - 1], i - 1
    return None, -1


class LandmarkTable:
    """Array form of one stream's landmark list, the layout lslam_associate works on:
    rows [M, 6] = a, b, pos_x, pos_y, end_x, end_y; life[M]; ids[M]; count; next_id."""

    def __init__(self, capacity):
        self.rows = np.zeros((capacity, 6))
        self.life = np.zeros(capacity, np.int32)
        self.ids = np.zeros(capacity, np.int32)
        self.count = 0
        self.next_id = 0

    def to_list(self):
        out = []
        for m in range(self.count):
            r = self.rows[m]
            lm = Landmark(r[0], r[1], int(self.ids[m]), r[2], r[3], r[4], r[5])
            lm.life = int(self.life[m])
            out.append(lm)
        return out

    def from_list(self, lms):
        self.count = len(lms)
        for m, lm in enumerate(lms):
            self.rows[m] = [lm.a, lm.b, lm.pos[0], lm.pos[1], lm.end[0], lm.end[1]]
            self.life[m] = lm.life
            self.ids[m] = lm.id
        self.rows[self.count:] = 0
        self.life[self.count:] = 0
        self.ids[self.count:] = 0


def associate_stream(table, queries, valid=None, append=True, reference_exact=True):
    """Feed one stream's queries [Q, 6] through ``associate`` in order, the way check_ransac
    does (ransac_functions.py:73-78): a miss appends the new landmark with id = the running
    extraction counter, which advances on EVERY extraction.  Returns per query
    (match[Q], match_id[Q], proj_ab[Q, 2]) with match = list index of the hit (after removals),
    -1 miss + appended, -3 miss but table full, -2 query skipped (valid[q] <= 0).
    proj_ab = the (a, b) used to project the inliers (:46 / :49,53)."""
    lms = table.to_list()
    Q = len(queries)
    match = np.full(Q, -2, np.int32)
    match_id = np.full(Q, -1, np.int32)
    proj = np.full((Q, 2), np.nan)
    for q in range(Q):
        if valid is not None and valid[q] <= 0:
            continue
        r = queries[q]
        new = Landmark(r[0], r[1], table.next_id, r[2], r[3], r[4], r[5])
        hit, idx = associate(lms, new, reference_exact)
        if hit is not None:
            match[q], match_id[q], proj[q] = idx, hit.id, (hit.a, hit.b)
        else:
            proj[q] = (new.a, new.b)
            match_id[q] = new.id
            if append and len(lms) < table.rows.shape[0]:
                lms.append(new)
                match[q] = -1
            else:
                match[q] = -1 if not append else -3
        table.next_id += 1
    table.from_list(lms)
    return match, match_id, proj
