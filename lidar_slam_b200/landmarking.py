"""Drop-in for the reference's ``landmarking`` module (/root/reference/landmarking.py).

Same names and meaning: ``LIFE, TOLERANCE_A, TOLERANCE_B, TOLERANCE`` (:3-6), ``Landmark`` with
its getters, life bookkeeping, the two tip/centroid distances and ``is_equal`` (:9-77), and
``landmarks_track`` (:79-82).  The record keeps the reference's attribute names (a, b, id, life,
pos, end, timesObserved) because ``ransac_functions`` and user code read them directly.

The association hot loop does not run here: ``ransac_functions.landmark_extraction`` packs the
Python list into a device table and calls ``lslam_associate`` (csrc/assoc.cuh).  The methods below
are the scalar convenience API the reference exposes; ``table_from_landmarks`` /
``apply_table_to_landmarks`` are the packing helpers between the list and the device layout.
"""
import ctypes
import ctypes.util
import math

import numpy as np

LIFE = 40
TOLERANCE_A = 0.1
TOLERANCE_B = 10
TOLERANCE = 100


if hasattr(math, "fma"):                    # Python >= 3.13
    _fma = math.fma
else:                                       # the C library's correctly rounded fma
    _libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    _libm.fma.restype = ctypes.c_double
    _libm.fma.argtypes = [ctypes.c_double] * 3
    _fma = _libm.fma


class Landmark:
    spec = "line"

    def __init__(self, a, b, ID, x, y, tipX, tipY):
        self.a, self.b, self.id = a, b, ID
        self.life = LIFE
        self.timesObserved = 0
        self.pos = np.array([x, y])          # inlier centroid (LineModelND origin)
        self.end = np.array([tipX, tipY])    # last inlier, projected on the line

    def __str__(self):
        return ("Landmark ID: {}\n(x, y): ({}, {})\nequation: {} * x + {}\n"
                .format(self.id, self.pos[0], self.pos[1], self.a, self.b))

    # -- getters (reference :27-43) --
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

    # -- life cycle (reference :45-55) --
    def observed(self):
        self.timesObserved += 1

    def decrease_life(self):
        """One missed observation.  True when the landmark has expired, otherwise None (the
        reference's truthiness contract, :48-52)."""
        self.life = max(self.life - 1, 0) if self.life > 0 else self.life
        return True if self.life == 0 else None

    def reset_life(self):
        self.life = LIFE

    # -- geometry (reference :57-77) --
    @staticmethod
    def _gap(p, q):
        """np.linalg.norm(p - q) of the reference (:58, :63) with the rounding the device path pins:
        sqrt(fma(vy, vy, vx * vx)) - what NumPy's BLAS dot gives here (tests/golden/landmarking_golden.npz holds a
        vector sitting on the 100 mm gate), so Landmark.is_equal on the host and lslam_associate agree to the bit."""
        vx, vy = float(p[0]) - float(q[0]), float(p[1]) - float(q[1])
        return math.sqrt(_fma(vy, vy, vx * vx))

    def distance_end_origin(self, landmark):
        return self._gap(self.end, landmark.get_pos())

    def distance_origin_end(self, landmark):
        return self._gap(self.pos, landmark.get_end())

    def is_equal(self, landmark):
        """Stored landmark (self) against a new observation: slopes within TOLERANCE_A, intercepts
        within TOLERANCE_B, and one's tip within TOLERANCE of the other's centroid (either way)."""
        if abs(self.a - landmark.get_a()) > TOLERANCE_A or abs(self.b - landmark.get_b()) > TOLERANCE_B:
            return False
        return bool(self.distance_end_origin(landmark) <= TOLERANCE or
                    self.distance_origin_end(landmark) <= TOLERANCE)


def landmarks_track(landmarks):
    """Kept for API parity.  The reference's version (:79-82) compares the bound method
    ``get_life`` with 0, which is never true, so it removes nothing; neither does this."""
    for landmark in list(landmarks):
        if landmark.get_life == 0:          # sic: the method object, as in the reference
            landmarks.remove(landmark)


# ---- list <-> device table (layout of lslam_associate: rows a, b, pos_x, pos_y, end_x, end_y) ----
def table_from_landmarks(landmarks, capacity=None):
    """NumPy arrays (rows [M,6], life [M], ids [M], count) for a list of Landmark.  ``ids`` are the
    list POSITIONS, so the kernel's compacted id column says which objects survived, in order."""
    n = len(landmarks)
    M = max(capacity or n, 1)
    rows = np.zeros((M, 6), dtype=np.float64)
    life = np.zeros(M, dtype=np.int32)
    ids = np.zeros(M, dtype=np.int32)
    for k, lm in enumerate(landmarks):
        rows[k] = (lm.a, lm.b, lm.pos[0], lm.pos[1], lm.end[0], lm.end[1])
        life[k] = lm.life
        ids[k] = k
    return rows, life, ids, n


def apply_table_to_landmarks(landmarks, life, ids, count):
    """Mutate the caller's list IN PLACE to what the kernel left: surviving objects in table order
    with their new life (the reference mutates the same list, ransac_functions.py:36-45)."""
    keep = [landmarks[int(ids[k])] for k in range(int(count))]
    for k, lm in enumerate(keep):
        lm.life = int(life[k])
    landmarks[:] = keep
