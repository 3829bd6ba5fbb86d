"""Drop-in for the reference's ``ransac_functions`` module (/root/reference/ransac_functions.py).

``landmark_extraction(pointsToBeFitted, landmarkNumber, landmarks)`` keeps the reference's
signature, return value and side effects (:15-59): it consumes ``pointsToBeFitted[0]`` and EMPTIES
the caller's list, mutates ``landmarks`` in place (life decrements, removals with the reference's
skip-after-remove behaviour, life reset on a hit), and returns ``(points, fittedLine,
newLandmark)``.  The arithmetic runs on the GPU through the C ABI: ``lslam_ransac_lines`` replaces
``skimage.measure.ransac(data, LineModelND, ...)`` + slope/intercept/tip (:23-30) and
``lslam_associate`` replaces the is_equal / decrease_life / remove loop (:34-45).  There is no CPU
fallback: without liblslam.so or a CUDA device the call raises.

Differences, all forced by the environment and kept at the edges:
  * samples come from the counter-based Philox stream (seed ``SEED``, scan index =
    ``landmarkNumber``) instead of an unseeded MT19937, so a run is reproducible;
  * the returned points are ``PyQt5.QtCore.QPointF`` when PyQt5 is importable, otherwise an
    ``ndarray[n, 2]`` (headless: no graphics stack on the bench box);
  * the ``print`` calls of the reference (:33,41,51,82) only happen with ``COMPAT_PRINTS = True``.

``check_ransac`` / ``ransac_core`` (:63-120) are the thread/queue glue around it; they are kept
API-compatible.  ``RevolutionAdapter`` is the batched replacement for that glue: it collects one
revolution's chunks and issues ONE launch per revolution.
"""
import threading
import time

import numpy as np

from . import batch
from .landmarking import *  # noqa: F401,F403  (the reference re-exports landmarking, :4)
from .landmarking import LIFE, TOLERANCE, TOLERANCE_A, TOLERANCE_B, Landmark, apply_table_to_landmarks, \
    table_from_landmarks

try:                                   # :6 - optional here
    from PyQt5.QtCore import QPointF
except Exception:                      # headless
    QPointF = None

THRESHOLD = 20      # :9  inlier distance, mm, strict '<'
MAX_TRIALS = 100    # :10
MIN_SAMPLES = 2     # :11
MIN_POINTS = 100    # :12 (unused by the reference too)

SEED = 0            # Philox key of the pair sampler
COMPAT_PRINTS = False
DEVICE = None       # None = current CUDA device


def _say(*a):
    if COMPAT_PRINTS:
        print(*a)


def _mask_to_bool(words, n):
    k = np.arange(n)
    return ((words[k >> 5].astype(np.uint32) >> (k & 31).astype(np.uint32)) & 1).astype(bool)


def landmark_extraction(pointsToBeFitted, landmarkNumber, landmarks):
    import torch
    eng = batch.engine(DEVICE)
    data = np.array(pointsToBeFitted[0][:], dtype=np.float64)
    del pointsToBeFitted[:]                                         # :22
    if MIN_SAMPLES != 2:
        raise ValueError("the CUDA path implements min_samples=2 (LineModelND from two points)")
    if data.ndim != 2 or data.shape[1] != 2 or data.shape[0] <= MIN_SAMPLES:
        # upstream's check and message: min_samples must be smaller than the number of samples (n = 2 included)
        raise ValueError("`min_samples` must be in range (0, <number-of-samples>)")
    n = int(data.shape[0])
    # ---- RANSAC + refit + slope/intercept/tip: one chunk = one slot of a B=1 batch ----
    xy_dev = torch.from_numpy(data[None]).to(eng.device)
    r = eng.ransac_lines(xy_dev, chunk=n, trials=MAX_TRIALS, thr=THRESHOLD, seed=SEED, scan_base=int(landmarkNumber))
    ninl = int(r["ninl"][0, 0])
    if ninl < 2:
        if ninl == batch._lib.LSLAM_ERR_NONFINITE:
            raise ValueError("chunk contains non-finite coordinates")
        # upstream returns (None, None) and the reference dies on `model_robust.params` (:25)
        raise AttributeError("'NoneType' object has no attribute 'params'")
    line = r["line"][0, 0].cpu().numpy()
    a, b, ox, oy, tipX, tipY = line[0], line[1], line[2], line[3], line[6], line[7]
    fittedLine = Landmark(float(a), float(b), landmarkNumber, float(ox), float(oy), float(tipX), float(tipY))
    _say("ransacking...")
    # ---- association against the caller's list (first match in list order, reference-exact) ----
    proj_ab, newLandmark = None, True                              # None: the new line's own a, b (:49, :53)
    if len(landmarks) > 0:
        rows, life, ids, count = table_from_landmarks(landmarks)
        dev = eng.device
        tables = {"rows": torch.from_numpy(rows[None]).to(dev), "life": torch.from_numpy(life[None]).to(dev),
                  "ids": torch.from_numpy(ids[None]).to(dev),
                  "count": torch.tensor([count], dtype=torch.int32, device=dev),
                  "next_id": torch.zeros(1, dtype=torch.int32, device=dev)}
        before = [str(lm) for lm in landmarks] if COMPAT_PRINTS else None
        out = eng.associate(r["line"][:, :1], tables, mode=0, append=False, tolA=TOLERANCE_A, tolB=TOLERANCE_B,
                            tol=TOLERANCE, life_init=LIFE)
        match = int(out["match"][0, 0])
        new_count = int(tables["count"][0])
        new_ids = tables["ids"][0].cpu().numpy()
        if COMPAT_PRINTS:
            for k in sorted(set(range(count)) - set(int(v) for v in new_ids[:new_count])):
                print("Excluded landmark: {}".format(before[k]))
        apply_table_to_landmarks(landmarks, tables["life"][0].cpu().numpy(), new_ids, new_count)
        if match >= 0:                                              # :44-47 the STORED landmark's a, b
            proj_ab = out["proj_ab"]
            newLandmark = False
        else:
            _say("New landmark found! Landmarks: {}".format(len(landmarks)))
    # xBase = data[inliers, 0], yBase = a' * xBase + b' (:28, :46 / :49 / :53) on the device: lslam_project_inliers
    pts, _ = eng.project_inliers(xy_dev, r, ab=proj_ab, chunk=n)
    pts = pts.cpu().numpy()
    if QPointF is not None:
        qPointsList = [QPointF(p[0], p[1]) for p in pts]            # :55
    else:
        qPointsList = pts
    return qPointsList, fittedLine, newLandmark


def _concat(parts):
    return np.concatenate([np.asarray(p, dtype=object if QPointF is not None else np.float64) for p in parts], axis=0)


def check_ransac(pairInliers, tempPoints, allPoints, pointsToBeFitted, landmarks, threadEvent, stop=None):
    """The worker loop of the reference (:63-93), same shared-list protocol: a chunk at
    ``pointsToBeFitted[0]`` is fitted; the sentinel 0 publishes the revolution to ``pairInliers`` /
    ``allPoints`` and sets ``threadEvent``.  ``stop`` (a threading.Event, optional) ends the loop,
    which the reference can only do by being killed."""
    inliersList = []
    landmarkNumber = 0
    while stop is None or not stop.is_set():
        if pointsToBeFitted != []:
            if not (isinstance(pointsToBeFitted[0], int) and pointsToBeFitted[0] == 0):
                tempList, extracted, newLandmark = landmark_extraction(pointsToBeFitted, landmarkNumber, landmarks)
                inliersList.append(tempList)
                if newLandmark:
                    landmarks.append(extracted)
                landmarkNumber += 1
            elif inliersList != []:
                _say("TROQUEEEEEEEEEEEEEEEEEEEEI\n\n\n\n\n")
                pairInliers.append(_concat(inliersList))
                if tempPoints:
                    allPoints.append(_concat(tempPoints))
                threadEvent.set()
                del inliersList[:]
                del pointsToBeFitted[:]
                del tempPoints[:]
            else:
                del pointsToBeFitted[:]
        else:
            time.sleep(0)          # yield instead of a hot spin


def ransac_core(rawPoints, stop=None, sink=None):
    """Process entry of the reference (:97-120): drains the queue of chunks into the worker's list.
    The Qt plotting thread (mainWindow.ploting, :105) is replaced by ``sink(pairInliers, allPoints,
    threadEvent)`` when given (headless: there is no graphics stack here)."""
    pairInliers, pointsToBeFitted, allPoints, tempPoints, landmarks = [], [], [], [], []
    threadEvent = threading.Event()
    stop = stop or threading.Event()
    worker = threading.Thread(target=check_ransac, args=(pairInliers, tempPoints, allPoints, pointsToBeFitted,
                                                         landmarks, threadEvent, stop), daemon=True)
    worker.start()
    if sink is not None:
        threading.Thread(target=sink, args=(pairInliers, allPoints, threadEvent), daemon=True).start()
    try:
        while not stop.is_set():
            time.sleep(0.000005)
            try:
                temp = rawPoints.get(True, 0.05)
            except Exception:
                continue
            pointsToBeFitted.append(temp)
            if not (isinstance(temp, int) and temp == 0):
                tempPoints.append([QPointF(p[0], p[1]) for p in temp] if QPointF is not None
                                  else np.asarray(temp, dtype=np.float64))
    except KeyboardInterrupt:
        pass
    finally:
        stop.set()
    return pairInliers, allPoints, landmarks


class RevolutionAdapter:
    """Batched replacement for the check_ransac glue: feed it what ``scanning``
    puts on the queue (chunks, then the sentinel 0, functions.py:64-76); on the sentinel it runs ONE
    lslam_ransac_lines launch over the revolution's chunks and ONE lslam_associate launch over its
    lines in order, and returns the revolution's projected inlier points [n, 2] (what the reference
    hands to the plot) together with the line records.  The landmark table lives on the device."""

    def __init__(self, capacity=64, chunk=100, device=None):
        import torch
        self.eng = batch.engine(device)
        self.chunk = chunk
        self.tables = self.eng.alloc_tables(1, capacity)
        self.pending = []
        self.scan_index = 0
        self._torch = torch

    def put(self, item):
        if isinstance(item, int) and item == 0:
            return self.flush()
        self.pending.append(np.asarray(item, dtype=np.float64))
        return None

    def flush(self):
        torch = self._torch
        if not self.pending:
            return None
        eng, ck = self.eng, self.chunk
        if any(len(c) > ck for c in self.pending):
            raise ValueError("chunk longer than the adapter's chunk size")
        # a revolution = its chunks back to back, padded per chunk slot
        C = len(self.pending)
        xy = np.zeros((C, ck, 2))
        npts = np.zeros(C, np.int32)
        for c, pts in enumerate(self.pending):
            xy[c, :len(pts)] = pts
            npts[c] = len(pts)
        self.pending = []
        xy_dev, npts_dev = torch.from_numpy(xy).to(eng.device), torch.from_numpy(npts).to(eng.device)
        r = eng.ransac_lines(xy_dev, npts=npts_dev, chunk=ck, trials=MAX_TRIALS, thr=THRESHOLD, seed=SEED,
                             scan_base=self.scan_index)
        self.scan_index += C
        line = r["line"].reshape(1, C, 8)
        a = eng.associate(line, self.tables, q_valid=r["ninl"].reshape(1, C), mode=0, append=True,
                          tolA=TOLERANCE_A, tolB=TOLERANCE_B, tol=TOLERANCE, life_init=LIFE)
        # the revolution's projected inlier points, compacted in chunk order, on the device (lslam_project_inliers)
        pts, _ = eng.project_inliers(xy_dev, r, ab=a["proj_ab"], npts=npts_dev, chunk=ck)
        ninl = r["ninl"].reshape(C).cpu().numpy()
        return {"points": pts.cpu().numpy(), "line": line[0].cpu().numpy(),
                "ninl": ninl, "match": a["match"][0].cpu().numpy(), "match_id": a["match_id"][0].cpu().numpy()}
