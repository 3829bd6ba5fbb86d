"""CPU oracle for the LiDAR_SLAM hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or the
thing timed as "the CPU", never as a fallback for the CUDA path.

PARITY UNPINNED for the RANSAC and UKF arithmetic: the reference
(Farofeiro231/LiDAR_SLAM) delegates it to scikit-image (``skimage.measure.ransac``,
``LineModelND``) and filterpy (``UnscentedKalmanFilter``,
``MerweScaledSigmaPoints``), neither vendored, neither version-pinned, neither
installable in the build container, and the reference ships no tests, fixtures
or golden vectors.  The oracle therefore restates the published upstream
algorithms (era: skimage <= 0.15, filterpy 1.4.x) driven exactly as the
reference's call sites drive them (each function cites file:line).

PINNED against the reference itself where the reference runs:
``landmarking.py`` (Landmark, is_equal, life bookkeeping) imports with numpy
only, so ``tests/golden/landmarking_golden.npz`` was generated from the real
module (``tests/golden/make_landmarking_golden.py``) and ``oracle/assoc.py`` is
checked against it.
"""
