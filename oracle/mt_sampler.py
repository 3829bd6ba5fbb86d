"""Sample pairs the way upstream draws them (TEST INFRASTRUCTURE, like everything under oracle/).

skimage.measure.ransac (the call at /root/reference/ransac_functions.py:23-24) draws, per trial,
``random_state.choice(num_samples, min_samples, replace=False)``:
  * scikit-image <= 0.18 (the reference's era: cpython-3.5 bytecode): ``random_state`` is a
    ``np.random.RandomState`` (MT19937) made by ``check_random_state(seed)``  -> ``legacy=True``
  * scikit-image >= 0.19: ``np.random.default_rng(seed)`` (PCG64)                 -> ``legacy=False``
NumPy is here, scikit-image is not; with these pairs a seeded upstream run can be replayed pair for pair
through ``oracle.ransac_line.ransac_line(pairs=...)``, ``oracle.cport.ransac_lines(pairs=...)`` and
``lslam_ransac_lines(pairs=...)``.
"""
import numpy as np


def upstream_pairs(seed, n, trials, legacy=True):
    """[trials, 2] int32: the index pairs a seeded skimage.measure.ransac run draws for an n-point chunk
    (one ``choice(n, 2, replace=False)`` per trial, in order, from ONE generator per ransac() call)."""
    rng = np.random.RandomState(seed) if legacy else np.random.default_rng(seed)
    out = np.empty((trials, 2), np.int32)
    for t in range(trials):
        out[t] = rng.choice(n, 2, replace=False)
    return out


def upstream_pairs_batch(seeds, npts_per_slot, trials, legacy=True):
    """[B, C, trials, 2] for a batch: slot (b, c) is one ransac() call seeded with seeds[b, c] on a chunk of
    npts_per_slot[b, c] points (slots with fewer than 2 points get zeros)."""
    seeds = np.asarray(seeds)
    B, C = seeds.shape
    out = np.zeros((B, C, trials, 2), np.int32)
    for b in range(B):
        for c in range(C):
            n = int(npts_per_slot[b, c])
            if n >= 2:
                out[b, c] = upstream_pairs(int(seeds[b, c]), n, trials, legacy)
    return out
