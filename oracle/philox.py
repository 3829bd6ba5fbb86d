"""Philox4x32-10 counter-based RNG and the 2-point pair sampler (oracle side).

TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference draws its RANSAC samples inside ``skimage.measure.ransac`` with
``random_state=None`` (call site /root/reference/ransac_functions.py:23-24), i.e. an
unseeded MT19937 ``RandomState.choice(N, 2, replace=False)``: the reference has no
reproducible sample sequence at all.  The contract of this build (BASELINE.json
north_star) is a counter-based generator shared bit-for-bit by the
oracle and the CUDA kernels:

    key     = (seed_lo, seed_hi)
    counter = (trial >> 1, chunk, scan_lo, scan_hi)   scan = GLOBAL scan index
    r       = philox4x32_10(counter, key)             one block serves two trials
    (u, v)  = (r[0], r[1]) if trial is even else (r[2], r[3])
    i       = floor(u * N       / 2**32)
    j       = floor(v * (N - 1) / 2**32);  j += (j >= i)        -> i != j, both in [0, N)

The algorithm is Salmon et al., "Parallel random numbers: as easy as 1, 2, 3" (SC'11);
the three known-answer vectors in tests/test_oracle_philox.py are the Random123
``kat_vectors`` entries for philox4x32-10.
"""
import numpy as np

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = np.uint32(0x9E3779B9)
PHILOX_W1 = np.uint32(0xBB67AE85)
_MASK32 = np.uint64(0xFFFFFFFF)
_SH32 = np.uint64(32)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  All inputs broadcastable integer arrays (taken mod 2**32).

    Returns four uint32 arrays.
    """
    c0, c1, c2, c3, k0, k1 = np.broadcast_arrays(
        *[np.asarray(v).astype(np.uint64) & _MASK32 for v in (c0, c1, c2, c3, k0, k1)])
    c0, c1, c2, c3 = (c.copy() for c in (c0, c1, c2, c3))
    k0 = k0.copy()
    k1 = k1.copy()
    for rnd in range(10):
        if rnd > 0:
            k0 = (k0 + np.uint64(PHILOX_W0)) & _MASK32
            k1 = (k1 + np.uint64(PHILOX_W1)) & _MASK32
        p0 = PHILOX_M0 * c0          # < 2**64, no overflow
        p1 = PHILOX_M1 * c2
        hi0, lo0 = p0 >> _SH32, p0 & _MASK32
        hi1, lo1 = p1 >> _SH32, p1 & _MASK32
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
    return (c0.astype(np.uint32), c1.astype(np.uint32),
            c2.astype(np.uint32), c3.astype(np.uint32))


def sample_pairs(seed, scan, chunk, trials, npts):
    """Index pairs (i, j), i != j, for trials 0..trials-1 of one (scan, chunk).

    ``seed`` and ``scan`` are 64-bit unsigned; ``npts`` is the chunk's point count (>= 2).
    Returns int32 arrays ``i[trials]``, ``j[trials]``.
    """
    if npts < 2:
        raise ValueError("need at least 2 points to sample a pair")
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    scan = int(scan) & 0xFFFFFFFFFFFFFFFF
    t = np.arange(trials, dtype=np.uint64)
    r0, r1, r2, r3 = philox4x32_10(t >> np.uint64(1), chunk, scan & 0xFFFFFFFF, scan >> 32,
                                   seed & 0xFFFFFFFF, seed >> 32)
    odd = (t & np.uint64(1)).astype(bool)
    u = np.where(odd, r2, r0).astype(np.uint64)
    v = np.where(odd, r3, r1).astype(np.uint64)
    i = (u * np.uint64(npts)) >> _SH32
    j = (v * np.uint64(npts - 1)) >> _SH32
    j = j + (j >= i).astype(np.uint64)
    return i.astype(np.int32), j.astype(np.int32)
