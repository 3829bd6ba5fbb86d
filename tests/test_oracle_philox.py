"""Philox4x32-10: Random123 known-answer vectors, NumPy vs C, pair sampler properties."""
import ctypes

import numpy as np

from oracle import philox

# Random123 kat_vectors, philox4x32 10 rounds: (counter[4], key[2]) -> output[4]
KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_kat_numpy():
    for ctr, key, want in KAT:
        got = philox.philox4x32_10(*ctr, *key)
        assert tuple(int(v) for v in got) == want


def test_kat_c(cport):
    lib = cport.lib()
    for ctr, key, want in KAT:
        c = (ctypes.c_uint32 * 4)(*ctr)
        k = (ctypes.c_uint32 * 2)(*key)
        o = (ctypes.c_uint32 * 4)()
        lib.oracle_philox4x32_10(c, k, o)
        assert tuple(o) == want


def test_pairs_distinct_in_range_and_uniformish():
    for n in (2, 3, 60, 100, 410):
        i, j = philox.sample_pairs(seed=12345, scan=7, chunk=3, trials=4000, npts=n)
        assert (i != j).all()
        assert i.min() >= 0 and j.min() >= 0 and i.max() < n and j.max() < n
        if n >= 60:
            hist = np.bincount(np.concatenate([i, j]), minlength=n)
            assert hist.min() > 0.4 * 8000 / n and hist.max() < 1.8 * 8000 / n


def test_pairs_depend_on_every_counter_field():
    base = philox.sample_pairs(1, 2, 3, 64, 100)
    for kw in ({"seed": 2}, {"scan": 3}, {"chunk": 4}, {"seed": 1 + (1 << 32)}, {"scan": 2 + (1 << 32)}):
        args = {"seed": 1, "scan": 2, "chunk": 3, "trials": 64, "npts": 100}
        args.update(kw)
        other = philox.sample_pairs(**args)
        assert not (np.array_equal(base[0], other[0]) and np.array_equal(base[1], other[1]))


def test_two_trials_share_one_block():
    # contract: counter word 0 is trial >> 1; even trials take words 0,1 and odd trials words 2,3
    r = philox.philox4x32_10(5, 9, 77, 0, 42, 0)
    i, j = philox.sample_pairs(seed=42, scan=77, chunk=9, trials=12, npts=100)
    for t, (u, v) in ((10, (r[0], r[1])), (11, (r[2], r[3]))):
        a = (int(u) * 100) >> 32
        b = (int(v) * 99) >> 32
        b += b >= a
        assert (i[t], j[t]) == (a, b)
