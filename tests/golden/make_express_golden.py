"""Generates tests/golden/express_golden.npz from the REAL reference decoder: /root/reference/lidar.py is
imported with a stub standing in for pyserial (its only missing import), and ExpressPacket.decode
(lidar.py:66-91) + Lidar._process_express_scan (lidar.py:179-187) are run on random valid packets.

    python tests/golden/make_express_golden.py        (needs /root/reference; the .npz is committed)
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.modules.setdefault("serial", types.ModuleType("serial"))
sys.path.insert(0, "/root/reference")
import lidar as ref_lidar  # noqa: E402
from oracle import express  # noqa: E402


def main():
    rng = np.random.default_rng(42)
    n = 64
    packets = np.zeros((n, 84), np.uint8)
    dist = np.zeros((n, 32), np.int64)
    dth = np.zeros((n, 32))
    start = np.zeros(n)
    new = np.zeros(n, np.int64)
    ang = np.zeros((n, 32))
    for k in range(n):
        d = rng.integers(0, 1 << 14, 32)
        code = rng.integers(0, 64, 32)                 # all 6-bit codes, incl. the ones >= 32
        sq6 = int(rng.integers(0, 360 * 64))
        raw = express.encode_packet(d, code, sq6, int(rng.integers(0, 2)))
        packets[k] = np.frombuffer(raw, np.uint8)
        dec = ref_lidar.ExpressPacket.decode(raw)
        dist[k], dth[k], new[k], start[k] = dec.distance, dec.angle, dec.new_scan, dec.start_angle
    for k in range(n):
        nxt = start[(k + 1) % n]
        dec = ref_lidar.ExpressPacket.decode(packets[k].tobytes())
        for trame in range(1, 33):
            _, _, a, dd = ref_lidar.Lidar._process_express_scan(None, dec, nxt, trame)
            ang[k, trame - 1] = a
            assert dd == dist[k, trame - 1]
    bad = packets[0].copy()
    bad[10] ^= 1
    for raw in (bad.tobytes(), bytes([0x00]) + packets[0].tobytes()[1:]):
        try:
            ref_lidar.ExpressPacket.decode(raw)
            raise SystemExit("reference accepted a corrupt packet")
        except ValueError:
            pass
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "express_golden.npz")
    np.savez(out, packets=packets, distance=dist, dtheta=dth, start_angle=start, new_scan=new, angle=ang)
    print("wrote", out)


if __name__ == "__main__":
    main()
