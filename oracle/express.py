"""RPLidar express-scan packets: decode, measure reconstruction and a synthetic encoder (oracle side).

TEST INFRASTRUCTURE (see oracle/__init__.py).  PINNED against the reference itself:
/root/reference/lidar.py imports with a stub for pyserial, so tests/golden/express_golden.npz holds
outputs of the REAL ``ExpressPacket.decode`` (lidar.py:66-91) and ``Lidar._process_express_scan``
(lidar.py:179-187) on random valid packets (generator: tests/golden/make_express_golden.py), and
tests/test_oracle_express.py checks this restatement against them bit for bit.

Packet (84 bytes): byte0 = 0xA<<4 | checksum low nibble, byte1 = 0x5<<4 | checksum high nibble,
checksum = XOR of bytes 2..83; bytes 2,3 = start angle in q6 (15 bits) + new-scan flag (bit 7 of
byte 3); then 16 cabins of 5 bytes, two measures each: distance (mm, 14 bits) and an angle
correction in q3 that the reference assembles from 6 bits and sign-folds at bit 4
(``twos_comp(value, 5)``, lidar.py:54-57,86,89 - kept exactly, including values >= 32).
Measure i = 1..32 of a packet: angle = (start + ((next_start - start) % 360) / 32 * i - dtheta_i) % 360,
distance = d_i (lidar.py:185-186).  ``scanning`` then converts with functions.py:59-60.
"""
import numpy as np

PACKET = 84


def twos_comp(val, bits):                          # lidar.py:54-57
    if (val & (1 << (bits - 1))) != 0:
        val = val - (1 << bits)
    return val


def decode_packet(data):
    """ExpressPacket.decode (lidar.py:66-91) -> (distance[32] int, dtheta[32] float, new_scan, start_angle)."""
    p = bytearray(data)
    if (p[0] >> 4) != 0xA or (p[1] >> 4) != 0x5:
        raise ValueError("bad sync")
    checksum = 0
    for b in p[2:]:
        checksum ^= b
    if checksum != (p[0] & 0x0F) + ((p[1] & 0x0F) << 4):
        raise ValueError("bad checksum")
    new_scan = p[3] >> 7
    start = (p[2] + ((p[3] & 0x7F) << 8)) / 64.0
    d, a = [], []
    for i in range(0, 80, 5):
        d.append((p[i + 4] >> 2) + (p[i + 5] << 6))
        a.append(twos_comp((p[i + 8] & 0x0F) + ((p[i + 4] & 0x03) << 4), 5) / 8.0)
        d.append((p[i + 6] >> 2) + (p[i + 7] << 6))
        a.append(twos_comp((p[i + 8] >> 4 & 0x0F) + ((p[i + 6] & 0x03) << 4), 5) / 8.0)
    return np.array(d, np.int64), np.array(a, np.float64), new_scan, start


def measures(dist, dtheta, start, next_start):
    """Lidar._process_express_scan (lidar.py:185-186) for trame = 1..32 -> (angle_deg[32], dist[32])."""
    i = np.arange(1, 33, dtype=np.float64)
    ang = (start + (((next_start - start) % 360) / 32) * i - dtheta) % 360
    return ang, dist.astype(np.float64)


def scan_xy(packets, beams):
    """One revolution given as P packets [P, 84] (the start angle after the last packet is the
    first packet's: steady rotation) -> xy[beams, 2] as scanning() computes it (functions.py:59-60)."""
    P = len(packets)
    dec = [decode_packet(pk) for pk in packets]
    ang, dist = [], []
    for k in range(P):
        a, d = measures(dec[k][0], dec[k][1], dec[k][3], dec[(k + 1) % P][3])
        ang.append(a)
        dist.append(d)
    ang = np.concatenate(ang)[:beams]
    dist = np.concatenate(dist)[:beams]
    t = -ang * (np.pi / 180) + np.pi / 2
    return np.stack([dist * np.cos(t), dist * np.sin(t)], -1)


# ---- synthetic encoder (there is no encoder in the reference; this is the inverse of decode) ----
def encode_packet(dist, dtheta_code, start_q6, new_scan=0):
    """dist[32] ints < 2**14, dtheta_code[32] 6-bit codes as the decoder assembles them, start angle in q6."""
    p = bytearray(PACKET)
    p[2] = start_q6 & 0xFF
    p[3] = ((start_q6 >> 8) & 0x7F) | ((new_scan & 1) << 7)
    for c in range(16):
        i = 5 * c
        d1, d2 = int(dist[2 * c]), int(dist[2 * c + 1])
        c1, c2 = int(dtheta_code[2 * c]), int(dtheta_code[2 * c + 1])
        p[i + 4] = ((d1 & 0x3F) << 2) | ((c1 >> 4) & 0x03)
        p[i + 5] = (d1 >> 6) & 0xFF
        p[i + 6] = ((d2 & 0x3F) << 2) | ((c2 >> 4) & 0x03)
        p[i + 7] = (d2 >> 6) & 0xFF
        p[i + 8] = (c1 & 0x0F) | ((c2 & 0x0F) << 4)
    cs = 0
    for b in p[2:]:
        cs ^= b
    p[0] = 0xA0 | (cs & 0x0F)
    p[1] = 0x50 | (cs >> 4)
    return bytes(p)


def dtheta_from_code(code):
    return twos_comp(int(code), 5) / 8.0
