"""Headless synthetic scan generator (host side, NumPy).

Replaces the capture process for tests and benchmarks: the reference reads an RPLidar over
/dev/ttyUSB0 (/root/reference/functions.py:47-56, lidar.py) which the bench box does not
have.  Output format is exactly what ``scanning`` hands to the hot path: sensor-frame
cartesian points in millimetres,

    dX = d * cos(-angle_deg * pi/180 + pi/2),  dY = d * sin(-angle_deg * pi/180 + pi/2)

(/root/reference/functions.py:59-60), no range filter (a dropped return, d = 0, becomes
the point (0, 0)), one revolution cut into chunks of MIN_NEIGHBOORS = 100 points
(functions.py:14,64).  The CUDA twin used by bench.py to fill 1M-scan batches on the
device is ``lslam_synth_scans`` (csrc/synth.cu): same scene model, same Philox streams,
device libm, so the two agree to rounding, not bit for bit.
"""
import numpy as np

ANGLE_TO_RAD = np.pi / 180          # functions.py:13
MIN_NEIGHBOORS = 100                # functions.py:14


def polar_to_cart(dist_mm, angle_deg):
    """functions.py:59-60 for arrays."""
    dist_mm = np.asarray(dist_mm, dtype=np.float64)
    ang = -np.asarray(angle_deg, dtype=np.float64) * ANGLE_TO_RAD + np.pi / 2
    return np.stack([dist_mm * np.cos(ang), dist_mm * np.sin(ang)], axis=-1)


def room_ranges(pose, room_w_mm, room_h_mm, angle_deg):
    """Range (mm) from ``pose`` = (x, y, heading_rad) to the walls of an axis-aligned
    room [-w/2, w/2] x [-h/2, h/2] along each beam (sensor angles as the RPLidar reports
    them: degrees, clockwise from the sensor's +y axis)."""
    px, py, th = pose
    ang = -np.asarray(angle_deg, dtype=np.float64) * ANGLE_TO_RAD + np.pi / 2 + th
    ux, uy = np.cos(ang), np.sin(ang)
    with np.errstate(divide="ignore", invalid="ignore"):
        tx = np.where(ux > 0, (room_w_mm / 2 - px) / ux,
                      np.where(ux < 0, (-room_w_mm / 2 - px) / ux, np.inf))
        ty = np.where(uy > 0, (room_h_mm / 2 - py) / uy,
                      np.where(uy < 0, (-room_h_mm / 2 - py) / uy, np.inf))
    return np.minimum(tx, ty)


def room_scan(pose=(300.0, -200.0, np.deg2rad(10.0)), room_w_mm=4000.0, room_h_mm=3000.0,
              beams=360, noise_mm=5.0, dropout=0.0, rng=None):
    """One revolution in a rectangular room -> xy[beams, 2] (sensor frame, mm).

    Defaults are BASELINE.json configs[0] / SURVEY.md 8(d) C1: 360 beams at 1 degree, robot
    at (0.3 m, -0.2 m, 10 deg) in a 4 m x 3 m room, N(0, 5 mm) range noise,
    ``np.random.default_rng(0)``.
    """
    if rng is None:
        rng = np.random.default_rng(0)
    angle = np.arange(beams, dtype=np.float64) * (360.0 / beams)
    d = room_ranges(pose, room_w_mm, room_h_mm, angle)
    d = d + noise_mm * rng.standard_normal(beams)
    if dropout > 0:
        d = np.where(rng.random(beams) < dropout, 0.0, d)
    return polar_to_cart(d, angle)


def random_room_scans(nscans, beams=360, seed=0, noise_mm=5.0, dropout=0.02):
    """SURVEY.md 8(d) C2-style batch: room size ~U(2, 6 m), pose ~U(interior, 0.3 m from
    the walls), heading ~U(-pi, pi), 2 % dropouts.  Returns xy[nscans, beams, 2]."""
    out = np.empty((nscans, beams, 2), dtype=np.float64)
    for s in range(nscans):
        rng = np.random.default_rng([seed, s])
        w, h = rng.uniform(2000.0, 6000.0, size=2)
        px = rng.uniform(-w / 2 + 300.0, w / 2 - 300.0)
        py = rng.uniform(-h / 2 + 300.0, h / 2 - 300.0)
        th = rng.uniform(-np.pi, np.pi)
        out[s] = room_scan((px, py, th), w, h, beams, noise_mm, dropout, rng)
    return out


def polygon_scan(beams=8192, nsides=24, radius_mm=5000.0, noise_mm=5.0, pose=(0.0, 0.0, 0.0),
                 rng=None):
    """SURVEY.md 8(d) C3-style scan: a convex polygonal room with ``nsides`` wall segments
    (>= 20 so every 410-point chunk of an 8192-beam scan sees about one wall)."""
    if rng is None:
        rng = np.random.default_rng(0)
    angle = np.arange(beams, dtype=np.float64) * (360.0 / beams)
    px, py, th = pose
    ang = -angle * ANGLE_TO_RAD + np.pi / 2 + th
    ux, uy = np.cos(ang), np.sin(ang)
    best = np.full(beams, np.inf)
    for k in range(nsides):
        # wall k: points p with n.p = apothem, n = outward normal at angle (k+.5)*2pi/nsides
        phi = (k + 0.5) * 2 * np.pi / nsides
        nx, ny = np.cos(phi), np.sin(phi)
        apo = radius_mm * np.cos(np.pi / nsides)
        den = nx * ux + ny * uy
        with np.errstate(divide="ignore", invalid="ignore"):
            t = np.where(den > 1e-12, (apo - (nx * px + ny * py)) / den, np.inf)
        best = np.minimum(best, t)
    d = best + noise_mm * rng.standard_normal(beams)
    return polar_to_cart(d, angle)


def chunk_revolution(xy, chunk=MIN_NEIGHBOORS):
    """What ``scanning`` puts on the queue for one revolution (functions.py:64-76): lists of
    [x, y] pairs, ``chunk`` points each, the remainder only if it has more than 2 points,
    then the sentinel 0."""
    xy = np.asarray(xy)
    out = []
    s = 0
    while s + chunk <= len(xy):
        out.append(xy[s:s + chunk].tolist())
        s += chunk
    if len(xy) - s > 2:
        out.append(xy[s:].tolist())
    out.append(0)
    return out


# ---- raw sensor bytes: RPLidar express packets (lidar.py:66-91 is the decoder) --------------------
def pack_express(dist, code, start_q6, new_scan):
    """Vectorised packet assembly: dist [..., 32] ints < 2**14, code [..., 32] 6-bit angle-correction
    codes as the reference's decoder assembles them (low nibble in the cabin's fifth byte, bits 4-5 in
    the two low bits of the distance byte), start_q6 / new_scan [...] -> uint8 [..., 84] with sync
    nibbles and XOR checksum."""
    dist = np.asarray(dist, dtype=np.int64)
    code = np.asarray(code, dtype=np.int64)
    lead = dist.shape[:-1]
    p = np.zeros(lead + (84,), dtype=np.uint8)
    sq = np.asarray(start_q6, dtype=np.int64)
    p[..., 2] = sq & 0xFF
    p[..., 3] = ((sq >> 8) & 0x7F) | ((np.asarray(new_scan, dtype=np.int64) & 1) << 7)
    d1, d2, c1, c2 = dist[..., 0::2], dist[..., 1::2], code[..., 0::2], code[..., 1::2]
    cab = np.zeros(lead + (16, 5), dtype=np.uint8)
    cab[..., 0] = ((d1 & 0x3F) << 2) | ((c1 >> 4) & 3)
    cab[..., 1] = (d1 >> 6) & 0xFF
    cab[..., 2] = ((d2 & 0x3F) << 2) | ((c2 >> 4) & 3)
    cab[..., 3] = (d2 >> 6) & 0xFF
    cab[..., 4] = (c1 & 0xF) | ((c2 & 0xF) << 4)
    p[..., 4:] = cab.reshape(lead + (80,))
    cs = np.bitwise_xor.reduce(p[..., 2:], axis=-1)
    p[..., 0] = 0xA0 | (cs & 0x0F)
    p[..., 1] = 0x50 | (cs >> 4)
    return p


def random_room_express(nscans, packets_per_scan=12, seed=0, noise_mm=5.0, dropout=0.01):
    """The random-room scene as the sensor would send it: [nscans, P, 84] uint8.  Uniform rotation
    (P*32 measures per revolution), random start phase, correction codes within +-0.5 degree, integer
    millimetre ranges.  Host twin of lslam_synth_express (same scene model, NumPy streams)."""
    P = packets_per_scan
    out = np.empty((nscans, P, 84), dtype=np.uint8)
    i = np.arange(1, 33, dtype=np.float64)
    for s in range(nscans):
        rng = np.random.default_rng([seed, s, 7])
        w, h = rng.uniform(2000.0, 6000.0, size=2)
        px = rng.uniform(-w / 2 + 300.0, w / 2 - 300.0)
        py = rng.uniform(-h / 2 + 300.0, h / 2 - 300.0)
        th = rng.uniform(-np.pi, np.pi)
        phase = int(rng.integers(0, 360 * 64))
        sq = (phase + (np.arange(P + 1) % P) * 360 * 64 // P) % (360 * 64)
        start, nxt = sq[:P] / 64.0, sq[1:] / 64.0
        r3 = rng.integers(0, 8, size=(P, 32))
        code = np.where(r3 & 4, 28 + (r3 & 3), r3 & 3)
        dth = np.where(code & 16, code - 32, code) / 8.0
        ang = (start[:, None] + (((nxt - start) % 360) / 32)[:, None] * i[None, :] - dth) % 360
        d = room_ranges((px, py, th), w, h, ang) + noise_mm * rng.standard_normal((P, 32))
        if dropout > 0:
            d = np.where(rng.random((P, 32)) < dropout, 0.0, d)
        di = np.clip(np.rint(d), 0, 16383).astype(np.int64)
        out[s] = pack_express(di, code, sq[:P], (np.arange(P) == 0).astype(np.int64))
    return out


def random_aug_filters(F, L, seed=0, spread=3000.0):
    """F augmented-state filters (BASELINE.json configs[3]: pose + L landmarks in the state, n = 3 + 2L):
    the true pose near the origin, landmarks 1-4 m away, the filter's state = truth + prior-sized noise,
    P = diag(.1, .1, .05, 25, ...) (systemClass.py:27, 5 mm landmark sigma) + a little correlation,
    u ~ U(0.5, 2)^2, z = the true (range, bearing) of every landmark + sensor noise.
    Returns x [F, n], P [F, n, n], u [F, 2], z [F, L, 2] (float64)."""
    rng = np.random.default_rng(seed)
    n = 3 + 2 * L
    pose = np.column_stack([rng.normal(0, 50, F), rng.normal(0, 50, F), rng.uniform(-np.pi, np.pi, F)])
    ang = rng.uniform(-np.pi, np.pi, (F, L))
    rad = rng.uniform(1000.0, 1000.0 + spread, (F, L))
    lm = np.stack([pose[:, :1] + rad * np.cos(ang), pose[:, 1:2] + rad * np.sin(ang)], axis=2)
    x = np.zeros((F, n))
    x[:, :3] = pose + rng.normal(0, 1, (F, 3)) * np.array([0.3, 0.3, 0.02])
    x[:, 3:] = (lm + rng.normal(0, 5.0, lm.shape)).reshape(F, -1)
    d = np.array([0.1, 0.1, 0.05] + [25.0] * (2 * L))
    A = rng.normal(0, 0.05, (F, n, n))
    P = np.broadcast_to(np.diag(d), (F, n, n)) + np.einsum("fij,fkj->fik", A, A)
    u = rng.uniform(0.5, 2.0, (F, 2))
    dx, dy = lm[..., 0] - pose[:, :1], lm[..., 1] - pose[:, 1:2]
    rg = np.hypot(dx, dy) + rng.normal(0, 0.5, (F, L))
    br = np.arctan2(dy, dx) - pose[:, 2:3] + rng.normal(0, 0.05, (F, L))
    br = (br + np.pi) % (2 * np.pi) - np.pi
    return x, P, u, np.stack([rg, br], axis=2)
