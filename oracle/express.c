/* express.c - C twin of oracle/express.py (TEST INFRASTRUCTURE, see oracle/__init__.py): RPLidar
 * express packets -> (angle, distance) -> cartesian, as the reference computes them:
 *   ExpressPacket.decode           /root/reference/lidar.py:66-91
 *   Lidar._process_express_scan    /root/reference/lidar.py:179-187
 *   scanning() conversion          /root/reference/functions.py:59-60
 * Pinned through oracle/express.py, which tests/test_oracle_express.py checks bit for bit against
 * outputs of the real reference decoder (tests/golden/express_golden.npz). */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static double py_mod(double a, double m)         /* Python float %, m > 0 */
{
    double r = fmod(a, m);
    if (r != 0.0) { if (r < 0.0) r += m; } else r = 0.0;
    return r;
}

static double start_angle(const uint8_t *p) { return (double)(p[2] + ((p[3] & 0x7F) << 8)) / 64.0; }

/* returns 0 ok, 1 bad sync, 2 bad checksum (the reference raises ValueError, lidar.py:70-77) */
static int packet_check(const uint8_t *p)
{
    if ((p[0] >> 4) != 0xA || (p[1] >> 4) != 0x5) return 1;
    unsigned cs = 0;
    for (int i = 2; i < 84; ++i) cs ^= p[i];
    return cs == (unsigned)((p[0] & 0x0F) + ((p[1] & 0x0F) << 4)) ? 0 : 2;
}

/* packets [nscans][P][84]; xy [nscans][beams][2]; meas nullable [nscans][beams][2]; status nullable */
void oracle_express_to_cart(const uint8_t *packets, long long nscans, int P, int beams, double *xy, double *meas,
                            int32_t *status, int32_t nthreads)
{
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
    const double kPi = 3.141592653589793, kAngleToRad = kPi / 180.0;
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long long s = 0; s < nscans; ++s) {
        const uint8_t *sp = packets + (size_t)s * P * 84;
        int bad = 0;
        for (int p = 0; p < P; ++p) {
            const uint8_t *pk = sp + (size_t)p * 84;
            const int st = packet_check(pk);
            bad |= st;
            const double start = start_angle(pk), nstart = start_angle(sp + (size_t)((p + 1) % P) * 84);
            const double step = py_mod(nstart - start, 360.0) / 32.0;
            for (int m = 0; m < 32; ++m) {
                const int k = p * 32 + m;
                if (k >= beams) break;
                const int i = 5 * (m >> 1);
                const int lo = (m & 1) ? i + 6 : i + 4;
                const int dist = (pk[lo] >> 2) + (pk[lo + 1] << 6);
                int code = (((m & 1) ? (pk[i + 8] >> 4) : pk[i + 8]) & 0x0F) + ((pk[lo] & 0x03) << 4);
                if (code & 16) code -= 32;
                const double ang = py_mod(start + step * (double)(m + 1) - (double)code / 8.0, 360.0);
                const size_t o = ((size_t)s * beams + k) * 2;
                if (st == 0) {
                    const double t = -ang * kAngleToRad + kPi / 2.0;
                    xy[o] = (double)dist * cos(t);
                    xy[o + 1] = (double)dist * sin(t);
                    if (meas) { meas[o] = ang; meas[o + 1] = (double)dist; }
                } else {
                    xy[o] = xy[o + 1] = NAN;
                    if (meas) meas[o] = meas[o + 1] = NAN;
                }
            }
        }
        if (status) status[s] = bad;
    }
}
