// express.cuh - input stage on the device: RPLidar express-scan packets -> points.
//
// Replaces, for a batch of revolutions resident in HBM as raw sensor bytes:
//   ExpressPacket.decode            /root/reference/lidar.py:66-91   (sync, checksum, 32 x (distance, dtheta))
//   Lidar._process_express_scan     /root/reference/lidar.py:179-187 (angle interpolation between packets)
//   scanning() conversion           /root/reference/functions.py:59-60
// Layout: packets[nscans][P][84] bytes; one revolution = P consecutive packets, the start angle that
// follows the last packet is the first packet's (steady rotation).  One warp decodes one packet: 21
// lanes load one 32-bit word each (coalesced, 84 B per packet), the XOR checksum is one warp reduction,
// lane l is measure l (trame l+1).  All angle arithmetic is exact in binary floating point (multiples
// of 1/2048 degree below 2^11), so the angles are bit-identical to the reference's Python doubles;
// only cos/sin differ from glibc's by their last bits.
// HBM traffic per point: 84/32 = 2.6 B read, 16 B written (+16 B when `meas` is requested).
#pragma once
#include "ransac.cuh"      // LSLAM_SMEM_DECL
#include "aux_kernels.cuh" // u01

namespace lslam {

constexpr int kExpressBytes = 84;
constexpr int kExpressWords = 21;
constexpr int kExpressMaxWarps = 12;
constexpr int kExpressRow = kExpressWords + 3;                 // words per warp row in shared memory
constexpr int kExpressSmem = kExpressRow * kExpressMaxWarps * 4 + 16;

// Python's float `%` with a positive divisor: fmod, then shifted into [0, m).
__device__ __forceinline__ double py_mod(double a, double m)
{
    double r = fmod(a, m);
    if (r != 0.0) {
        if (r < 0.0) r = __dadd_rn(r, m);
    } else {
        r = 0.0;                                    // copysign(0, m)
    }
    return r;
}

__device__ __forceinline__ unsigned express_byte(const uint32_t *w, int off) { return (w[off >> 2] >> ((off & 3) * 8)) & 0xffu; }

// status bits per revolution
constexpr int kExpressBadSync = 1, kExpressBadChecksum = 2;

__global__ void express_to_cart_kernel(const uint32_t *__restrict__ packets, long long nscans, int P, int beams,
                                       double2 *__restrict__ xy, double2 *__restrict__ meas, int32_t *__restrict__ status)
{
    LSLAM_SMEM_DECL;
    uint32_t (*words)[kExpressRow] = reinterpret_cast<uint32_t (*)[kExpressRow]>(smem_raw);
    int &bad = *reinterpret_cast<int *>(smem_raw + sizeof(uint32_t) * kExpressRow * kExpressMaxWarps);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double kPi = 3.141592653589793, kAngleToRad = kPi / 180.0;
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    for (long long scan = blockIdx.x; scan < nscans; scan += gridDim.x) {
        if (threadIdx.x == 0) bad = 0;
        __syncthreads();
        const uint32_t *sp = packets + (size_t)scan * P * kExpressWords;
        for (int p = warp; p < P; p += nwarps) {
            const uint32_t *pk = sp + (size_t)p * kExpressWords;
            const uint32_t w = lane < kExpressWords ? pk[lane] : 0u;
            const uint32_t nw0 = sp[(size_t)((p + 1 == P) ? 0 : p + 1) * kExpressWords];   // next packet's first word
            if (lane < kExpressWords) words[warp][lane] = w;
            // checksum over bytes 2..83 = all words with the two sync bytes masked off
            uint32_t x = __reduce_xor_sync(0xffffffffu, lane == 0 ? (w & 0xffff0000u) : w);
            x ^= x >> 16;
            x = (x ^ (x >> 8)) & 0xffu;
            const uint32_t w0 = __shfl_sync(0xffffffffu, w, 0);
            int st = 0;
            if (((w0 >> 4) & 0xfu) != 0xAu || ((w0 >> 12) & 0xfu) != 0x5u) st |= kExpressBadSync;
            else if (x != ((w0 & 0xfu) | (((w0 >> 8) & 0xfu) << 4))) st |= kExpressBadChecksum;
            __syncwarp();
            const int k = p * 32 + lane;
            if (k < beams) {
                const uint32_t *ww = words[warp];
                const int i = 5 * (lane >> 1);
                const int lo = (lane & 1) ? i + 6 : i + 4;
                const unsigned b_lo = express_byte(ww, lo), b_hi = express_byte(ww, lo + 1), b_a = express_byte(ww, i + 8);
                const int dist = (int)(b_lo >> 2) + (int)(b_hi << 6);                       // lidar.py:85,88
                int code = (int)(((lane & 1) ? (b_a >> 4) : b_a) & 0xfu) + (int)((b_lo & 3u) << 4);
                if (code & 16) code -= 32;                                                   // twos_comp(., 5)
                const double dth = (double)code / 8.0;
                const double start = (double)(((w0 >> 16) & 0xffu) + (((w0 >> 24) & 0x7fu) << 8)) / 64.0;   // :80
                const double nstart = (double)(((nw0 >> 16) & 0xffu) + (((nw0 >> 24) & 0x7fu) << 8)) / 64.0;
                // lidar.py:185  (start + ((next - start) % 360)/32 * trame - dtheta) % 360
                const double step = py_mod(__dsub_rn(nstart, start), 360.0) / 32.0;
                const double ang = py_mod(__dsub_rn(__dadd_rn(start, __dmul_rn(step, (double)(lane + 1))), dth), 360.0);
                const double d = (double)dist;
                const size_t o = (size_t)scan * beams + k;
                if (st == 0) {
                    const double t = __dadd_rn(__dmul_rn(-ang, kAngleToRad), kPi / 2.0);     // functions.py:59-60
                    double s, c;
                    sincos(t, &s, &c);
                    xy[o] = make_double2(__dmul_rn(d, c), __dmul_rn(d, s));
                    if (meas) meas[o] = make_double2(ang, d);
                } else {
                    xy[o] = make_double2(nan, nan);
                    if (meas) meas[o] = make_double2(nan, nan);
                }
            }
            if (st != 0 && lane == 0) atomicOr(&bad, st);
            __syncwarp();
        }
        __syncthreads();
        if (threadIdx.x == 0 && status) status[scan] = bad;
        __syncthreads();
    }
}

// Headless stand-in for the sensor: revolution `scan_base + b` of the random-room scene of
// synth_scans_kernel, emitted as express packets.  P packets of 32 measures per revolution at a
// uniform rotation rate, a random start phase per revolution, angle-correction codes within +-0.5
// degree, ranges rounded to the sensor's integer millimetres.  One warp per packet.
__global__ void synth_express_kernel(uint32_t *__restrict__ packets, long long nscans, int P, uint64_t seed,
                                     uint64_t scan_base, double noise_mm, double dropout)
{
    LSLAM_SMEM_DECL;
    uint32_t (*words)[kExpressRow] = reinterpret_cast<uint32_t (*)[kExpressRow]>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double kPi = 3.141592653589793;
    for (long long sb = blockIdx.x; sb < nscans; sb += gridDim.x) {
        const uint64_t scan = scan_base + (uint64_t)sb;
        const uint32_t k0 = (uint32_t)seed ^ 0x5ca1ab1eu, k1 = (uint32_t)(seed >> 32);
        const Philox4 room = philox4x32_10(0u, 0xffffffffu, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
        const Philox4 head = philox4x32_10(1u, 0xffffffffu, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
        const double w = 2000.0 + 4000.0 * u01(room.r[0]), hgt = 2000.0 + 4000.0 * u01(room.r[1]);
        const double px = (-0.5 * w + 300.0) + (w - 600.0) * u01(room.r[2]);
        const double py = (-0.5 * hgt + 300.0) + (hgt - 600.0) * u01(room.r[3]);
        const double th = -kPi + 2.0 * kPi * u01(head.r[0]);
        const int phase_q6 = (int)(head.r[1] % (360u * 64u));
        for (int p = warp; p < P; p += nwarps) {
            // start angles in q6, uniform rotation: packet p starts at phase + p*360/P
            const int s_q6 = (phase_q6 + (int)(((long long)p * 360 * 64) / P)) % (360 * 64);
            const int n_q6 = (phase_q6 + (int)(((long long)((p + 1) % P) * 360 * 64) / P)) % (360 * 64);
            const int k = p * 32 + lane;
            const Philox4 nz = philox4x32_10((uint32_t)k, 0u, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
            const int r3 = (int)(nz.r[3] & 7u);
            const int code = (r3 & 4) ? 28 + (r3 & 3) : (r3 & 3);
            const double dth = (double)((code & 16) ? code - 32 : code) / 8.0;
            const double start = (double)s_q6 / 64.0, nstart = (double)n_q6 / 64.0;
            const double step = py_mod(nstart - start, 360.0) / 32.0;
            const double ang = py_mod(start + step * (double)(lane + 1) - dth, 360.0);
            const double sens = -ang * (kPi / 180.0) + kPi / 2.0;
            double us, uc;
            sincos(sens + th, &us, &uc);
            const double inf = __longlong_as_double(0x7ff0000000000000LL);
            const double tx = uc > 0.0 ? (0.5 * w - px) / uc : (uc < 0.0 ? (-0.5 * w - px) / uc : inf);
            const double ty = us > 0.0 ? (0.5 * hgt - py) / us : (us < 0.0 ? (-0.5 * hgt - py) / us : inf);
            double d = fmin(tx, ty);
            d += noise_mm * sqrt(-2.0 * log(u01(nz.r[0]))) * cos(2.0 * kPi * u01(nz.r[1]));
            if (u01(nz.r[2]) < dropout) d = 0.0;
            int di = (int)rint(d);
            di = di < 0 ? 0 : (di > 16383 ? 16383 : di);
            // cabin assembly: the even lane of a pair writes the cabin's 5 bytes
            const int d2 = __shfl_down_sync(0xffffffffu, di, 1), c2 = __shfl_down_sync(0xffffffffu, code, 1);
            if (lane < kExpressRow) words[warp][lane] = 0u;
            __syncwarp();
            unsigned char *bytes = reinterpret_cast<unsigned char *>(words[warp]);
            if ((lane & 1) == 0) {
                const int i = 5 * (lane >> 1);
                bytes[i + 4] = (unsigned char)(((di & 0x3f) << 2) | ((code >> 4) & 3));
                bytes[i + 5] = (unsigned char)((di >> 6) & 0xff);
                bytes[i + 6] = (unsigned char)(((d2 & 0x3f) << 2) | ((c2 >> 4) & 3));
                bytes[i + 7] = (unsigned char)((d2 >> 6) & 0xff);
                bytes[i + 8] = (unsigned char)((code & 0xf) | ((c2 & 0xf) << 4));
            }
            if (lane == 31) {
                bytes[2] = (unsigned char)(s_q6 & 0xff);
                bytes[3] = (unsigned char)(((s_q6 >> 8) & 0x7f) | ((p == 0) ? 0x80 : 0));   // new-scan flag on packet 0
            }
            __syncwarp();
            uint32_t wv = lane < kExpressWords ? words[warp][lane] : 0u;
            uint32_t x = __reduce_xor_sync(0xffffffffu, wv);
            x ^= x >> 16;
            x = (x ^ (x >> 8)) & 0xffu;
            if (lane == 0) wv |= (0xA0u | (x & 0xfu)) | ((0x50u | (x >> 4)) << 8);
            if (lane < kExpressWords) packets[((size_t)sb * P + p) * kExpressWords + lane] = wv;
            __syncwarp();
        }
    }
}

} // namespace lslam
