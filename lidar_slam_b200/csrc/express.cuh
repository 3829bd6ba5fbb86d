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
// cos/sin come from two small tables and the angle-sum formulas (about 1 ulp, like glibc's).
// HBM traffic per point: 84/32 = 2.6 B read, 16 B written (+16 B when `meas` is requested).
#pragma once
#include "ransac.cuh"      // LSLAM_SMEM_DECL
#include "aux_kernels.cuh" // u01

namespace lslam {

constexpr int kExpressBytes = 84;
constexpr int kExpressWords = 21;
constexpr int kExpressMaxWarps = 12;
constexpr int kExpressRow = kExpressWords + 3;                 // words per warp row in shared memory
constexpr int kExpressSmem = 2 * kExpressRow * kExpressMaxWarps * 4 + 16;   // two packet rows per warp
// sin/cos tables of the decoder: 360 whole degrees + 2048 steps of 1/2048 degree, (sin, cos) pairs
constexpr int kExpressTabPairs = 360 + 2048;
constexpr int kExpressTabDoubles = 2 * kExpressTabPairs;
constexpr int kExpressDecodeSmem = kExpressSmem + 16 + kExpressTabDoubles * 8;

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

// position of row r of the sub-degree table in shared memory (see the staging loop of express_to_cart_kernel)
__device__ __forceinline__ int express_swz(int r) { return r ^ ((r >> 7) & 7); }

// status bits per revolution
constexpr int kExpressBadSync = 1, kExpressBadChecksum = 2;

// Persistent warps, TWO consecutive packets per warp iteration, no block-level synchronisation: packets 2 j and 2 j + 1 of
// the batch go to warp j mod (warps in the grid); lanes 0-15 decode the first, lanes 16-31 the second, a lane per CABIN
// (5 bytes = two measures), so the byte extraction and the angle base are shared by two points and every lane stores 32
// contiguous bytes.  Addressing is incremental (a packet pointer, an output index and the position in the revolution advance
// by constants), the floor-mods are conditional additions (their operands are 15-bit fields).  The next pair's words are
// loaded before the current one is decoded (software pipelining in registers; a four-deep cp.async ring of 4-byte copies
// measured slower, 2.16 against 2.01 ms per 1M revolutions).  `status` must be zero on entry (the entry point clears it):
// a rejected packet ORs its bit into its revolution's word.
__global__ void __launch_bounds__(32 * kExpressMaxWarps, 3)
express_to_cart_kernel(const uint32_t *__restrict__ packets, long long nscans, int P, int beams, double2 *__restrict__ xy,
                       double2 *__restrict__ meas, int32_t *__restrict__ status, const double2 *__restrict__ trig)
{
    LSLAM_SMEM_DECL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int half = lane >> 4, cab = lane & 15;
    uint32_t *ww = reinterpret_cast<uint32_t *>(smem_raw) + warp * 2 * kExpressRow;   // this warp's two packets
    // (sin, cos) of d degrees [0, 360) and of b/2048 degree [0, 2048): staged once per CTA
    double2 *tab = reinterpret_cast<double2 *>(smem_raw + ((kExpressSmem + 15) & ~15));
    // The sub-degree table is read at (k2048 & 2047), and at a steady rotation that index moves by a multiple of 128 from one
    // measure to the next (30 degrees per packet = 1920 / 2048 degree per measure): rows 2 KB apart, all in the same four
    // banks.  XOR-ing bits 7-9 of the row into its low three bits spreads them over the eight 16-byte bank groups (express_swz).
    for (int k = threadIdx.x; k < kExpressTabPairs; k += blockDim.x) tab[k < 360 ? k : 360 + express_swz(k - 360)] = trig[k];
    __syncthreads();
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const long long stride = 2 * (long long)gridDim.x * wpb, first = 2 * ((long long)blockIdx.x * wpb + warp);
    const long long npk = nscans * P;
    if (first >= npk) return;
    const long long stride_scan = stride / P;
    const int stride_p = (int)(stride - stride_scan * P);
    long long scan = first / P;                                                // of the pair's FIRST packet (warp-uniform)
    int p = (int)(first - scan * P);
    const uint32_t *pkt = packets + (size_t)first * kExpressWords;
    const long long pkt_step = stride * kExpressWords;
    long long obase = scan * beams + (long long)p * 32;                        // output index of the first packet's first measure
    const long long o_step = stride_scan * beams + (long long)stride_p * 32, o_wrap = (long long)beams - (long long)P * 32;
    const unsigned char *wb = reinterpret_cast<const unsigned char *>(ww + half * kExpressRow) + 4 + 5 * cab;   // my cabin
    // words of both packets (42 consecutive words: lanes 0-20 the first, again the second) and, per half, the first word of
    // the packet that FOLLOWS mine in its revolution (its start angle; after the last packet: the first packet's)
    auto load = [&](const uint32_t *pk_, int pp, bool two, uint32_t &wa_, uint32_t &wb_, uint32_t &n_) {
        wa_ = lane < kExpressWords ? pk_[lane] : 0u;
        wb_ = (two && lane < kExpressWords) ? pk_[kExpressWords + lane] : 0u;
        int pm = pp + half;                                                    // my packet's position in its revolution
        if (pm == P) pm = 0;
        const uint32_t *mine = pk_ + half * kExpressWords;
        n_ = (half && !two) ? 0u : ((pm + 1 == P) ? *(mine - (ptrdiff_t)pm * kExpressWords) : mine[kExpressWords]);
    };
    uint32_t wA, wB, nw0;
    load(pkt, p, first + 1 < npk, wA, wB, nw0);
    for (long long pk = first; pk < npk; pk += stride) {
        const bool two = pk + 1 < npk;
        // prefetch the following pair of this warp
        long long scan_n = scan + stride_scan, obase_n = obase + o_step;
        int p_n = p + stride_p;
        if (p_n >= P) { p_n -= P; ++scan_n; obase_n += o_wrap; }
        const uint32_t *pkt_n = pkt + pkt_step;
        uint32_t wA_n = 0u, wB_n = 0u, nw0_n = 0u;
        if (pk + stride < npk) load(pkt_n, p_n, pk + stride + 1 < npk, wA_n, wB_n, nw0_n);

        if (lane < kExpressWords) { ww[lane] = wA; ww[kExpressRow + lane] = wB; }
        // checksum over bytes 2..83 = all words with the two sync bytes masked off
        const uint32_t xa = __reduce_xor_sync(0xffffffffu, lane == 0 ? (wA & 0xffff0000u) : wA);
        const uint32_t xb = __reduce_xor_sync(0xffffffffu, lane == 0 ? (wB & 0xffff0000u) : wB);
        const uint32_t w0a = __shfl_sync(0xffffffffu, wA, 0), w0b = __shfl_sync(0xffffffffu, wB, 0);
        uint32_t x = half ? xb : xa;
        const uint32_t w0 = half ? w0b : w0a;
        x ^= x >> 16;
        x = (x ^ (x >> 8)) & 0xffu;
        int st = 0;
        if (((w0 >> 4) & 0xfu) != 0xAu || ((w0 >> 12) & 0xfu) != 0x5u) st |= kExpressBadSync;
        else if (x != ((w0 & 0xfu) | (((w0 >> 8) & 0xfu) << 4))) st |= kExpressBadChecksum;
        __syncwarp();
        // my packet: position in its revolution, revolution, output index of its first measure
        int pm = p + half;
        long long scan_m = scan, o = obase + 32 * half;
        if (pm == P) { pm = 0; ++scan_m; o += o_wrap; }
        const int k = pm * 32 + 2 * cab;                                       // index of the cabin's first measure in the revolution
        if ((two || !half) && k < beams) {
            const unsigned b0 = wb[0], b1 = wb[1], b2 = wb[2], b3 = wb[3], b4 = wb[4];
            const int dist0 = (int)(b0 >> 2) + (int)(b1 << 6), dist1 = (int)(b2 >> 2) + (int)(b3 << 6);   // lidar.py:85,88
            int code0 = (int)(b4 & 0xfu) + (int)((b0 & 3u) << 4), code1 = (int)(b4 >> 4) + (int)((b2 & 3u) << 4);
            if (code0 & 16) code0 -= 32;                                                 // twos_comp(., 5)
            if (code1 & 16) code1 -= 32;
            // lidar.py:80,185  angle = (start + ((next - start) % 360)/32 * trame - dtheta) % 360 with
            // start = q6/64, dtheta = code/8.  Every term is a multiple of 1/2048 degree below 2^11, so
            // the reference's float arithmetic is exact and the same value comes out of integers in
            // units of 1/2048 degree (Python's % on floats = floor-mod).
            const int sq = (int)((w0 >> 16) & 0x7fffu);                                   // start, q6
            const int nq = (int)((nw0 >> 16) & 0x7fffu);                                  // next start, q6
            int dq = nq - sq;                                                             // in (-32768, 32768); mod 360 * 64
            if (dq < 0) dq += 23040;
            if (dq < 0) dq += 23040;
            if (dq >= 23040) dq -= 23040;
            const int base = sq * 32 + dq * (2 * cab + 1);                                // trame = measure index + 1
            o += 2 * cab;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (j == 1 && k + 1 >= beams) break;
                int k2048 = base + (j ? dq : 0) - (j ? code1 : code0) * 256;              // in (-4096, 3 * 360 * 2048); mod 360 * 2048
                if (k2048 >= 737280) k2048 -= 737280;
                if (k2048 >= 737280) k2048 -= 737280;
                if (k2048 < 0) k2048 += 737280;
                const double d = (double)(j ? dist1 : dist0);
                if (st == 0) {
                    // functions.py:59-60: x = d cos(pi/2 - theta), y = d sin(pi/2 - theta), theta = ang pi/180,
                    // i.e. x = d sin(theta), y = d cos(theta).  ang is an exact multiple of 1/2048 degree:
                    // theta = (a + b/2048) degrees, and the angle-sum formulas on two table entries give
                    // sin/cos to ~1 ulp without the ~100 fp64 instructions of a general sincos.
                    const double2 A = tab[k2048 >> 11], Bq = tab[360 + express_swz(k2048 & 2047)];   // (sin, cos)
                    const double sn = fma(A.x, Bq.y, A.y * Bq.x);                         // sin(a + b)
                    const double cs = fma(A.y, Bq.y, -(A.x * Bq.x));                      // cos(a + b)
                    xy[o + j] = make_double2(__dmul_rn(d, sn), __dmul_rn(d, cs));
                    if (meas) meas[o + j] = make_double2((double)k2048 * (1.0 / 2048.0), d);
                } else {
                    xy[o + j] = make_double2(nan, nan);
                    if (meas) meas[o + j] = make_double2(nan, nan);
                }
            }
        }
        if (st != 0 && cab == 0 && (two || !half) && status) atomicOr(status + scan_m, st);
        __syncwarp();
        wA = wA_n; wB = wB_n; nw0 = nw0_n; scan = scan_n; p = p_n; pkt = pkt_n; obase = obase_n;
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
