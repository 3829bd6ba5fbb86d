// aux_kernels.cuh - input stage (polar -> cartesian), synthetic scan generator, fp64 probe.
#pragma once
#include "lslam_core.cuh"

namespace lslam {

// K0: /root/reference/functions.py:59-60, one thread per measure, 16-byte stores.
//   ang = -angle_deg * (pi/180) + pi/2 ;  x = d*cos(ang) ; y = d*sin(ang)
__global__ void polar_to_cart_kernel(const double *dist, const double *ang_deg, double2 *xy, long long npts)
{
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= npts) return;
    const double kPi = 3.141592653589793;                  // np.pi
    const double kAngleToRad = kPi / 180.0;                // ANGLE_TO_RAD, functions.py:13
    const double d = dist[k];
    const double ang = __dadd_rn(__dmul_rn(-ang_deg[k], kAngleToRad), kPi / 2.0);
    double s, c;
    sincos(ang, &s, &c);
    xy[k] = make_double2(__dmul_rn(d, c), __dmul_rn(d, s));
}

// R6b: the projected inlier points landmark_extraction returns (/root/reference/ransac_functions.py:28,46,49,53,55):
//   xBase = data[inliers, 0];  yBase = a' * xBase + b'   (a', b' = the matched STORED landmark's on a hit, the new line's
//   on a miss: `proj_ab` of lslam_associate, or the line record itself) - a vertical projection, not an orthogonal one.
// One warp per chunk slot (grid-stride): lane l owns point 32 w + l of mask word w; an inlier's position among the
// slot's inliers is the popcount of the lower bits, so the points come out compacted in data order.  Output rows
// [first, first + ninl) with first = offsets[slot] (compact layout) or slot * chunk (padded layout).
__global__ void project_inliers_kernel(const double2 *xy, const int32_t *npts, long long nslots, int N, int chunk, int C, int W,
                                       const uint32_t *mask, const int32_t *ninl, const double *ab, int ab_stride,
                                       const long long *offsets, double2 *points)
{
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long s = warp0; s < nslots; s += nwarps) {
        if (ninl[s] < 2) continue;                                // unused slot or status code: nothing to project
        const long long b = s / C;
        const int c = (int)(s - b * C);
        int nb = npts ? npts[b] : N;
        nb = nb < 0 ? 0 : (nb > N ? N : nb);
        int n = nb - c * chunk;
        n = n < 0 ? 0 : (n > chunk ? chunk : n);
        const double2 *P = xy + b * N + (long long)c * chunk;
        const double pa = ab[s * ab_stride], pb = ab[s * ab_stride + 1];
        double2 *out = points + (offsets ? offsets[s] : s * chunk);
        int base = 0;
        for (int w = 0; w * 32 < n; ++w) {
            const unsigned word = mask[s * W + w];
            const int k = w * 32 + lane;
            if (k < n && ((word >> lane) & 1u)) {
                const double x = P[k].x;
                out[base + __popc(word & ((1u << lane) - 1u))] = make_double2(x, __dadd_rn(__dmul_rn(pa, x), pb));
            }
            base += __popc(word);
        }
    }
}

__device__ __forceinline__ double u01(uint32_t r) { return ((double)r + 0.5) * 0x1p-32; }

// Headless stand-in for the capture process: scan `scan_base + b` is one revolution in a
// random rectangular room (same scene model as lidar_slam_b200/synth.py random_room_scans):
// room w,h ~ U(2, 6 m), pose ~ U(interior, 0.3 m off the walls), heading ~ U(-pi, pi),
// N(0, noise_mm) range noise, `dropout` probability of a zero range.  One thread per beam.
__global__ void synth_scans_kernel(double2 *xy, long long total, int beams, uint64_t seed, uint64_t scan_base,
                                   double noise_mm, double dropout)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const uint64_t scan = scan_base + (uint64_t)(idx / beams);
    const int beam = (int)(idx % beams);
    const uint32_t k0 = (uint32_t)seed ^ 0x5ca1ab1eu, k1 = (uint32_t)(seed >> 32);
    const Philox4 room = philox4x32_10(0u, 0xffffffffu, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
    const Philox4 head = philox4x32_10(1u, 0xffffffffu, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
    const Philox4 nz = philox4x32_10((uint32_t)beam, 0u, (uint32_t)scan, (uint32_t)(scan >> 32), k0, k1);
    const double kPi = 3.141592653589793;
    const double w = 2000.0 + 4000.0 * u01(room.r[0]), hgt = 2000.0 + 4000.0 * u01(room.r[1]);
    const double px = (-0.5 * w + 300.0) + (w - 600.0) * u01(room.r[2]);
    const double py = (-0.5 * hgt + 300.0) + (hgt - 600.0) * u01(room.r[3]);
    const double th = -kPi + 2.0 * kPi * u01(head.r[0]);
    const double angle_deg = (double)beam * (360.0 / (double)beams);
    const double sens = -angle_deg * (kPi / 180.0) + kPi / 2.0;   // sensor-frame beam angle
    double us, uc;
    sincos(sens + th, &us, &uc);                                  // world-frame beam direction
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const double tx = uc > 0.0 ? (0.5 * w - px) / uc : (uc < 0.0 ? (-0.5 * w - px) / uc : inf);
    const double ty = us > 0.0 ? (0.5 * hgt - py) / us : (us < 0.0 ? (-0.5 * hgt - py) / us : inf);
    double d = fmin(tx, ty);
    const double gauss = sqrt(-2.0 * log(u01(nz.r[0]))) * cos(2.0 * kPi * u01(nz.r[1]));
    d += noise_mm * gauss;
    if (u01(nz.r[2]) < dropout) d = 0.0;
    double ss, sc;
    sincos(sens, &ss, &sc);
    xy[idx] = make_double2(d * sc, d * ss);
}

// fp64 issue-rate probe: kDfmaChains independent dependent-FMA chains per thread.
constexpr int kDfmaChains = 8;
constexpr int kDfmaUnroll = 16;
__global__ void dfma_probe_kernel(double *sink, int iters)
{
    double acc[kDfmaChains];
    const double a = 1.0000001, b = 1e-9 * (double)(threadIdx.x + 1);
#pragma unroll
    for (int c = 0; c < kDfmaChains; ++c) acc[c] = (double)c;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < kDfmaUnroll; ++u)
#pragma unroll
            for (int c = 0; c < kDfmaChains; ++c) acc[c] = fma(acc[c], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < kDfmaChains; ++c) s += acc[c];
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

} // namespace lslam
