// ukf.cuh - K4: UKF sigma-point predict / update for many independent n = 3 filters.
//
// Replaces filterpy's UnscentedKalmanFilter.predict/update + MerweScaledSigmaPoints driven by the
// callbacks of /root/reference/UKFMethods.py:10-71 with the configuration of
// /root/reference/systemClass.py:14-29.  Semantics and operation order = oracle/ukf.c (every
// product and sum below is an explicit round-to-nearest intrinsic or an explicit fma, in the
// order that file pins), so the only differences to the CPU oracle are the last-ulp differences
// of sin/cos/atan2 between the CUDA math library and libm.
//
// Work decomposition: a GROUP of G lanes (G = 4, 8, 16 or 32, the smallest that holds the L
// landmark slots) owns one filter, so a warp advances 32/G filters at once.
//   * everything of dimension n = 3 (Cholesky of P, the 7 sigma points, fx, the circular state
//     mean, the prior covariance) is computed redundantly by every lane of the group in
//     registers: replicated scalar work costs the same issue slots as doing it on one lane and
//     needs no communication;
//   * the measurement side is spread over the group: lane q owns the q-th valid landmark - its
//     7 hx evaluations, its circular mean, its residual columns and its 3x2 block of Pxz;
//   * S (2L x 2L) is built by row pairs (rows q and 2L-1-q have 2L+1 lower-triangle entries
//     together, so the lanes are balanced), factorised by a group-cooperative column Cholesky
//     in shared memory, and K, x, P follow from triangular solves and small products.
// Shared memory is interleaved across the groups of a warp (element e of group g at e*NG + g)
// so replicated reads are one broadcast wavefront and per-lane reads are conflict free.
#pragma once
#include "lslam_core.cuh"

namespace lslam {

enum { kUkfPredict = 1, kUkfUpdate = 2, kUkfRedraw = 4 };
enum { kUkfStSigma = 1, kUkfStRedraw = 2, kUkfStS = 4 };

constexpr int kUkfMaxPeers = 8;

struct UkfArgs {
    double *x, *P;                 // [F,3], [F,3,3] in/out
    const double *u;               // [F,2] or null (zero controls)
    const double *z;               // [F,L,zstride]
    const double *lm_pos;          // [F,L,2]
    const int32_t *lm_valid;       // [F,L] or null
    int32_t *status;               // [F] or null
    double *sigmas_f;              // [F,7,3] or null
    int F, L, zstride, zmode, valid_min, flags;
    double dt, half_r, r_over_l, cscale;     // R/2, R/L, n + lambda
    double wm0, wc0, wi;                     // weights
    double var_dist, var_angle;
    double Q[9];
    // fused all-gather of the step's results over peer memory (lslam_ukf_step_gather); world == 0: off
    int gather_world, gather_L;
    long long gather_first;        // record index of this call's filter 0 in the gather buffers
    const int32_t *gather_ids;     // [F, gather_L] local per-slot ids to publish, or null
    double *gather_state_mc;       // multicast address of the [world, F, 12] state buffers, or null
    int32_t *gather_ids_mc;        // multicast address of the [world, F, gather_L] id buffers, or null
    double *gather_state[kUkfMaxPeers];     // every rank's state buffer as mapped on this GPU (unicast path)
    int32_t *gather_idbuf[kUkfMaxPeers];
};

// One store that NVSwitch replicates into the same offset of every rank's buffer (multicast object).
__device__ __forceinline__ void multimem_store(double *mc_addr, double v)
{
#ifndef LSLAM_EMU
    asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(mc_addr), "d"(v) : "memory");
#else
    *mc_addr = v;
#endif
}
__device__ __forceinline__ void multimem_store(int32_t *mc_addr, int32_t v)
{
#ifndef LSLAM_EMU
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
#else
    *mc_addr = v;
#endif
}

constexpr int kUkfHandoff = 33;               // doubles per filter handed from the predict to the update: x 3, P 9, sigma points 21
__host__ __device__ inline size_t ukf_smem_doubles_per_warp(int G)
{
    return (size_t)(kUkfHandoff + 1) * 32 + (size_t)(15 * 2 * G + (G >= 16 ? 4 * G * G : 0)) * (32 / G);
}

__device__ __forceinline__ double ukf_normalize_angle(double a)
{
    const double kPi = 3.141592653589793, kTwoPi = 2 * 3.141592653589793;
    // fmod is exact, and for |a| < 2 pi it returns a itself: the common case skips the ~50-instruction routine
    double m = fabs(a) < kTwoPi ? a : fmod(a, kTwoPi);   // Python's float % then fixes the sign
    if (m != 0.0) { if (m < 0.0) m = __dadd_rn(m, kTwoPi); }
    else m = 0.0;
    if (m > kPi) m = __dsub_rn(m, kTwoPi);
    return m;
}

// upper Cholesky of c*P (3x3, upper triangle read) and the 7 sigma points; false if not PD
__device__ __forceinline__ bool ukf_sigma_points(const double x[3], const double P[9], double c, double sg[7][3])
{
    double U[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        double s = __dmul_rn(c, P[j * 3 + j]);
#pragma unroll
        for (int p = 0; p < j; ++p) s = __fma_rn(-U[p * 3 + j], U[p * 3 + j], s);
        if (!(s > 0.0)) return false;
        const double d = __dsqrt_rn(s);
        U[j * 3 + j] = d;
#pragma unroll
        for (int i = j + 1; i < 3; ++i) {
            double t = __dmul_rn(c, P[j * 3 + i]);
#pragma unroll
            for (int p = 0; p < j; ++p) t = __fma_rn(-U[p * 3 + j], U[p * 3 + i], t);
            U[j * 3 + i] = __ddiv_rn(t, d);
        }
#pragma unroll
        for (int i = 0; i < j; ++i) U[j * 3 + i] = 0.0;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) sg[0][i] = x[i];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            sg[k + 1][i] = __dadd_rn(x[i], U[k * 3 + i]);
            sg[3 + k + 1][i] = __dsub_rn(x[i], U[k * 3 + i]);
        }
    return true;
}

template <int G>
__global__ void __launch_bounds__(128) ukf_kernel(const UkfArgs a)
{
    LSLAM_SMEM_DECL;
    constexpr int NG = 32 / G;                 // groups (filters) per warp and pass
    constexpr int DZ = 2 * G;                  // leading dimension of the measurement arrays
    // shared memory per filter of a pass: E [7][DZ] residual columns; Aug [4][DZ] Pxz rows 0..2 and y (the right-hand sides, then
    // W and v); Row [2][DZ] the current row of U, double-buffered (G <= 8) or the diagonal of S and 1 / diag U (G >= 16);
    // Res [2*DZ] the nine final sums; S [DZ][DZ] only for G >= 16
    constexpr int oE = 0, oAug = 7 * DZ, oPxz = oAug, oY = oAug + 3 * DZ, oRow = 11 * DZ, oSd = oRow, oRd = oRow + DZ, oRes = 13 * DZ, oS = 15 * DZ;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int grp = lane / G, g = lane % G;
    // per warp: the hand-off area of its 32 filters (x, P, sigma points; element e of filter l at e * 32 + l, then status and
    // flags as ints), then the scratch of the NG filters of a pass
    double *H = reinterpret_cast<double *>(smem_raw) + (size_t)warp * ukf_smem_doubles_per_warp(G);
    int *Hi = reinterpret_cast<int *>(H + kUkfHandoff * 32);
    double *smw = H + (kUkfHandoff + 1) * 32;
#define UKF_SM(e) smw[(e) * NG + grp]
    const long long fb = ((long long)blockIdx.x * nwarps + warp) * 32;    // the warp's first filter
    const unsigned gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << (grp * G);

    double Wm[7], Wc[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) { Wm[k] = a.wi; Wc[k] = a.wi; }
    Wm[0] = a.wm0; Wc[0] = a.wc0;

    // ================= predict: THREAD per filter (lane l owns filter fb + l) ===================
    {
        const long long f = fb + lane;
        const bool live = f < a.F;
        double x[3] = {0.0, 0.0, 0.0}, P[9] = {1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0};
        double sf[7][3];
        int st = 0;
        bool run = live, wrote_prior = false;
        if (live) {
#pragma unroll
            for (int i = 0; i < 3; ++i) x[i] = a.x[f * 3 + i];
#pragma unroll
            for (int i = 0; i < 9; ++i) P[i] = a.P[f * 9 + i];
        }
#pragma unroll
        for (int k = 0; k < 7; ++k) { sf[k][0] = 0.0; sf[k][1] = 0.0; sf[k][2] = 0.0; }
        if (a.flags & kUkfPredict) {
            double sg[7][3];
            if (run && !ukf_sigma_points(x, P, a.cscale, sg)) { st = kUkfStSigma; run = false; }
            if (run) {
                const double u0 = a.u ? a.u[f * 2] : 0.0, u1 = a.u ? a.u[f * 2 + 1] : 0.0;
                const double wth = __dmul_rn(a.dt, __dadd_rn(__dmul_rn(-a.r_over_l, u0), __dmul_rn(a.r_over_l, u1)));
                double a0 = 0.0, a1 = 0.0, ss = 0.0, sc = 0.0;
#pragma unroll
                for (int k = 0; k < 7; ++k) {                   // UKFMethods.py:17-24
                    double sn, cs;
                    sincos(sg[k][2], &sn, &cs);
                    const double rc = __dmul_rn(a.half_r, cs), rs = __dmul_rn(a.half_r, sn);
                    sf[k][0] = __dadd_rn(sg[k][0], __dmul_rn(a.dt, __dadd_rn(__dmul_rn(rc, u0), __dmul_rn(rc, u1))));
                    sf[k][1] = __dadd_rn(sg[k][1], __dmul_rn(a.dt, __dadd_rn(__dmul_rn(rs, u0), __dmul_rn(rs, u1))));
                    sf[k][2] = __dadd_rn(sg[k][2], wth);
                    sincos(sf[k][2], &sn, &cs);                 // UKFMethods.py:37-45
                    ss = __fma_rn(sn, Wm[k], ss);
                    sc = __fma_rn(cs, Wm[k], sc);
                    a0 = __fma_rn(sf[k][0], Wm[k], a0);
                    a1 = __fma_rn(sf[k][1], Wm[k], a1);
                }
                x[0] = a0; x[1] = a1; x[2] = atan2(ss, sc);
                double Pn[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // lower triangle 00 10 11 20 21 22
#pragma unroll
                for (int k = 0; k < 7; ++k) {
                    const double y0 = __dsub_rn(sf[k][0], x[0]), y1 = __dsub_rn(sf[k][1], x[1]);
                    const double y2 = ukf_normalize_angle(__dsub_rn(sf[k][2], x[2]));
                    Pn[0] = __fma_rn(Wc[k], __dmul_rn(y0, y0), Pn[0]);
                    Pn[1] = __fma_rn(Wc[k], __dmul_rn(y1, y0), Pn[1]);
                    Pn[2] = __fma_rn(Wc[k], __dmul_rn(y1, y1), Pn[2]);
                    Pn[3] = __fma_rn(Wc[k], __dmul_rn(y2, y0), Pn[3]);
                    Pn[4] = __fma_rn(Wc[k], __dmul_rn(y2, y1), Pn[4]);
                    Pn[5] = __fma_rn(Wc[k], __dmul_rn(y2, y2), Pn[5]);
                }
                P[0] = __dadd_rn(Pn[0], a.Q[0]); P[1] = __dadd_rn(Pn[1], a.Q[1]); P[2] = __dadd_rn(Pn[3], a.Q[2]);
                P[3] = __dadd_rn(Pn[1], a.Q[3]); P[4] = __dadd_rn(Pn[2], a.Q[4]); P[5] = __dadd_rn(Pn[4], a.Q[5]);
                P[6] = __dadd_rn(Pn[3], a.Q[6]); P[7] = __dadd_rn(Pn[4], a.Q[7]); P[8] = __dadd_rn(Pn[5], a.Q[8]);
                if (a.flags & kUkfRedraw) {                     // a failed re-draw: the prior stands, the update is skipped
                    if (!ukf_sigma_points(x, P, a.cscale, sf)) { st = kUkfStRedraw; run = false; wrote_prior = true; }
                }
                if (run && a.sigmas_f != nullptr) {
#pragma unroll
                    for (int k = 0; k < 7; ++k)
#pragma unroll
                        for (int i = 0; i < 3; ++i) a.sigmas_f[(f * 7 + k) * 3 + i] = sf[k][i];
                }
            }
        } else if (a.flags & kUkfUpdate) {
            if (run) {
                if (a.sigmas_f != nullptr) {
#pragma unroll
                    for (int k = 0; k < 7; ++k)
#pragma unroll
                        for (int i = 0; i < 3; ++i) sf[k][i] = a.sigmas_f[(f * 7 + k) * 3 + i];
                } else if (!ukf_sigma_points(x, P, a.cscale, sf)) { st = kUkfStRedraw; run = false; }
            }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) H[i * 32 + lane] = x[i];
#pragma unroll
        for (int i = 0; i < 9; ++i) H[(3 + i) * 32 + lane] = P[i];
#pragma unroll
        for (int k = 0; k < 7; ++k)
#pragma unroll
            for (int i = 0; i < 3; ++i) H[(12 + k * 3 + i) * 32 + lane] = sf[k][i];
        Hi[lane] = st;
        Hi[32 + lane] = (run ? 1 : 0) | (wrote_prior ? 2 : 0);
    }
    __syncwarp();

    // ================= update: a GROUP of G lanes per filter, NG filters per pass ================
    if (a.flags & kUkfUpdate) {
#pragma unroll 1
    for (int pass = 0; pass < G; ++pass) {
    if (fb + (long long)pass * NG >= a.F) break;
    const int fl = pass * NG + grp;
    const long long f = fb + fl;
    const bool run = (Hi[32 + fl] & 1) != 0;
    int st = 0;
    double x[3], P[9], sf[7][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) x[i] = H[i * 32 + fl];
#pragma unroll
    for (int i = 0; i < 9; ++i) P[i] = H[(3 + i) * 32 + fl];
#pragma unroll
    for (int k = 0; k < 7; ++k)
#pragma unroll
        for (int i = 0; i < 3; ++i) sf[k][i] = H[(12 + k * 3 + i) * 32 + fl];
    // ================= update ===================================================================
    int m = 0, sel = 0;
    {
        const bool ok = run && g < a.L && (a.lm_valid == nullptr || a.lm_valid[f * a.L + g] >= a.valid_min);
        const unsigned bits = (__ballot_sync(LSLAM_FULL_MASK, ok) & gmask) >> (grp * G);
        m = __popc(bits);
        sel = (g < m) ? (int)__fns(bits, 0, g + 1) : 0;         // the g-th valid slot
    }
    const int dz = 2 * m;
    const int dzmax = __reduce_max_sync(LSLAM_FULL_MASK, dz);   // warp-uniform loop bound
    if (dzmax > 0) {
        const bool mine = g < m;                                // this lane owns measurement q = g
        const int q = g;
        // ---- hx, z_mean, residual columns of the lane's landmark (registers) ----
        double e_d[7], e_b[7], zz_d = 0.0, zz_b = 0.0, zp_d = 0.0, zp_b = 0.0;
#pragma unroll
        for (int k = 0; k < 7; ++k) { e_d[k] = 0.0; e_b[k] = 0.0; }
        if (mine) {
            const double px = a.lm_pos[(f * a.L + sel) * 2], py = a.lm_pos[(f * a.L + sel) * 2 + 1];
            double acc = 0.0, ss = 0.0, sc = 0.0;
#pragma unroll
            for (int k = 0; k < 7; ++k) {                       // UKFMethods.py:26-34, 47-57
                const double dx = __dsub_rn(px, sf[k][0]), dy = __dsub_rn(py, sf[k][1]);
                e_d[k] = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
                e_b[k] = ukf_normalize_angle(__dsub_rn(atan2(dy, dx), sf[k][2]));
                double sn, cs;
                sincos(e_b[k], &sn, &cs);
                ss = __fma_rn(sn, Wm[k], ss);
                sc = __fma_rn(cs, Wm[k], sc);
                acc = __fma_rn(e_d[k], Wm[k], acc);
            }
            zp_d = acc;
            zp_b = atan2(ss, sc);
            const double *zs = a.z + (f * a.L + sel) * (long long)a.zstride;
            if (a.zmode == 0) { zz_d = zs[0]; zz_b = zs[1]; }
            else {
                const double ox = zs[2], oy = zs[3];
                zz_d = __dsqrt_rn(__dadd_rn(__dmul_rn(ox, ox), __dmul_rn(oy, oy)));
                zz_b = atan2(oy, ox);
            }
#pragma unroll
            for (int k = 0; k < 7; ++k) {                       // residual_h (UKFMethods.py:66-71)
                e_d[k] = __dsub_rn(e_d[k], zp_d);
                e_b[k] = ukf_normalize_angle(__dsub_rn(e_b[k], zp_b));
                UKF_SM(oE + k * DZ + 2 * q) = e_d[k];
                UKF_SM(oE + k * DZ + 2 * q + 1) = e_b[k];
            }
            UKF_SM(oY + 2 * q) = __dsub_rn(zz_d, zp_d);
            UKF_SM(oY + 2 * q + 1) = ukf_normalize_angle(__dsub_rn(zz_b, zp_b));
            // ---- the lane's 3x2 block of Pxz: sum_k Wc_k * y_k[i] * e_k[j] ----
            double pd[3] = {0.0, 0.0, 0.0}, pb[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                const double y0 = __dsub_rn(sf[k][0], x[0]), y1 = __dsub_rn(sf[k][1], x[1]);
                const double y2 = ukf_normalize_angle(__dsub_rn(sf[k][2], x[2]));
                pd[0] = __fma_rn(Wc[k], __dmul_rn(y0, e_d[k]), pd[0]);
                pd[1] = __fma_rn(Wc[k], __dmul_rn(y1, e_d[k]), pd[1]);
                pd[2] = __fma_rn(Wc[k], __dmul_rn(y2, e_d[k]), pd[2]);
                pb[0] = __fma_rn(Wc[k], __dmul_rn(y0, e_b[k]), pb[0]);
                pb[1] = __fma_rn(Wc[k], __dmul_rn(y1, e_b[k]), pb[1]);
                pb[2] = __fma_rn(Wc[k], __dmul_rn(y2, e_b[k]), pb[2]);
            }
#pragma unroll
            for (int i = 0; i < 3; ++i) { UKF_SM(oPxz + i * DZ + 2 * q) = pd[i]; UKF_SM(oPxz + i * DZ + 2 * q + 1) = pb[i]; }
        }
        __syncwarp();
        // The gain K = Pxz S^-1 is never formed: with S = U^T U and the forward solves U^T [w_0 w_1 w_2 v] = [Pxz^T y],
        // K y = W^T v and K S K^T = W^T W (oracle/ukf.c pins exactly this order).
        bool spd = true;
        if constexpr (G <= 8) {
            // ---- register-resident, right-looking: lane g owns columns c0 = g and c1 = g + G of the upper triangle
            //      (entries j <= c) and, for g < 4, one right-hand side (Pxz rows 0..2, y) as an extra column ----
            const int c0 = g, c1 = g + G;
            double A0[DZ], A1[DZ], A2[DZ];                      // A0 holds G entries; the rest is never touched
            {
                double we0[7], we1[7];
#pragma unroll
                for (int k = 0; k < 7; ++k) {
                    we0[k] = c0 < dz ? __dmul_rn(Wc[k], UKF_SM(oE + k * DZ + c0)) : 0.0;
                    we1[k] = c1 < dz ? __dmul_rn(Wc[k], UKF_SM(oE + k * DZ + c1)) : 0.0;
                }
#pragma unroll
                for (int j = 0; j < DZ; ++j) {
                    double s0 = 0.0, s1 = 0.0;
                    if (j < dzmax) {
#pragma unroll
                        for (int k = 0; k < 7; ++k) {
                            const double ej = UKF_SM(oE + k * DZ + j);
                            if (j < G) s0 = __fma_rn(we0[k], ej, s0);
                            s1 = __fma_rn(we1[k], ej, s1);
                        }
                    }
                    const double rv = (j & 1) ? a.var_angle : a.var_dist;
                    if (j < G) A0[j] = (j <= c0 && c0 < dz) ? (j == c0 ? __dadd_rn(s0, rv) : s0) : 0.0;
                    A1[j] = (j <= c1 && c1 < dz) ? (j == c1 ? __dadd_rn(s1, rv) : s1) : 0.0;
                    A2[j] = (g < 4 && j < dz) ? UKF_SM(oAug + g * DZ + j) : 0.0;
                }
            }
#pragma unroll
            for (int p = 0; p < DZ; ++p) {
                if (p < dzmax) {
                    const double piv = __shfl_sync(LSLAM_FULL_MASK, p < G ? A0[p] : A1[p], (lane & ~(G - 1)) | (p % G));
                    const bool ok = piv > 0.0;
                    if (p < dz && !ok) spd = false;
                    const double rd = __ddiv_rn(1.0, __dsqrt_rn(ok ? piv : 1.0));
                    // row p of U: the lane's entries, scaled, and published for the other lanes' updates
                    if (p < G) { if (c0 > p) { A0[p] = __dmul_rn(A0[p], rd); UKF_SM(oRow + (p & 1) * DZ + c0) = A0[p]; } }
                    if (c1 > p) { A1[p] = __dmul_rn(A1[p], rd); UKF_SM(oRow + (p & 1) * DZ + c1) = A1[p]; }
                    A2[p] = __dmul_rn(A2[p], rd);
                    __syncwarp();
#pragma unroll
                    for (int j = p + 1; j < DZ; ++j) {
                        if (j < dzmax) {
                            const double u = UKF_SM(oRow + (p & 1) * DZ + j);
                            if (j < G) { if (j <= c0) A0[j] = __fma_rn(-u, A0[p], A0[j]); }
                            if (j <= c1) A1[j] = __fma_rn(-u, A1[p], A1[j]);
                            A2[j] = __fma_rn(-u, A2[p], A2[j]);
                        }
                    }
                }
            }
            if (g < 4) {
#pragma unroll
                for (int j = 0; j < DZ; ++j)
                    if (j < dz) UKF_SM(oAug + g * DZ + j) = A2[j];
            }
        } else {
            // ---- S = sum_k (Wc_k e_ki) e_kj + R: rows q and dz-1-q by lane q (lower triangle, mirrored) ----
            if (mine) {
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {
                    const int i = half ? dz - 1 - q : q;
                    double ei[7];
#pragma unroll
                    for (int k = 0; k < 7; ++k) ei[k] = __dmul_rn(Wc[k], UKF_SM(oE + k * DZ + i));
                    for (int j = 0; j <= i; ++j) {
                        double s = 0.0;
#pragma unroll
                        for (int k = 0; k < 7; ++k) s = __fma_rn(ei[k], UKF_SM(oE + k * DZ + j), s);
                        if (i == j) {
                            s = __dadd_rn(s, (i & 1) ? a.var_angle : a.var_dist);
                            UKF_SM(oSd + i) = s;
                        }
                        UKF_SM(oS + i * DZ + j) = s;            // strict lower keeps S, upper will hold U
                        UKF_SM(oS + j * DZ + i) = s;
                    }
                }
            }
            __syncwarp();
            // ---- upper Cholesky U^T U = S in the upper triangle, cooperative; row j scaled by 1 / U_jj ----
            for (int j = 0; j < dzmax; ++j) {
                if (j < dz) {
                    double s = UKF_SM(oSd + j);                 // every lane of the group: the pivot
                    for (int p = 0; p < j; ++p) { const double v = UKF_SM(oS + p * DZ + j); s = __fma_rn(-v, v, s); }
                    if (!(s > 0.0)) spd = false;
                    const double rd = __ddiv_rn(1.0, __dsqrt_rn(spd ? s : 1.0));
                    for (int i = j + 1 + g; i < dz; i += G) {   // row j of U, columns dealt to the lanes
                        double t = UKF_SM(oS + i * DZ + j);     // S[j][i] = S[i][j], read from the lower triangle
                        for (int p = 0; p < j; ++p) t = __fma_rn(-UKF_SM(oS + p * DZ + j), UKF_SM(oS + p * DZ + i), t);
                        UKF_SM(oS + j * DZ + i) = __dmul_rn(t, rd);
                    }
                    if (g == 0) UKF_SM(oRd + j) = rd;
                }
                __syncwarp();
            }
            // ---- forward solves U^T w = rhs (lanes 0..3 of the group: Pxz rows 0..2, y), right-looking in registers: as soon
            //      as w_p is known it is subtracted from every later right-hand side ----
            if (m > 0 && spd && g < 4) {
                double w[DZ];
#pragma unroll
                for (int i = 0; i < DZ; ++i) w[i] = i < dz ? UKF_SM(oAug + g * DZ + i) : 0.0;
#pragma unroll
                for (int p = 0; p < DZ; ++p) {
                    if (p < dz) {
                        w[p] = __dmul_rn(w[p], UKF_SM(oRd + p));
#pragma unroll
                        for (int i = p + 1; i < DZ; ++i)
                            if (i < dz) w[i] = __fma_rn(-UKF_SM(oS + p * DZ + i), w[p], w[i]);
                    }
                }
#pragma unroll
                for (int i = 0; i < DZ; ++i)
                    if (i < dz) UKF_SM(oAug + g * DZ + i) = w[i];
            }
        }
        if (m > 0 && !spd) { st |= kUkfStS; }
        const bool upd = m > 0 && spd;
        __syncwarp();
        // ---- the nine sums W^T W (6, symmetric) and W^T v (3), dealt to the lanes of the group ----
        for (int t = g; t < 9; t += G) {
            const int ri = t < 3 ? 0 : (t < 5 ? 1 : (t < 6 ? 2 : t - 6));
            const int rj = t < 3 ? t : (t < 5 ? t - 2 : (t < 6 ? 2 : 3));
            double s = 0.0;
            if (upd)
                for (int j = 0; j < dz; ++j) s = __fma_rn(UKF_SM(oAug + ri * DZ + j), UKF_SM(oAug + rj * DZ + j), s);
            UKF_SM(oRes + t) = s;
        }
        __syncwarp();
        // ---- x = x + W^T v ; P = P - W^T W (replicated) ----
        if (upd) {
#pragma unroll
            for (int i = 0; i < 3; ++i) x[i] = __dadd_rn(x[i], UKF_SM(oRes + 6 + i));
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const int lo = i < r ? i : r, hi = i < r ? r : i;
                    P[i * 3 + r] = __dsub_rn(P[i * 3 + r], UKF_SM(oRes + (lo == 0 ? hi : (lo == 1 ? 2 + hi : 5))));
                }
        }
        if (g == 0 && (upd || st != 0)) {                       // the posterior (and the status) back into the hand-off area
            if (upd) {
#pragma unroll
                for (int i = 0; i < 3; ++i) H[i * 32 + fl] = x[i];
#pragma unroll
                for (int i = 0; i < 9; ++i) H[(3 + i) * 32 + fl] = P[i];
            }
            Hi[fl] |= st;
        }
    }
    __syncwarp();
    }   // pass
    }   // update
    __syncwarp();
    // ================= outputs: thread per filter ================================================
    {
        const long long f = fb + lane;
        if (f < a.F) {
            const int fl = Hi[32 + lane];
            if (fl != 0) {                                      // ran, or stopped after writing the prior
#pragma unroll
                for (int i = 0; i < 3; ++i) a.x[f * 3 + i] = H[i * 32 + lane];
#pragma unroll
                for (int i = 0; i < 9; ++i) a.P[f * 9 + i] = H[(3 + i) * 32 + lane];
            }
            if (a.status != nullptr) a.status[f] = Hi[lane];
        }
    }
    // ================= fused all-gather over peer memory =========================================
    // The warp publishes what this call left in (x, P) of its 32 filters - and the slots' landmark ids - into records
    // gather_first + f of EVERY rank's gather buffer, consecutive lanes to consecutive addresses: one multimem store per
    // element when the buffers have a multicast mapping (NVSwitch replicates it), else one peer store per rank over
    // NVLink.  The stores are fire-and-forget and overlap the other warps' arithmetic; the caller orders them against the
    // readers with a cross-rank barrier after the kernel (lidar_slam_b200/dist.py).
    if (a.gather_world > 0) {
        const long long left = a.F - fb;
        const int nf = left >= 32 ? 32 : (left > 0 ? (int)left : 0);
        const size_t rec0 = (size_t)a.gather_first + (size_t)fb;
        for (int idx = lane; idx < nf * 12; idx += 32) {
            const double v = H[(idx % 12) * 32 + idx / 12];
            if (a.gather_state_mc != nullptr) multimem_store(a.gather_state_mc + rec0 * 12 + idx, v);
            else
                for (int p = 0; p < a.gather_world; ++p) a.gather_state[p][rec0 * 12 + idx] = v;
        }
        if (a.gather_ids != nullptr) {
            for (int idx = lane; idx < nf * a.gather_L; idx += 32) {
                const int32_t v = a.gather_ids[fb * a.gather_L + idx];
                if (a.gather_ids_mc != nullptr) multimem_store(a.gather_ids_mc + rec0 * a.gather_L + idx, v);
                else
                    for (int p = 0; p < a.gather_world; ++p) a.gather_idbuf[p][rec0 * a.gather_L + idx] = v;
            }
        }
    }
#undef UKF_SM
}

} // namespace lslam
