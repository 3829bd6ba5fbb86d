// ukf_aug.cuh - augmented-state UKF: pose + L landmarks in the state,
// n = 3 + 2L (403 at L = 200), 2n + 1 sigma points, dim_z = 2L.
//
// The reference's filter is a localiser with n = 3 (ukf.cuh).  BASELINE.json configs[3] asks for the same
// filterpy engine (UnscentedKalmanFilter + MerweScaledSigmaPoints, systemClass.py:20-21) with the landmark
// positions INSIDE the state: fx leaves them fixed, hx reads them from the state vector
// (oracle/ukf.py transfer_function_augmented), x_mean / residual_x treat index 2 as an angle and the rest
// linearly.  At this size one filter step is dense linear algebra, ~1.3 GFLOP of fp64, and the matrices
// (P: 1.3 MB) do not fit on chip, so the step is a sequence of batched kernels over global-memory
// matrices, all filters of a batch side by side:
//   scale+Cholesky  U^T U = (n + lambda) P        blocked right-looking: panel kernel + GEMM update
//   sigma points    X = x, x + U[k], x - U[k]; fx on the pose part
//   mean, residuals xbar (circular mean for the heading), Y = residual_x(X, xbar)
//   Pbar            = Y^T diag(Wc) Y + Q          batched GEMM
//   redraw          Cholesky of (n + lambda) Pbar, X = sigma points of (xbar, Pbar)      (filterpy >= 1.4.5)
//   hx              H[k] = (range, bearing) of every landmark of sigma point k; zhat; E = residual_h(H, zhat)
//   S               = E^T diag(Wc) E + R,   Pzx = E^T diag(Wc) Y                          batched GEMM
//   gain            K^T = S^-1 Pzx through the Cholesky of S and two blocked triangular solves
//                   (filterpy forms inv(S) explicitly; same K to rounding)
//   update          x = xbar + K residual_h(z, zhat);  P = Pbar - K (S K^T)               batched GEMM
// One GEMM form serves all of them: C[i,j] = beta C[i,j] + alpha sum_k w_k A[k,i] B[k,j] with A and B
// stored k-major ("TN"), which is how the sigma-point arrays and the factor rows lie in memory.
#pragma once
#include "lslam_core.cuh"
#include "ransac.cuh"
#include "ukf.cuh"

namespace lslam {

// ---- batched TN GEMM -------------------------------------------------------------------------
// C (M x N, ldc) = beta * C + alpha * sum_{k < K} w[k] * A[k, i] * B[k, j];  A: K x M (lda), B: K x N (ldb).
// grid = (ceil(N/64), ceil(M/64), batch); 256 threads, 4 x 4 outputs per thread; tiles staged in shared
// memory k-major so that the global loads are contiguous along i / j.
struct AugGemm {
    const double *A, *B, *w;        // w may be null (all ones); w is shared by the batch
    double *C;
    long long sA, sB, sC;           // element strides between consecutive filters
    int M, N, K, lda, ldb, ldc;
    double alpha, beta;
    const int32_t *skip;            // per-filter status: non-zero -> leave C alone (may be null)
    int upper;                      // 1: C is square and only its upper triangle (j >= i) is computed / stored
};

constexpr int kAugTile = 64, kAugBK = 16, kAugLd = kAugTile + 1;
constexpr size_t kAugGemmSmem = 2 * kAugBK * kAugLd * sizeof(double);
constexpr size_t kAugBlockSmem = (32 * 33 + 2) * sizeof(double);

__global__ void __launch_bounds__(256) aug_gemm_tn_kernel(const AugGemm g)
{
    LSLAM_SMEM_DECL;
    double (*As)[kAugLd] = reinterpret_cast<double (*)[kAugLd]>(smem_raw);
    double (*Bs)[kAugLd] = As + kAugBK;
    const int f = blockIdx.z;
    if (g.skip != nullptr && g.skip[f] != 0) return;
    if (g.upper && blockIdx.y > blockIdx.x) return;
    const double *A = g.A + (size_t)f * g.sA, *B = g.B + (size_t)f * g.sB;
    double *C = g.C + (size_t)f * g.sC;
    const int i0 = blockIdx.y * kAugTile, j0 = blockIdx.x * kAugTile;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;          // 16 x 16 threads, 4 x 4 each
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    for (int k0 = 0; k0 < g.K; k0 += kAugBK) {
        // 16 x 64 elements per tile, 256 threads: 4 elements each, contiguous along the row
        for (int e = threadIdx.x; e < kAugBK * kAugTile; e += 256) {
            const int kk = e >> 6, c = e & 63;
            const int k = k0 + kk;
            double av = 0.0, bv = 0.0;
            if (k < g.K) {
                const double wk = g.w ? g.w[k] : 1.0;
                if (i0 + c < g.M) av = wk * A[(size_t)k * g.lda + i0 + c];
                if (j0 + c < g.N) bv = B[(size_t)k * g.ldb + j0 + c];
            }
            As[kk][c] = av; Bs[kk][c] = bv;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < kAugBK; ++kk) {
            double a4[4], b4[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) a4[a] = As[kk][ty + 16 * a];
#pragma unroll
            for (int b = 0; b < 4; ++b) b4[b] = Bs[kk][tx + 16 * b];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fma(a4[a], b4[b], acc[a][b]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int i = i0 + ty + 16 * a;
        if (i >= g.M) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = j0 + tx + 16 * b;
            if (j >= g.N || (g.upper && j < i)) continue;
            double *c = C + (size_t)i * g.ldc + j;
            *c = (g.beta == 0.0 ? 0.0 : g.beta * *c) + g.alpha * acc[a][b];
        }
    }
}

// ---- blocked Cholesky (upper, U^T U = A, in place; only the upper triangle is read / written) --------
// Panel step jb (width nb <= 32), one CTA per matrix: factor the nb x nb diagonal block, then the panel
// row U[jb:jb+nb, jb+nb:n] = D^-T A[jb:jb+nb, jb+nb:n].  The trailing update A22 -= panel^T panel is the
// GEMM above.  A non-positive pivot sets bit `fail_bit` of status[f] (scipy raises LinAlgError) and the
// matrix is left alone from then on.
__global__ void __launch_bounds__(256) aug_chol_panel_kernel(double *Abase, long long sA, int n, int lda, int jb, int nb,
                                                             int32_t *status, int fail_bit)
{
    LSLAM_SMEM_DECL;
    double (*D)[33] = reinterpret_cast<double (*)[33]>(smem_raw);
    int &bad = *reinterpret_cast<int *>(smem_raw + 32 * 33 * sizeof(double));
    const int f = blockIdx.x;
    if (status[f] != 0) return;
    double *A = Abase + (size_t)f * sA;
    const int tid = threadIdx.x;
    if (tid == 0) bad = 0;
    for (int e = tid; e < nb * nb; e += blockDim.x) {
        const int r = e / nb, c = e % nb;
        D[r][c] = c >= r ? A[(size_t)(jb + r) * lda + jb + c] : 0.0;
    }
    __syncthreads();
    // unblocked factorisation of the diagonal block by the first warp (row by row, lanes over columns)
    if (tid < 32) {
        for (int j = 0; j < nb; ++j) {
            double s = D[j][j];
            for (int k = 0; k < j; ++k) s = fma(-D[k][j], D[k][j], s);
            if (!(s > 0.0)) { if (tid == 0) bad = 1; break; }
            const double d = sqrt(s);
            __syncwarp();
            if (tid == 0) D[j][j] = d;
            const int c = j + 1 + tid;
            if (c < nb) {
                double v = D[j][c];
                for (int k = 0; k < j; ++k) v = fma(-D[k][j], D[k][c], v);
                D[j][c] = v / d;
            }
            __syncwarp();
        }
    }
    __syncthreads();
    if (bad) { if (tid == 0) status[f] |= fail_bit; return; }
    for (int e = tid; e < nb * nb; e += blockDim.x) {
        const int r = e / nb, c = e % nb;
        if (c >= r) A[(size_t)(jb + r) * lda + jb + c] = D[r][c];
    }
    // panel row: column c of the panel solves D^T u = a  (forward substitution down the nb rows)
    for (int c = jb + nb + tid; c < n; c += blockDim.x) {
        double u[32];
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            if (r < nb) {
                double v = A[(size_t)(jb + r) * lda + c];
                for (int k = 0; k < r; ++k) v = fma(-D[k][r], u[k], v);
                u[r] = v / D[r][r];
            }
        }
#pragma unroll
        for (int r = 0; r < 32; ++r)
            if (r < nb) A[(size_t)(jb + r) * lda + c] = u[r];
    }
}

// ---- blocked triangular solves for K^T = S^-1 Pzx with S = U^T U ----------------------------------
// forward (U^T W = Pzx) block step: rows [jb, jb+nb) of W, after the GEMM subtracted the rows above:
//   W[jb+r, :] = (W[jb+r, :] - sum_{k<r} U[jb+k, jb+r] W[jb+k, :]) / U[jb+r, jb+r]
// backward (U K^T = W) block step, rows descending inside the block:
//   X[jb+r, :] = (X[jb+r, :] - sum_{k>r} U[jb+r, jb+k] X[jb+k, :]) / U[jb+r, jb+r]
__global__ void __launch_bounds__(256) aug_trsm_block_kernel(const double *Ubase, long long sU, int ldu, double *Xbase,
                                                             long long sX, int ldx, int ncols, int jb, int nb, int backward,
                                                             const int32_t *status)
{
    LSLAM_SMEM_DECL;
    double (*D)[33] = reinterpret_cast<double (*)[33]>(smem_raw);
    const int f = blockIdx.y;
    if (status[f] != 0) return;
    const double *U = Ubase + (size_t)f * sU;
    double *X = Xbase + (size_t)f * sX;
    for (int e = threadIdx.x; e < nb * nb; e += blockDim.x) {
        const int r = e / nb, c = e % nb;
        D[r][c] = c >= r ? U[(size_t)(jb + r) * ldu + jb + c] : 0.0;
    }
    __syncthreads();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncols) return;
    double x[32];
    if (!backward) {
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            if (r < nb) {
                double v = X[(size_t)(jb + r) * ldx + c];
                for (int k = 0; k < r; ++k) v = fma(-D[k][r], x[k], v);
                x[r] = v / D[r][r];
            }
        }
    } else {
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) {
            const int r = nb - 1 - rr;
            if (r >= 0) {
                double v = X[(size_t)(jb + r) * ldx + c];
                for (int k = r + 1; k < nb; ++k) v = fma(-D[r][k], x[k], v);
                x[r] = v / D[r][r];
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 32; ++r)
        if (r < nb) X[(size_t)(jb + r) * ldx + c] = x[r];
}

// ---- elementwise stages -----------------------------------------------------------------------
struct AugArgs {
    int F, L, n, m, dz;             // filters in this batch, landmarks, 3 + 2L, 2n + 1, 2L
    double *x, *P;                  // [F, n], [F, n, n] in / out
    const double *u, *z;            // [F, 2], [F, L, 2]
    int32_t *status;                // [F]
    double *U, *X, *Y, *H, *S, *ST, *Kt, *T, *Pbar, *xbar, *zhat, *wm, *wc;   // workspace
    double dt, half_r, r_over_l, cscale, var_dist, var_angle;
    double Q[9];
};

// Merwe weights (MerweScaledSigmaPoints._compute_weights): wi everywhere, wm0 / wc0 on the central point
__global__ void aug_weights_kernel(double *wm, double *wc, int m, double wi, double wm0, double wc0)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < m) { wm[k] = k ? wi : wm0; wc[k] = k ? wi : wc0; }
}

// dst = scale * src on the upper triangle (lower zeroed), n x n
__global__ void aug_scale_upper_kernel(const double *src, long long s_src, double *dst, long long s_dst, int n, double scale,
                                       const int32_t *status)
{
    const int f = blockIdx.y;
    if (status[f] != 0) return;
    src += (size_t)f * s_src; dst += (size_t)f * s_dst;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)n * n; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / n), c = (int)(e % n);
        dst[e] = c >= r ? __dmul_rn(scale, src[e]) : 0.0;
    }
}

// sigma points of (x0, U) [MerweScaledSigmaPoints.sigma_points]: rows x0, x0 + U[k], x0 - U[k]
__global__ void aug_sigmas_kernel(const AugArgs a, const double *x0)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int n = a.n;
    const double *x = x0 + (size_t)f * n, *U = a.U + (size_t)f * n * n;
    double *X = a.X + (size_t)f * a.m * n;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)a.m * n; e += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(e / n), c = (int)(e % n);
        double v = x[c];
        if (k >= 1 && k <= n) v = __dadd_rn(v, U[(size_t)(k - 1) * n + c]);
        else if (k > n) v = __dsub_rn(v, U[(size_t)(k - n - 1) * n + c]);
        X[e] = v;
    }
}

// fx on the pose columns of every sigma point: the differential-drive transition (UKFMethods.py:17-24);
// landmarks unchanged
__global__ void aug_fx_kernel(const AugArgs a)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a.m) return;
    double *s = a.X + ((size_t)f * a.m + k) * a.n;
    const double ul = a.u ? a.u[2 * f] : 0.0, ur = a.u ? a.u[2 * f + 1] : 0.0;
    double sn, cs;
    sincos(s[2], &sn, &cs);
    // x' = x + dt * B(theta) u,  B = [[R/2 cos, R/2 cos], [R/2 sin, R/2 sin], [-R/L, R/L]]  (oracle/ukf.py order)
    const double b0 = __dmul_rn(a.half_r, cs), b1 = __dmul_rn(a.half_r, sn);
    const double d0 = __dadd_rn(__dmul_rn(b0, ul), __dmul_rn(b0, ur));
    const double d1 = __dadd_rn(__dmul_rn(b1, ul), __dmul_rn(b1, ur));
    const double d2 = __dadd_rn(__dmul_rn(-a.r_over_l, ul), __dmul_rn(a.r_over_l, ur));
    s[0] = __dadd_rn(s[0], __dmul_rn(a.dt, d0));
    s[1] = __dadd_rn(s[1], __dmul_rn(a.dt, d1));
    s[2] = __dadd_rn(s[2], __dmul_rn(a.dt, d2));
}

// weighted column means of V [m x cols] with circular means on the angle columns, then residuals in place
// or into R.  angle columns: state -> {2}; measurement -> odd columns.  One CTA per filter.
__global__ void __launch_bounds__(256) aug_mean_resid_kernel(const AugArgs a, const double *Vbase, int cols, int is_meas,
                                                             int mean_given, double *mean_base, double *Rbase)
{
    const int f = blockIdx.x;
    if (a.status[f] != 0) return;
    const double *V = Vbase + (size_t)f * a.m * cols;
    double *mean = mean_base + (size_t)f * cols, *R = Rbase + (size_t)f * a.m * cols;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
        const bool ang = is_meas ? (c & 1) : (c == 2);
        double mu;
        if (mean_given) {
            mu = mean[c];
        } else {
            double s0 = 0.0, s1 = 0.0;
            for (int k = 0; k < a.m; ++k) {
                const double v = V[(size_t)k * cols + c], w = a.wm[k];
                if (ang) { double sn, cs; sincos(v, &sn, &cs); s0 = fma(sn, w, s0); s1 = fma(cs, w, s1); }
                else s0 = fma(v, w, s0);
            }
            mu = ang ? atan2(s0, s1) : s0;
            mean[c] = mu;
        }
        for (int k = 0; k < a.m; ++k) {
            double d = __dsub_rn(V[(size_t)k * cols + c], mu);
            if (ang) d = ukf_normalize_angle(d);
            R[(size_t)k * cols + c] = d;
        }
    }
}

// hx of every sigma point: (range, bearing) of every landmark read from the state (oracle/ukf.py
// transfer_function_augmented, UKFMethods.py:26-34)
__global__ void aug_hx_kernel(const AugArgs a)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)a.m * a.L; e += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(e / a.L), l = (int)(e % a.L);
        const double *s = a.X + ((size_t)f * a.m + k) * a.n;
        const double dx = __dsub_rn(s[3 + 2 * l], s[0]), dy = __dsub_rn(s[4 + 2 * l], s[1]);
        double *h = a.H + ((size_t)f * a.m + k) * a.dz + 2 * l;
        h[0] = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
        h[1] = ukf_normalize_angle(__dsub_rn(atan2(dy, dx), s[2]));
    }
}

// C = diag block: Pbar <- Q on the pose block, zero elsewhere; S <- R = diag(var_dist, var_angle, ...)
__global__ void aug_fill_noise_kernel(const AugArgs a, double *Cbase, int dim, int is_meas)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    double *C = Cbase + (size_t)f * dim * dim;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)dim * dim; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / dim), c = (int)(e % dim);
        double v = 0.0;
        if (is_meas) { if (r == c) v = (r & 1) ? a.var_angle : a.var_dist; }
        else if (r < 3 && c < 3) v = a.Q[r * 3 + c];
        C[e] = v;
    }
}

__global__ void aug_transpose_kernel(const double *src, long long s_src, double *dst, long long s_dst, int dim,
                                     const int32_t *status)
{
    const int f = blockIdx.y;
    if (status[f] != 0) return;
    src += (size_t)f * s_src; dst += (size_t)f * s_dst;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)dim * dim; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / dim), c = (int)(e % dim);
        dst[(size_t)c * dim + r] = src[e];
    }
}

// x = xbar + K residual_h(z, zhat) (K^T stored: Kt [dz x n]); one CTA per filter
__global__ void __launch_bounds__(256) aug_update_x_kernel(const AugArgs a, double *xnew_base)
{
    LSLAM_SMEM_DECL;
    double *ysh = reinterpret_cast<double *>(smem_raw);
    const int f = blockIdx.x;
    if (a.status[f] != 0) return;
    const double *zh = a.zhat + (size_t)f * a.dz, *z = a.z + (size_t)f * a.dz;
    for (int j = threadIdx.x; j < a.dz; j += blockDim.x) {
        double d = __dsub_rn(z[j], zh[j]);
        if (j & 1) d = ukf_normalize_angle(d);
        ysh[j] = d;
    }
    __syncthreads();
    const double *Kt = a.Kt + (size_t)f * a.dz * a.n, *xb = a.xbar + (size_t)f * a.n;
    double *xn = xnew_base + (size_t)f * a.n;
    for (int i = threadIdx.x; i < a.n; i += blockDim.x) {
        double s = 0.0;
        for (int j = 0; j < a.dz; ++j) s = fma(Kt[(size_t)j * a.n + i], ysh[j], s);
        xn[i] = __dadd_rn(xb[i], s);
    }
}

// commit: filters whose status is still 0 take (xnew, Pnew); the others keep their state
__global__ void aug_commit_kernel(const AugArgs a, const double *xnew, const double *Pnew)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const long long nn = (long long)a.n * a.n;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nn; e += (long long)gridDim.x * blockDim.x) {
        a.P[(size_t)f * nn + e] = Pnew[(size_t)f * nn + e];
        if (e < a.n) a.x[(size_t)f * a.n + e] = xnew[(size_t)f * a.n + e];
    }
}

} // namespace lslam
