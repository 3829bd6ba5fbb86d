/* C restatement of the UKF predict/update of the path, with the OPERATION ORDER PINNED.
 *
 * TEST INFRASTRUCTURE (see oracle/__init__.py): checker for the CUDA kernel (csrc/ukf.cuh) and
 * part of the timed CPU baseline.  PARITY UNPINNED against filterpy (absent, un-pinned, and the
 * reference's UKF modules do not parse); oracle/ukf.py is the literal NumPy restatement of
 * filterpy 1.4.x + the repaired callbacks of /root/reference/UKFMethods.py:10-71, and this file
 * is the same algorithm with every sum, product and solve written out so that the GPU can follow
 * it operation for operation (tests/test_oracle_ukf.py checks this file against oracle/ukf.py).
 *
 * Pinned choices (n = 3 state, L landmarks, dz = 2L, sigma index k = 0..6):
 *   weights    lam = alpha^2 (n + kappa) - n ; Wm_k = Wc_k = 0.5/(n + lam) ; Wm_0 = lam/(n + lam) ;
 *              Wc_0 = Wm_0 + (1 - alpha^2 + beta)
 *   sigma pts  A = (n + lam) P ; U = upper Cholesky read from the UPPER triangle of A (scipy),
 *              sums s = fma(-U[p][j], U[p][i], s) for p ascending ; sigma_0 = x ;
 *              sigma_{k+1} = x + U[k,:] ; sigma_{n+k+1} = x - U[k,:]
 *   fx         UKFMethods.py:17-24: rc = (R/2) cos th ; rs = (R/2) sin th ; every product and sum
 *              rounded on its own: x' = x + dt (rc vl + rc vr), y' = y + dt (rs vl + rs vr),
 *              th' = th + dt ((-R/L) vl + (R/L) vr)
 *   means      acc = fma(value_k, Wm_k, acc) for k ascending from acc = 0 ; angles through
 *              atan2(sum Wm sin, sum Wm cos)                                  (UKFMethods.py:37-57)
 *   residuals  a - b, angles through normalize_angle = Python floor-mod 2pi then fold to (-pi, pi]
 *   covariances  C_ij = fma(Wc_k, v_i w_j, C_ij) for k ascending from 0 for P and Pxz ; S with the weight folded into the
 *              row's residual first, S_ij = fma(Wc_k e_ki, e_kj, S_ij) for j <= i, mirrored ; + Q / + R after the sum
 *   gain       never formed.  S = U^T U (upper Cholesky, the recurrence above, row j scaled by rd_j = 1 / U_jj) ; one
 *              forward solve U^T [w_0 w_1 w_2 v] = [Pxz_0 Pxz_1 Pxz_2 y]^T, s = fma(-U_pj, w_p, s) for p ascending, then
 *              w_j = s rd_j.  With K = Pxz S^-1 = W^T U^-T:  K y = W^T v  and  K S K^T = W^T W  exactly, so
 *   update     x = x + W^T v ; P = P - W^T W, ascending fma sums over the measurement index (filterpy's
 *              x + K y, P - K S K^T without the back substitution and the two products S K^T, K (S K^T))
 * libm: sin, cos, atan2, sqrt, fmod.  sqrt and fmod are exact in IEEE terms; sin/cos/atan2 differ
 * from the CUDA math library by an ulp or so, which is what the 1e-5 parity tolerance absorbs.
 *
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off: nothing is fused unless written as fma()).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define UKF_N 3
#define UKF_NS 7
#define UKF_LMAX 32

enum { UKF_PREDICT = 1, UKF_UPDATE = 2, UKF_REDRAW = 4 };
enum { UKF_ST_SIGMA = 1, UKF_ST_REDRAW = 2, UKF_ST_S = 4 };

typedef struct {
    double dt, wheel_r, base_l, alpha, beta, kappa, var_dist, var_angle, Q[9];
} ukf_params;

static const double kPi = 3.141592653589793;

static double normalize_angle(double a)
{
    const double two_pi = 2 * kPi;
    double m = fmod(a, two_pi);                      /* Python float %: fmod, then fix the sign */
    if (m != 0.0) { if (m < 0.0) m += two_pi; }
    else m = 0.0;                                    /* copysign(0, two_pi) */
    if (m > kPi) m -= two_pi;
    return m;
}

/* upper Cholesky of the n x n matrix c*A (row-major, lda), upper triangle read; U row-major n x n */
static int chol_upper_scaled(const double *A, int n, int lda, double c, double *U, int ldu)
{
    for (int j = 0; j < n; ++j) {
        double s = c * A[j * lda + j];
        for (int p = 0; p < j; ++p) s = fma(-U[p * ldu + j], U[p * ldu + j], s);
        if (!(s > 0.0)) return 1;
        const double d = sqrt(s);
        U[j * ldu + j] = d;
        for (int i = j + 1; i < n; ++i) {
            double t = c * A[j * lda + i];
            for (int p = 0; p < j; ++p) t = fma(-U[p * ldu + j], U[p * ldu + i], t);
            U[j * ldu + i] = t / d;
        }
        for (int i = 0; i < j; ++i) U[j * ldu + i] = 0.0;
    }
    return 0;
}

/* upper Cholesky S = U^T U of the n x n matrix A (upper triangle read); row j is scaled by rd[j] = 1 / U[j][j] */
static int chol_upper_rd(const double *A, int n, int lda, double *U, int ldu, double *rd)
{
    for (int j = 0; j < n; ++j) {
        double s = A[j * lda + j];
        for (int p = 0; p < j; ++p) s = fma(-U[p * ldu + j], U[p * ldu + j], s);
        if (!(s > 0.0)) return 1;
        const double d = sqrt(s);
        U[j * ldu + j] = d;
        rd[j] = 1.0 / d;
        for (int i = j + 1; i < n; ++i) {
            double t = A[j * lda + i];
            for (int p = 0; p < j; ++p) t = fma(-U[p * ldu + j], U[p * ldu + i], t);
            U[j * ldu + i] = t * rd[j];
        }
    }
    return 0;
}

static int sigma_points(const double x[3], const double P[9], double c, double sg[UKF_NS][3])
{
    double U[9];
    if (chol_upper_scaled(P, 3, 3, c, U, 3)) return 1;
    for (int i = 0; i < 3; ++i) sg[0][i] = x[i];
    for (int k = 0; k < 3; ++k)
        for (int i = 0; i < 3; ++i) {
            sg[k + 1][i] = x[i] + U[k * 3 + i];
            sg[3 + k + 1][i] = x[i] - U[k * 3 + i];
        }
    return 0;
}

static void ukf_one(double x[3], double P[9], const double u[2], const double *z, int zstride, int zmode,
                    const double *lm_pos, const int32_t *lm_valid, int valid_min, int L,
                    const ukf_params *pr, int flags, int32_t *status, double *sigmas_io)
{
    const int n = UKF_N;
    const double lam = pr->alpha * pr->alpha * (n + pr->kappa) - n;
    const double c = n + lam;
    double Wm[UKF_NS], Wc[UKF_NS];
    for (int k = 0; k < UKF_NS; ++k) Wm[k] = Wc[k] = 0.5 / c;
    Wm[0] = lam / c;
    Wc[0] = lam / c + (1 - pr->alpha * pr->alpha + pr->beta);
    double sf[UKF_NS][3];
    int st = 0;
    *status = 0;

    if (flags & UKF_PREDICT) {
        double sg[UKF_NS][3];
        if (sigma_points(x, P, c, sg)) { *status = UKF_ST_SIGMA; return; }
        const double h = pr->wheel_r / 2.0, w = pr->wheel_r / pr->base_l;
        for (int k = 0; k < UKF_NS; ++k) {
            const double rc = h * cos(sg[k][2]), rs = h * sin(sg[k][2]);
            sf[k][0] = sg[k][0] + pr->dt * (rc * u[0] + rc * u[1]);
            sf[k][1] = sg[k][1] + pr->dt * (rs * u[0] + rs * u[1]);
            sf[k][2] = sg[k][2] + pr->dt * ((-1.0 * w) * u[0] + (1.0 * w) * u[1]);
        }
        double xm[3], a0 = 0.0, a1 = 0.0, ss = 0.0, sc = 0.0;
        for (int k = 0; k < UKF_NS; ++k) {
            ss = fma(sin(sf[k][2]), Wm[k], ss);
            sc = fma(cos(sf[k][2]), Wm[k], sc);
            a0 = fma(sf[k][0], Wm[k], a0);
            a1 = fma(sf[k][1], Wm[k], a1);
        }
        xm[0] = a0; xm[1] = a1; xm[2] = atan2(ss, sc);
        double Pn[9] = {0};
        for (int k = 0; k < UKF_NS; ++k) {
            double y[3] = { sf[k][0] - xm[0], sf[k][1] - xm[1], normalize_angle(sf[k][2] - xm[2]) };
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j <= i; ++j) Pn[i * 3 + j] = fma(Wc[k], y[i] * y[j], Pn[i * 3 + j]);
        }
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < i; ++j) Pn[j * 3 + i] = Pn[i * 3 + j];
        for (int i = 0; i < 9; ++i) Pn[i] = Pn[i] + pr->Q[i];
        memcpy(x, xm, sizeof xm);
        memcpy(P, Pn, sizeof Pn);
        if (flags & UKF_REDRAW) {
            if (sigma_points(x, P, c, sf)) { *status = UKF_ST_REDRAW; return; }
        }
        if (sigmas_io) memcpy(sigmas_io, sf, sizeof sf);
    } else if (flags & UKF_UPDATE) {
        if (sigmas_io) memcpy(sf, sigmas_io, sizeof sf);
        else if (sigma_points(x, P, c, sf)) { *status = UKF_ST_REDRAW; return; }
    }
    if (!(flags & UKF_UPDATE)) return;

    /* valid measurement slots, in order */
    int sel[UKF_LMAX], m = 0;
    for (int l = 0; l < L && m < UKF_LMAX; ++l)
        if (lm_valid == 0 || lm_valid[l] >= valid_min) sel[m++] = l;
    if (m == 0) return;
    const int dz = 2 * m;
    double sh[UKF_NS][2 * UKF_LMAX], zp[2 * UKF_LMAX], zz[2 * UKF_LMAX];
    for (int q = 0; q < m; ++q) {
        const int l = sel[q];
        const double px = lm_pos[2 * l], py = lm_pos[2 * l + 1];
        for (int k = 0; k < UKF_NS; ++k) {                       /* UKFMethods.py:29-33 */
            const double dx = px - sf[k][0], dy = py - sf[k][1];
            sh[k][2 * q] = sqrt(dx * dx + dy * dy);
            sh[k][2 * q + 1] = normalize_angle(atan2(dy, dx) - sf[k][2]);
        }
        double a = 0.0, ss = 0.0, sc = 0.0;                     /* UKFMethods.py:51-56 */
        for (int k = 0; k < UKF_NS; ++k) {
            ss = fma(sin(sh[k][2 * q + 1]), Wm[k], ss);
            sc = fma(cos(sh[k][2 * q + 1]), Wm[k], sc);
            a = fma(sh[k][2 * q], Wm[k], a);
        }
        zp[2 * q] = a;
        zp[2 * q + 1] = atan2(ss, sc);
        const double *zs = z + (size_t)l * zstride;
        if (zmode == 0) { zz[2 * q] = zs[0]; zz[2 * q + 1] = zs[1]; }
        else {              /* a line record: measured range/bearing of its inlier centroid (ox, oy) */
            const double ox = zs[2], oy = zs[3];
            zz[2 * q] = sqrt(ox * ox + oy * oy);
            zz[2 * q + 1] = atan2(oy, ox);
        }
    }
    /* residuals e_k = residual_h(sh_k, zp), y_k = residual_x(sf_k, x) */
    double e[UKF_NS][2 * UKF_LMAX], yk[UKF_NS][3];
    for (int k = 0; k < UKF_NS; ++k) {
        for (int q = 0; q < m; ++q) {
            e[k][2 * q] = sh[k][2 * q] - zp[2 * q];
            e[k][2 * q + 1] = normalize_angle(sh[k][2 * q + 1] - zp[2 * q + 1]);
        }
        yk[k][0] = sf[k][0] - x[0];
        yk[k][1] = sf[k][1] - x[1];
        yk[k][2] = normalize_angle(sf[k][2] - x[2]);
    }
    static const int LD = 2 * UKF_LMAX;
    double S[2 * UKF_LMAX * 2 * UKF_LMAX], G[2 * UKF_LMAX * 2 * UKF_LMAX], Pxz[3 * 2 * UKF_LMAX];
    for (int i = 0; i < dz; ++i) {
        double we[UKF_NS];
        for (int k = 0; k < UKF_NS; ++k) we[k] = Wc[k] * e[k][i];
        for (int j = 0; j <= i; ++j) {
            double s = 0.0;
            for (int k = 0; k < UKF_NS; ++k) s = fma(we[k], e[k][j], s);
            if (i == j) s = s + ((i & 1) ? pr->var_angle : pr->var_dist);
            S[i * LD + j] = s;
            S[j * LD + i] = s;
        }
    }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < dz; ++j) {
            double s = 0.0;
            for (int k = 0; k < UKF_NS; ++k) s = fma(Wc[k], yk[k][i] * e[k][j], s);
            Pxz[i * LD + j] = s;
        }
    double yr[2 * UKF_LMAX];
    for (int q = 0; q < m; ++q) {
        yr[2 * q] = zz[2 * q] - zp[2 * q];
        yr[2 * q + 1] = normalize_angle(zz[2 * q + 1] - zp[2 * q + 1]);
    }
    /* S = U^T U; forward solves U^T w_r = Pxz_r (r = 0..2) and U^T v = y: K is never formed (header) */
    double rd[2 * UKF_LMAX];
    if (chol_upper_rd(S, dz, LD, G, LD, rd)) { *status = st | UKF_ST_S; return; }
    double w[4][2 * UKF_LMAX];
    for (int r = 0; r < 4; ++r)
        for (int i = 0; i < dz; ++i) {
            double s = r < 3 ? Pxz[r * LD + i] : yr[i];
            for (int p = 0; p < i; ++p) s = fma(-G[p * LD + i], w[r][p], s);
            w[r][i] = s * rd[i];
        }
    double xn[3], Pn[9];
    for (int i = 0; i < 3; ++i) {
        double s = 0.0;
        for (int j = 0; j < dz; ++j) s = fma(w[i][j], w[3][j], s);
        xn[i] = x[i] + s;
    }
    for (int i = 0; i < 3; ++i)
        for (int r = 0; r < 3; ++r) {
            double s = 0.0;
            for (int j = 0; j < dz; ++j) s = fma(w[i][j], w[r][j], s);
            Pn[i * 3 + r] = P[i * 3 + r] - s;
        }
    memcpy(x, xn, sizeof xn);
    memcpy(P, Pn, sizeof Pn);
}

/* Batched entry: F independent filters.  params = dt, wheel_r, base_l, alpha, beta, kappa,
 * var_dist, var_angle, Q[9] (17 doubles).  Same argument meaning as lslam_ukf_step. */
int oracle_ukf_step(double *x, double *P, const double *u, const double *z, int32_t zstride, int32_t zmode,
                    const double *lm_pos, const int32_t *lm_valid, int32_t valid_min, int32_t F, int32_t L,
                    const double *params, int32_t flags, int32_t *status, double *sigmas_f, int32_t nthreads)
{
    if (F < 0 || L < 0 || L > UKF_LMAX || !x || !P || !params || !status) return -1;
    ukf_params pr;
    memcpy(&pr, params, sizeof pr);
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
#else
    (void)nthreads;
#endif
    for (int f = 0; f < F; ++f) {
        static const double zero2[2] = {0.0, 0.0};
        ukf_one(x + 3 * (size_t)f, P + 9 * (size_t)f, u ? u + 2 * (size_t)f : zero2,
                z ? z + (size_t)f * L * zstride : 0, zstride, zmode,
                lm_pos ? lm_pos + (size_t)f * L * 2 : 0, lm_valid ? lm_valid + (size_t)f * L : 0, valid_min, L,
                &pr, flags, status + f, sigmas_f ? sigmas_f + (size_t)f * UKF_NS * 3 : 0);
    }
    return 0;
}
