/* ransac_ols.c - C twin of oracle/ransac_ols.py (TEST INFRASTRUCTURE, see oracle/__init__.py):
 * the OLS-flavour RANSAC of /root/reference/line_detection.py:51-73, i.e. scikit-learn's
 * RANSACRegressor(residual_threshold=3) with a LinearRegression base estimator, per window of
 * PNT_NBR points.  Same pinned operation order as the Python restatement (bit-identical to it) and as
 * the CUDA kernel (csrc/ransac_ols.cuh); pinned to the real scikit-learn through
 * tests/golden/ols_golden.npz (tests/test_oracle_ols.py). */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

static void ols_sample_pair(uint64_t seed, uint64_t scan, uint32_t window, uint32_t trial, uint32_t n, int *i, int *j)
{
    /* the library's sampler: identical stream layout to the LineModelND flavour (ransac_line.c) */
    uint32_t ctr[4] = { trial >> 1, window, (uint32_t)scan, (uint32_t)(scan >> 32) };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) }, r[4];
    oracle_philox4x32_10(ctr, key, r);
    uint32_t u = (trial & 1u) ? r[2] : r[0], v = (trial & 1u) ? r[3] : r[1];
    uint32_t a = (uint32_t)(((uint64_t)u * n) >> 32), b = (uint32_t)(((uint64_t)v * (n - 1u)) >> 32);
    if (b >= a) ++b;
    *i = (int)a; *j = (int)b;
}

/* fixed 32-lane summation tree (oracle/ransac_line.py tree32_sum) over f(k) for k with use[k] */
#define TREE32(result, n, use_expr, value_expr)                                        \
    do {                                                                                \
        double lane_[32];                                                               \
        for (int l_ = 0; l_ < 32; ++l_) lane_[l_] = 0.0;                                \
        for (int k = 0; k < (n); ++k) if (use_expr) lane_[k & 31] = lane_[k & 31] + (value_expr); \
        for (int h_ = 16; h_ >= 1; h_ >>= 1)                                            \
            for (int l_ = 0; l_ < h_; ++l_) lane_[l_] = lane_[l_] + lane_[l_ + h_];     \
        (result) = lane_[0];                                                            \
    } while (0)

static double dynamic_max_trials(int n_inliers, int n_samples, double probability)
{
    double ratio = (double)n_inliers / (double)n_samples;
    double nom = fmax(DBL_EPSILON, 1 - probability);
    double denom = fmax(DBL_EPSILON, 1 - ratio * ratio);
    if (nom == 1) return 0;
    if (denom == 1) return INFINITY;
    return fabs(ceil(log(nom) / log(denom)));
}

#define BIT(mask, k) (((mask)[(k) >> 5] >> ((k) & 31)) & 1u)

/* one window; p = interleaved x,y; returns n_inliers or -3 (sklearn raises ValueError) */
static int ols_window(const double *p, int n, int trials, double thr, double probability, uint64_t seed,
                      uint64_t scan, uint32_t window, const int32_t *pairs, double fit[8], int32_t *ntrials,
                      int32_t *best_trial, uint32_t *mask, uint32_t *cur)
{
    const int W = (n + 31) / 32;
    int n_best = 1, best = -1, n_trials = 0, have = 0;
    double score_best = -INFINITY, limit = (double)trials;
    while ((double)n_trials < limit) {
        int i, j;
        if (pairs) { i = pairs[2 * n_trials]; j = pairs[2 * n_trials + 1]; }
        else ols_sample_pair(seed, scan, window, (uint32_t)n_trials, (uint32_t)n, &i, &j);
        ++n_trials;
        const double xi = p[2 * i], yi = p[2 * i + 1], xj = p[2 * j], yj = p[2 * j + 1];
        const double x0 = (xi + xj) / 2.0, y0 = (yi + yj) / 2.0;
        const double xci = xi - x0, xcj = xj - x0, yci = yi - y0, ycj = yj - y0;
        const double den = xci * xci + xcj * xcj, num = xci * yci + xcj * ycj;
        const double coef = den != 0 ? num / den : 0.0, icpt = y0 - x0 * coef;
        int cnt = 0;
        memset(cur, 0, sizeof(uint32_t) * (size_t)W);
        for (int k = 0; k < n; ++k) {
            double r = fabs(p[2 * k + 1] - (p[2 * k] * coef + icpt));
            if (r <= thr) { ++cnt; cur[k >> 5] |= 1u << (k & 31); }
        }
        if (cnt < n_best) continue;
        double score;
        if (cnt < 2) score = NAN;
        else {
            double sy, numr, denr;
            TREE32(sy, n, BIT(cur, k), p[2 * k + 1]);
            const double ym = sy / cnt;
            TREE32(numr, n, BIT(cur, k), (p[2 * k + 1] - (p[2 * k] * coef + icpt)) * (p[2 * k + 1] - (p[2 * k] * coef + icpt)));
            TREE32(denr, n, BIT(cur, k), (p[2 * k + 1] - ym) * (p[2 * k + 1] - ym));
            score = (numr != 0 && denr != 0) ? 1.0 - numr / denr : (numr != 0 ? 0.0 : 1.0);
        }
        if (cnt == n_best && score < score_best) continue;
        n_best = cnt; score_best = score; best = n_trials - 1; have = 1;
        memcpy(mask, cur, sizeof(uint32_t) * (size_t)W);
        limit = fmin(limit, dynamic_max_trials(n_best, n, probability));
    }
    *ntrials = n_trials;
    double xmin = INFINITY, xmax = -INFINITY;
    int anynan = 0;
    for (int k = 0; k < n; ++k) { if (p[2 * k] != p[2 * k]) anynan = 1; xmin = fmin(xmin, p[2 * k]); xmax = fmax(xmax, p[2 * k]); }
    if (anynan) xmin = xmax = NAN;                       /* np.min / np.max propagate NaN (line_detection.py:66) */
    if (!have) {
        for (int k = 0; k < 8; ++k) fit[k] = NAN;
        fit[3] = xmin; fit[4] = xmax;
        *best_trial = -1;
        memset(mask, 0, sizeof(uint32_t) * (size_t)W);
        return -3;
    }
    double sx, sy, den, num;
    TREE32(sx, n, BIT(mask, k), p[2 * k]);
    TREE32(sy, n, BIT(mask, k), p[2 * k + 1]);
    const double x0 = sx / n_best, y0 = sy / n_best;
    TREE32(den, n, BIT(mask, k), (p[2 * k] - x0) * (p[2 * k] - x0));
    TREE32(num, n, BIT(mask, k), (p[2 * k] - x0) * (p[2 * k + 1] - y0));
    const double coef = den != 0 ? num / den : 0.0;
    fit[0] = coef; fit[1] = y0 - x0 * coef; fit[2] = score_best; fit[3] = xmin; fit[4] = xmax;
    fit[5] = x0; fit[6] = y0; fit[7] = 0.0;
    *best_trial = best;
    return n_best;
}

/* Batched driver, argument meaning of lslam_ransac_ols (include/lslam.h).  xy [B,N,2]; windows of
 * `window` points, C = ceil(N/window) slots per scan; pairs nullable [B,C,trials,2]. */
int oracle_ransac_ols(const double *xy, const int32_t *npts, int32_t B, int32_t N, int32_t window, int32_t trials,
                      double thr, double probability, uint64_t seed, uint64_t scan_base, const int32_t *pairs,
                      double *fit, int32_t *ninl, int32_t *ntrials, int32_t *best_trial, uint32_t *inl_mask,
                      int32_t nthreads)
{
    const int C = (N + window - 1) / window, W = (window + 31) / 32;
    if (W > 64) return -1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads)
    for (int64_t bc = 0; bc < (int64_t)B * C; ++bc) {
        const int b = (int)(bc / C), c = (int)(bc % C);
        const int nb = npts ? npts[b] : N;
        const int s = c * window, e = s + window < nb ? s + window : nb, n = e > s ? e - s : 0;
        uint32_t *mk = inl_mask + (size_t)bc * W, cur[64];
        memset(mk, 0, sizeof(uint32_t) * (size_t)W);
        if (n < 2) {                 /* no window (0) or sklearn's "min_samples may not be larger" ValueError (-2) */
            for (int k = 0; k < 8; ++k) fit[bc * 8 + k] = NAN;
            ninl[bc] = n == 0 ? 0 : -2; ntrials[bc] = 0; best_trial[bc] = -1;
            continue;
        }
        ninl[bc] = ols_window(xy + ((size_t)b * N + s) * 2, n, trials, thr, probability, seed, scan_base + (uint64_t)b,
                              (uint32_t)c, pairs ? pairs + (size_t)bc * trials * 2 : NULL, fit + bc * 8, ntrials + bc,
                              best_trial + bc, mk, cur);
    }
    return 0;
}
