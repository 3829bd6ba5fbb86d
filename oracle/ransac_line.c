/* C restatement of the RANSAC line / landmark extraction path (oracle side).
 *
 * TEST INFRASTRUCTURE (see oracle/__init__.py): checker for the CUDA path and the timed
 * CPU baseline of bench.py.  Never linked into, or called from, the product library.
 * PARITY UNPINNED (scikit-image is absent and un-pinned): this follows the published
 * algorithm of skimage.measure.ransac + LineModelND exactly as oracle/ransac_line.py
 * states it (operation order in that file's header), driven the way
 * /root/reference/ransac_functions.py:20-31 drives it, one call per chunk cut as
 * /root/reference/functions.py:64-76 cuts a revolution.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC   (oracle/Makefile)
 * -ffp-contract=off matters: every product and sum below must round on its own.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- Philox4x32-10 (Salmon et al. SC'11), same contract as oracle/philox.py ---- */
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 10; ++r) {
        if (r) { k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    }
}

static inline void sample_pair(uint64_t seed, uint64_t scan, uint32_t chunk, uint32_t trial,
                               uint32_t n, int32_t *i, int32_t *j)
{
    uint32_t c[4] = { trial >> 1, chunk, (uint32_t)scan, (uint32_t)(scan >> 32) };
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint32_t u = (trial & 1u) ? c[2] : c[0], v = (trial & 1u) ? c[3] : c[1];
    uint32_t a = (uint32_t)(((uint64_t)u * n) >> 32);
    uint32_t b = (uint32_t)(((uint64_t)v * (n - 1)) >> 32);
    if (b >= a) ++b;
    *i = (int32_t)a; *j = (int32_t)b;
}

void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c[4] = { ctr[0], ctr[1], ctr[2], ctr[3] };
    philox4x32_10(c, key[0], key[1]);
    memcpy(out, c, sizeof c);
}

/* ---- LineModelND, 2-point estimate and residual (oracle/ransac_line.py header) ---- */
static inline void pair_model(const double *p, int i, int j, double m[4])
{
    double xi = p[2 * i], yi = p[2 * i + 1], xj = p[2 * j], yj = p[2 * j + 1];
    double ox = (xi + xj) / 2.0, oy = (yi + yj) / 2.0;
    double cix = xi - ox, ciy = yi - oy, cjx = xj - ox, cjy = yj - oy;
    double dx = cjx - cix, dy = cjy - ciy;
    double n = sqrt(dx * dx + dy * dy);
    if (n != 0) { dx = dx / n; dy = dy / n; }
    m[0] = ox; m[1] = oy; m[2] = dx; m[3] = dy;
}

static inline double residual(const double m[4], double x, double y)
{
    double vx = x - m[0], vy = y - m[1];
    double t = vx * m[2] + vy * m[3];
    double rx = vx - t * m[2], ry = vy - t * m[3];
    return sqrt(rx * rx + ry * ry);
}

/* count of r < thr, tree32 sum of r*r, optional inlier bitmask */
static int score(const double *p, int n, const double m[4], double thr, double *ssum,
                 uint32_t *mask)
{
    double lane[32];
    int cnt = 0;
    for (int l = 0; l < 32; ++l) lane[l] = 0.0;
    if (mask) memset(mask, 0, sizeof(uint32_t) * (size_t)((n + 31) / 32));
    for (int k = 0; k < n; ++k) {
        double r = fabs(residual(m, p[2 * k], p[2 * k + 1]));
        if (r < thr) { ++cnt; if (mask) mask[k >> 5] |= 1u << (k & 31); }
        lane[k & 31] = lane[k & 31] + r * r;
    }
    for (int h = 16; h >= 1; h >>= 1)
        for (int l = 0; l < h; ++l) lane[l] = lane[l] + lane[l + h];
    *ssum = lane[0];
    return cnt;
}

/* LineModelND.estimate on the inliers: 2 points -> pair model in data order; more ->
 * principal axis of the centred points.  Upstream takes it from np.linalg.svd; here the
 * closed form of the 2x2 scatter matrix (same axis, sign normalised to d_x > 0). */
static void refit(const double *p, int n, const uint32_t *mask, int cnt, double m[4])
{
    if (cnt == 2) {
        int idx[2], c = 0;
        for (int k = 0; k < n && c < 2; ++k) if (mask[k >> 5] >> (k & 31) & 1u) idx[c++] = k;
        pair_model(p, idx[0], idx[1], m);
    } else {
        double sx = 0.0, sy = 0.0;
        for (int k = 0; k < n; ++k) if (mask[k >> 5] >> (k & 31) & 1u) { sx += p[2 * k]; sy += p[2 * k + 1]; }
        double ox = sx / cnt, oy = sy / cnt;
        double sxx = 0.0, sxy = 0.0, syy = 0.0;
        for (int k = 0; k < n; ++k) if (mask[k >> 5] >> (k & 31) & 1u) {
            double cx = p[2 * k] - ox, cy = p[2 * k + 1] - oy;
            sxx += cx * cx; sxy += cx * cy; syy += cy * cy;
        }
        double del = 0.5 * (sxx - syy), h = hypot(del, sxy), dx, dy;
        if (h == 0) { dx = 1.0; dy = 0.0; }
        else if (del >= 0) { dx = del + h; dy = sxy; }
        else { dx = sxy; dy = h - del; }
        double nn = hypot(dx, dy);
        if (nn != 0) { dx /= nn; dy /= nn; } else { dx = 1.0; dy = 0.0; }
        m[0] = ox; m[1] = oy; m[2] = dx; m[3] = dy;
    }
    if (m[2] < 0 || (m[2] == 0 && m[3] < 0)) { m[2] = -m[2]; m[3] = -m[3]; }
}

/* One chunk: the ransac loop + refit + slope/intercept/tip (ransac_functions.py:23-30).
 * Returns inlier count (>= 2) or -3 (LSLAM_ERR_NO_MODEL).  line = a,b,ox,oy,dx,dy,tipx,tipy */
/* pairs: NULL = the Philox sampler, else [trials][2] explicit sample indices (clamped into the chunk), e.g. what a
 * seeded upstream run's random_state.choice(n, 2, replace=False) drew. */
int oracle_ransac_chunk_pairs(const double *p, int n, int trials, double thr, uint64_t seed,
                              uint64_t scan, uint32_t chunk, const int32_t *pairs, double line[8],
                              int32_t *win_trial, int32_t win_pair[2], double *ssum_out, uint32_t *mask);
int oracle_ransac_chunk(const double *p, int n, int trials, double thr, uint64_t seed,
                        uint64_t scan, uint32_t chunk, double line[8], int32_t *win_trial,
                        int32_t win_pair[2], double *ssum_out, uint32_t *mask)
{
    return oracle_ransac_chunk_pairs(p, n, trials, thr, seed, scan, chunk, NULL, line, win_trial, win_pair, ssum_out, mask);
}

int oracle_ransac_chunk_pairs(const double *p, int n, int trials, double thr, uint64_t seed,
                              uint64_t scan, uint32_t chunk, const int32_t *pairs, double line[8],
                              int32_t *win_trial, int32_t win_pair[2], double *ssum_out, uint32_t *mask)
{
    int best_n = 0, win = -1, wi = -1, wj = -1;
    double best_ssum = INFINITY, bm[4] = { 0, 0, 0, 0 };
    for (int t = 0; t < trials; ++t) {
        int32_t i, j;
        double m[4], ssum;
        if (pairs) {
            i = pairs[2 * t]; j = pairs[2 * t + 1];
            i = i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
            j = j < 0 ? 0 : (j > n - 1 ? n - 1 : j);
        } else {
            sample_pair(seed, scan, chunk, (uint32_t)t, (uint32_t)n, &i, &j);
        }
        pair_model(p, i, j, m);
        int cnt = score(p, n, m, thr, &ssum, NULL);
        if (cnt > best_n || (cnt == best_n && ssum < best_ssum)) {
            best_n = cnt; best_ssum = ssum; win = t; wi = i; wj = j;
            memcpy(bm, m, sizeof bm);
            if (best_ssum <= 0) break;
        }
    }
    *win_trial = win; win_pair[0] = wi; win_pair[1] = wj; *ssum_out = best_ssum;
    /* fewer than 2 inliers: upstream's final LineModelND.estimate raises ValueError (or ransac
     * returned (None, None) and ransac_functions.py:25 raises) -> LSLAM_ERR_NO_MODEL */
    if (win < 0 || best_n < 2) {
        for (int k = 0; k < 8; ++k) line[k] = NAN;
        *win_trial = -1; win_pair[0] = win_pair[1] = -1; *ssum_out = NAN;
        return -3;
    }
    double s2, m[4];
    score(p, n, bm, thr, &s2, mask);
    refit(p, n, mask, best_n, m);
    double a = m[3] / m[2];
    double b = m[1] - a * m[0];
    int last = -1;
    for (int k = n - 1; k >= 0; --k) if (mask[k >> 5] >> (k & 31) & 1u) { last = k; break; }
    double tipx = p[2 * last];
    line[0] = a; line[1] = b; line[2] = m[0]; line[3] = m[1]; line[4] = m[2]; line[5] = m[3];
    line[6] = tipx; line[7] = tipx * a + b;
    return best_n;
}

/* Batched driver with the same argument meaning as lslam_ransac_lines (include/lslam.h).
 * xy[B,N,2]; npts[B] or NULL (all N); C = ceil(N/chunk) chunk slots per scan; a slot is
 * processed when it is a full chunk or a remainder of more than 2 points
 * (functions.py:71), otherwise ninl = 0 / win_trial = -1. */
int oracle_ransac_lines_pairs(const double *xy, const int32_t *npts, int32_t B, int32_t N,
                              int32_t chunk, int32_t trials, double thr, uint64_t seed,
                              uint64_t scan_base, const int32_t *pairs, double *line, int32_t *ninl,
                              int32_t *win_trial, int32_t *win_pair, double *ssum, uint32_t *inl_mask,
                              int32_t nthreads);
int oracle_ransac_lines(const double *xy, const int32_t *npts, int32_t B, int32_t N,
                        int32_t chunk, int32_t trials, double thr, uint64_t seed,
                        uint64_t scan_base, double *line, int32_t *ninl, int32_t *win_trial,
                        int32_t *win_pair, double *ssum, uint32_t *inl_mask, int32_t nthreads)
{
    return oracle_ransac_lines_pairs(xy, npts, B, N, chunk, trials, thr, seed, scan_base, NULL, line, ninl, win_trial,
                                     win_pair, ssum, inl_mask, nthreads);
}

/* pairs: NULL or [B,C,trials,2] (include/lslam.h lslam_ransac_lines) */
int oracle_ransac_lines_pairs(const double *xy, const int32_t *npts, int32_t B, int32_t N,
                              int32_t chunk, int32_t trials, double thr, uint64_t seed,
                              uint64_t scan_base, const int32_t *pairs, double *line, int32_t *ninl,
                              int32_t *win_trial, int32_t *win_pair, double *ssum, uint32_t *inl_mask,
                              int32_t nthreads)
{
    int C = (N + chunk - 1) / chunk, W = (chunk + 31) / 32;
    if (nthreads <= 0) {
#ifdef _OPENMP
        nthreads = omp_get_max_threads();
#else
        nthreads = 1;
#endif
    }
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads)
    for (int64_t bc = 0; bc < (int64_t)B * C; ++bc) {
        int b = (int)(bc / C), c = (int)(bc % C);
        int nb = npts ? npts[b] : N;
        int s = c * chunk, e = s + chunk < nb ? s + chunk : nb, n = e - s;
        uint32_t *mk = inl_mask + (size_t)bc * W;
        memset(mk, 0, sizeof(uint32_t) * (size_t)W);
        if (n < chunk && n <= 2) {
            for (int k = 0; k < 8; ++k) line[bc * 8 + k] = NAN;
            ninl[bc] = 0; win_trial[bc] = -1; win_pair[2 * bc] = win_pair[2 * bc + 1] = -1;
            ssum[bc] = NAN;
            continue;
        }
        ninl[bc] = oracle_ransac_chunk_pairs(xy + ((size_t)b * N + s) * 2, n, trials, thr, seed,
                                             scan_base + (uint64_t)b, (uint32_t)c,
                                             pairs ? pairs + (size_t)bc * trials * 2 : NULL, line + bc * 8,
                                             win_trial + bc, win_pair + 2 * bc, ssum + bc, mk);
    }
    return 0;
}

/* omp_set_num_threads for every later oracle_* call with nthreads <= 0 (torchrun exports OMP_NUM_THREADS=1) */
void oracle_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
