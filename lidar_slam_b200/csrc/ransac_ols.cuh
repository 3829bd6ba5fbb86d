// ransac_ols.cuh - OLS-flavour RANSAC.
//
// Replaces, per window of PNT_NBR points, what /root/reference/line_detection.py:57-58 asks of
// scikit-learn: RANSACRegressor(residual_threshold=3).fit(x, y) - LinearRegression base estimator
// (y on x), min_samples 2, loss |y - y_hat|, inlier iff loss <= thr (non-strict), candidates ranked by
// inlier count then by R^2 on their inliers, max_trials shrinking with the best inlier ratio
// (stop_probability 0.99), final least-squares refit on the accepted inliers.  The operation order is
// that of oracle/ransac_ols.py / oracle/ransac_ols.c, which are pinned to the real scikit-learn.
//
// One warp per window.  Upstream's trial loop is sequential and its length depends on the data (the dynamic bound
// usually ends it after 2 to ~50 trials on wall-like windows).  The kernel keeps that rule bit for bit but splits a trial
// in two: the inlier COUNT of 32 consecutive trials is computed speculatively, one trial per lane (the lane walks the
// points itself); the accept rule is then replayed in trial order over the 32 counts, and only trials that reach the best
// count so far are evaluated by the whole warp (lane l owns points l, l+32, ...: mask as ballots, R^2 sums as the fixed
// 32-lane tree shared with the oracle).  Round 1 ran every trial warp-wide (every lane repeating the sampler and the model):
// 6 x the instructions.  Points are staged once per window in shared memory (x[] and y[] planes).
// HBM traffic per window: 16 B per point in, 64 + 12 + 4*W bytes out.
#pragma once
#include "lslam_core.cuh"
#include "ransac.cuh"

namespace lslam {

constexpr int kStatusTooFew = -2;

struct OlsArgs {
    const double2 *xy;        // [B, N]
    const int32_t *npts;      // [B] or null
    const int32_t *pairs;     // [B, C, trials, 2] or null (Philox sampler)
    int B, N, window, C, W, trials;
    double thr, nom;          // nom = max(eps, 1 - stop_probability)
    uint64_t seed, scan_base;
    double *fit;              // [B, C, 8]
    int32_t *ninl, *ntrials, *best_trial;
    uint32_t *inl_mask;       // [B, C, W]
};

__device__ __forceinline__ double ols_dynamic_max_trials(int n_inliers, int n_samples, double nom)
{
    const double ratio = (double)n_inliers / (double)n_samples;
    const double eps = 2.220446049250313e-16;
    const double denom = fmax(eps, __dsub_rn(1.0, __dmul_rn(ratio, ratio)));
    if (nom == 1.0) return 0.0;
    if (denom == 1.0) return __longlong_as_double(0x7ff0000000000000LL);
    return fabs(ceil(log(nom) / log(denom)));
}

__global__ void __launch_bounds__(256, 4) ransac_ols_kernel(OlsArgs a)
{
    LSLAM_SMEM_DECL;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    double2 *sp = reinterpret_cast<double2 *>(smem_raw) + (size_t)warp * a.window;   // the window's points (one LDS.128 per point)
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const long long total = (long long)a.B * a.C;
    for (long long bc = (long long)blockIdx.x * nwarps + warp; bc < total; bc += (long long)gridDim.x * nwarps) {
        const int b = (int)(bc / a.C), c = (int)(bc % a.C);
        const int nb = a.npts ? a.npts[b] : a.N;
        const int s = c * a.window;
        const int e = s + a.window < nb ? s + a.window : nb;
        const int n = e > s ? e - s : 0;
        uint32_t *mk = a.inl_mask + (size_t)bc * a.W;
        double *fit = a.fit + (size_t)bc * 8;
        if (n < 2) {
            for (int w = lane; w < a.W; w += 32) mk[w] = 0u;
            if (lane < 8) fit[lane] = nan;
            if (lane == 0) { a.ninl[bc] = n == 0 ? 0 : kStatusTooFew; a.ntrials[bc] = 0; a.best_trial[bc] = -1; }
            continue;
        }
        __syncwarp();
        const double2 *src = a.xy + (size_t)b * a.N + s;
        double xmin = inf, xmax = -inf;
        bool anynan = false;
        for (int k = lane; k < n; k += 32) {
            const double2 p = src[k];
            sp[k] = p;
            anynan |= (p.x != p.x);
            xmin = fmin(xmin, p.x); xmax = fmax(xmax, p.x);
        }
        __syncwarp();
        const int32_t *pr = a.pairs ? a.pairs + (size_t)bc * a.trials * 2 : nullptr;
        // lane w keeps word w of the current and of the best inlier mask (W <= 32)
        uint32_t cur = 0u, best_mask = 0u;
        int n_best = 1, best = -1;
        bool have = false;
        double score_best = -inf, limit = (double)a.trials;
        // 32 trials at a time, one per lane: sample, 2-point model and inlier COUNT of trial c0 + lane, the lane walking the
        // window's points itself (shared-memory broadcasts).  Then upstream's sequential accept rule is replayed over the 32
        // counts in trial order; only a trial that reaches the best count so far - a handful per window - is looked at by the
        // whole warp (its mask as ballots, R^2 on its inliers), exactly as upstream would.  A trial upstream's loop would not
        // have run (index >= the shrinking max_trials) is never looked at: computing it was speculation, its result is dropped.
        for (int c0 = 0; (double)c0 < limit; c0 += 32) {
            const int t = c0 + lane;
            const bool act = t < a.trials;
            int cnt_t = -1;
            double coef_t = 0.0, icpt_t = 0.0;
            if (act) {
                int i, j;
                if (pr) { i = pr[2 * t]; j = pr[2 * t + 1]; }
                else sample_pair(a.seed, a.scan_base + (uint64_t)b, (uint32_t)c, (uint32_t)t, (uint32_t)n, i, j);
                const double2 pi = sp[i], pj = sp[j];
                const double xi = pi.x, yi = pi.y, xj = pj.x, yj = pj.y;
                const double x0 = __dadd_rn(xi, xj) / 2.0, y0 = __dadd_rn(yi, yj) / 2.0;
                const double xci = __dsub_rn(xi, x0), xcj = __dsub_rn(xj, x0), yci = __dsub_rn(yi, y0), ycj = __dsub_rn(yj, y0);
                const double den = __dadd_rn(__dmul_rn(xci, xci), __dmul_rn(xcj, xcj));
                const double num = __dadd_rn(__dmul_rn(xci, yci), __dmul_rn(xcj, ycj));
                coef_t = den != 0.0 ? num / den : 0.0;
                icpt_t = __dsub_rn(y0, __dmul_rn(x0, coef_t));
                int cc = 0;
#pragma unroll 4
                for (int k = 0; k < n; ++k) {
                    const double2 p = sp[k];
                    cc += fabs(__dsub_rn(p.y, __dadd_rn(__dmul_rn(p.x, coef_t), icpt_t))) <= a.thr ? 1 : 0;
                }
                cnt_t = cc;
            }
            int from = 0;                                         // lanes below it have been replayed
            for (;;) {
                const unsigned m = __ballot_sync(LSLAM_FULL_MASK, act && cnt_t >= n_best && lane >= from);
                if (m == 0u) break;
                const int u = __ffs((int)m) - 1;
                if (!((double)(c0 + u) < limit)) break;           // upstream's loop had ended before this trial
                from = u + 1;
                const double coef = __shfl_sync(LSLAM_FULL_MASK, coef_t, u), icpt = __shfl_sync(LSLAM_FULL_MASK, icpt_t, u);
                const int cnt = __shfl_sync(LSLAM_FULL_MASK, cnt_t, u);
                cur = 0u;
                for (int base = 0; base < n; base += 32) {
                    const int k = base + lane;
                    bool inl = false;
                    if (k < n) inl = fabs(__dsub_rn(sp[k].y, __dadd_rn(__dmul_rn(sp[k].x, coef), icpt))) <= a.thr;
                    const uint32_t word = __ballot_sync(LSLAM_FULL_MASK, inl);
                    if (lane == (base >> 5)) cur = word;
                }
                double score;
                if (cnt < 2) {
                    score = nan;
                } else {
                    double py = 0.0;
                    for (int base = 0; base < n; base += 32) {
                        const uint32_t word = __shfl_sync(LSLAM_FULL_MASK, cur, base >> 5);
                        const int k = base + lane;
                        if (k < n && ((word >> lane) & 1u)) py = __dadd_rn(py, sp[k].y);
                    }
                    const double ym = warp_tree32(py) / (double)cnt;
                    double pn = 0.0, pd = 0.0;
                    for (int base = 0; base < n; base += 32) {
                        const uint32_t word = __shfl_sync(LSLAM_FULL_MASK, cur, base >> 5);
                        const int k = base + lane;
                        if (k < n && ((word >> lane) & 1u)) {
                            const double r = __dsub_rn(sp[k].y, __dadd_rn(__dmul_rn(sp[k].x, coef), icpt));
                            const double d = __dsub_rn(sp[k].y, ym);
                            pn = __dadd_rn(pn, __dmul_rn(r, r));
                            pd = __dadd_rn(pd, __dmul_rn(d, d));
                        }
                    }
                    const double numr = warp_tree32(pn), denr = warp_tree32(pd);
                    score = (numr != 0.0 && denr != 0.0) ? __dsub_rn(1.0, numr / denr) : (numr != 0.0 ? 0.0 : 1.0);
                }
                if (cnt == n_best && score < score_best) continue;
                n_best = cnt; score_best = score; best = c0 + u; have = true;
                best_mask = cur;
                limit = fmin(limit, ols_dynamic_max_trials(n_best, n, a.nom));
            }
        }
        // trials upstream's loop runs: up to the (integer-valued) bound in force after the last accepted trial, at least that trial
        int n_trials = (int)limit;
        if (n_trials < best + 1) n_trials = best + 1;
        // np.min / np.max over the window's x (line_detection.py:66) propagate NaN
        for (int o = 16; o >= 1; o >>= 1) {
            xmin = fmin(xmin, __shfl_xor_sync(LSLAM_FULL_MASK, xmin, o));
            xmax = fmax(xmax, __shfl_xor_sync(LSLAM_FULL_MASK, xmax, o));
        }
        if (__any_sync(LSLAM_FULL_MASK, anynan)) xmin = xmax = nan;
        if (lane == 0) a.ntrials[bc] = n_trials;
        if (!have) {          // every trial skipped: upstream raises ValueError
            for (int w = lane; w < a.W; w += 32) mk[w] = 0u;
            if (lane < 8) fit[lane] = (lane == 3) ? xmin : (lane == 4 ? xmax : nan);
            if (lane == 0) { a.ninl[bc] = kStatusNoModel; a.best_trial[bc] = -1; }
            continue;
        }
        double px = 0.0, py = 0.0;
        for (int base = 0; base < n; base += 32) {
            const uint32_t word = __shfl_sync(LSLAM_FULL_MASK, best_mask, base >> 5);
            const int k = base + lane;
            if (k < n && ((word >> lane) & 1u)) { px = __dadd_rn(px, sp[k].x); py = __dadd_rn(py, sp[k].y); }
        }
        const double x0 = warp_tree32(px) / (double)n_best, y0 = warp_tree32(py) / (double)n_best;
        double pd = 0.0, pn = 0.0;
        for (int base = 0; base < n; base += 32) {
            const uint32_t word = __shfl_sync(LSLAM_FULL_MASK, best_mask, base >> 5);
            const int k = base + lane;
            if (k < n && ((word >> lane) & 1u)) {
                const double xc = __dsub_rn(sp[k].x, x0), yc = __dsub_rn(sp[k].y, y0);
                pd = __dadd_rn(pd, __dmul_rn(xc, xc));
                pn = __dadd_rn(pn, __dmul_rn(xc, yc));
            }
        }
        const double den = warp_tree32(pd), num = warp_tree32(pn);
        const double coef = den != 0.0 ? num / den : 0.0;
        const double icpt = __dsub_rn(y0, __dmul_rn(x0, coef));
        if (lane < a.W) mk[lane] = best_mask;
        if (lane == 0) {
            fit[0] = coef; fit[1] = icpt; fit[2] = score_best; fit[3] = xmin; fit[4] = xmax;
            fit[5] = x0; fit[6] = y0; fit[7] = 0.0;
            a.ninl[bc] = n_best; a.best_trial[bc] = best;
        }
    }
}

} // namespace lslam
