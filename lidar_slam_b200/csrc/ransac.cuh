// ransac.cuh - K1/K2: batched 2-point RANSAC line extraction + inlier refit (one launch).
//
// Replaces, per chunk, /root/reference/ransac_functions.py:23-30 (skimage.measure.ransac with
// LineModelND, then slope/intercept/tip).  Semantics = oracle/ransac_line.py.
//
// Work decomposition (one CTA = G consecutive chunk slots, staged once in shared memory):
//   pass 1  one THREAD per (chunk, block of TPT trials): the thread derives its hypotheses
//           from Philox (one block per two trials), then walks the chunk's points, which every
//           lane of the warp reads from the same shared-memory address (broadcast LDS.128).
//           Per point-hypothesis: 2 DFMA (signed distance in normal form), 1 DFMA (sum of
//           squares), 1 DADD (|s| - thr), then integer-pipe work only: the sign bit is
//           funnel-shifted into a 32-point bitmask (POPC every 32 points) and the high word of
//           |s| - thr feeds an FMNMX that tracks how close any point came to the threshold.
//           No shuffles, no reductions, no divergence in the hot loop.
//   pass 2  one WARP per chunk: trials that came within the re-check band of the threshold (or
//           had coincident sample points) are re-scored with the literal formula; among the
//           trials tied on the maximal inlier count, only those whose fast residual sum could
//           still be the minimum (rigorous bound, lslam_core.cuh) are re-scored literally, in
//           trial order, with the oracle's summation tree; the first best wins, exactly like
//           the sequential accept rule.  The winner's literal inlier mask then drives the
//           total-least-squares refit (principal axis of the 2x2 scatter), slope, intercept
//           and tip.
// The early exit of the sequential loop (sum of squares <= 0) never changes the winner, so all
// trials are evaluated.
#pragma once
#include "lslam_core.cuh"

#ifdef LSLAM_EMU
#define LSLAM_SMEM_DECL unsigned char *smem_raw = emu::block_smem()
#else
#define LSLAM_SMEM_DECL extern __shared__ __align__(16) unsigned char smem_raw[]
#endif

namespace lslam {

struct RansacArgs {
    const double2 *xy;
    const int32_t *npts;
    int B, N, chunk, C, W, trials;
    double thr;
    uint64_t seed, scan_base;
    double *line;
    int32_t *ninl, *win_trial, *win_pair;
    double *ssum;
    uint32_t *mask;
    unsigned long long *counters;   // [3]: literal re-scores, extra tie-break candidates, faults
    int G;                          // chunk slots per CTA
    int ipc;                        // work items (threads) per chunk = ceil(trials / TPT)
    long long nslots;               // B * C
};

struct SlotInfo { int off, n, scale_hi, valid; };

constexpr int kCntFlagged = -1;
constexpr int kCntExact = 0x40000000;
constexpr int kStatusNoModel = -3;
constexpr int kStatusNonFinite = -5;

__host__ __device__ inline size_t ransac_smem_bytes(int G, int chunk, int T, int W, int nwarps)
{
    size_t b = (size_t)G * chunk * 16;           // points
    b += (size_t)G * T * 8;                      // fast residual sums
    b += (((size_t)G * T * 4) + 15) & ~(size_t)15; // counts
    b += (size_t)G * sizeof(SlotInfo);
    b += (size_t)nwarps * 2 * W * 4;             // two mask buffers per warp
    return b;
}

// literal score of one hypothesis by a whole warp: inlier count, tree32 sum of r*r, and
// (optionally) the inlier bitmask words written to shared memory.
__device__ __forceinline__ void warp_literal_score(const double2 *P, int n, const LitLine m, double thr,
                                                   int lane, int &cnt, double &ssum, uint32_t *mask_out)
{
    double part = 0.0;
    int c = 0;
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int k = k0 + lane;
        bool in = false;
        if (k < n) {
            const double r = lit_residual(m, P[k]);
            in = r < thr;
            part = __dadd_rn(part, __dmul_rn(r, r));
        }
        const unsigned b = __ballot_sync(LSLAM_FULL_MASK, in);
        c += __popc(b);
        if (mask_out != nullptr && lane == 0) mask_out[k0 >> 5] = b;
    }
    cnt = c;
    ssum = warp_tree32(part);
}

__device__ __forceinline__ void write_empty_slot(const RansacArgs &a, long long s, int status, int lane)
{
    if (lane == 0) {
        const double qnan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
        for (int k = 0; k < 8; ++k) a.line[s * 8 + k] = qnan;
        a.ninl[s] = status;
        a.win_trial[s] = -1;
        a.win_pair[2 * s] = -1;
        a.win_pair[2 * s + 1] = -1;
        a.ssum[s] = qnan;
    }
    for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = 0u;
}

template <int TPT, int MAXT>
__global__ void __launch_bounds__(MAXT) ransac_lines_kernel(const RansacArgs a)
{
    LSLAM_SMEM_DECL;
    const int T = a.trials, G = a.G, chunk = a.chunk;
    double2 *pts = reinterpret_cast<double2 *>(smem_raw);
    double *ssf_s = reinterpret_cast<double *>(pts + (size_t)G * chunk);
    int *cnt_s = reinterpret_cast<int *>(ssf_s + (size_t)G * T);
    SlotInfo *info = reinterpret_cast<SlotInfo *>(
        reinterpret_cast<unsigned char *>(cnt_s) + ((((size_t)G * T * 4) + 15) & ~(size_t)15));
    uint32_t *wmask = reinterpret_cast<uint32_t *>(info + G);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const long long slot0 = (long long)blockIdx.x * G;
    const long long rem_slots = a.nslots - slot0;
    const int nsl = rem_slots < (long long)G ? (int)rem_slots : G;
    const int b0 = (int)(slot0 / a.C), c0 = (int)(slot0 % a.C);
    const long long gp0 = (long long)b0 * a.N + (long long)c0 * chunk;   // first staged point

    // ---- stage the CTA's points (contiguous in global memory), 16 B per thread, coalesced ----
    {
        const long long sl = slot0 + nsl - 1;
        const int bl = (int)(sl / a.C), cl = (int)(sl % a.C);
        const int cend = (cl + 1) * chunk < a.N ? (cl + 1) * chunk : a.N;
        const int nstage = (int)((long long)bl * a.N + cend - gp0);
        const double2 *src = a.xy + gp0;
        for (int k = tid; k < nstage; k += blockDim.x) pts[k] = __ldg(src + k);
    }
    __syncthreads();
    // ---- per-slot geometry and coordinate scale (one warp per slot) ----
    for (int q = warp; q < nsl; q += nwarps) {
        const long long s = slot0 + q;
        const int b = (int)(s / a.C), c = (int)(s % a.C);
        int nb = a.npts ? a.npts[b] : a.N;
        nb = nb < 0 ? 0 : (nb > a.N ? a.N : nb);
        int n = nb - c * chunk;
        n = n < 0 ? 0 : (n > chunk ? chunk : n);
        const int off = (int)((long long)b * a.N + (long long)c * chunk - gp0);
        int mh = 0;
        for (int k = lane; k < n; k += 32) {
            const double2 p = pts[off + k];
            const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
            mh = mh > hx ? mh : hx;
            mh = mh > hy ? mh : hy;
        }
        mh = __reduce_max_sync(LSLAM_FULL_MASK, mh);
        if (lane == 0) {
            SlotInfo si;
            si.off = off; si.n = n; si.scale_hi = mh;
            // functions.py:64-76: full chunks, or a remainder of MORE than 2 points
            si.valid = (n >= 2) && (n == chunk || n > 2);
            info[q] = si;
        }
    }
    __syncthreads();

    // ================= pass 1: thread per (chunk, TPT trials), loop over points =================
    const int items = nsl * a.ipc;
    for (int item = tid; item < items; item += blockDim.x) {
        const int q = item / a.ipc, tb = item - q * a.ipc;
        const SlotInfo si = info[q];
        const ChunkBand cb = chunk_band(si.scale_hi);
        const bool skip = !si.valid || cb.bad;
        const double2 *P = pts + si.off;
        const long long s = slot0 + q;
        const uint64_t scan = a.scan_base + (uint64_t)(s / a.C);
        const uint32_t c = (uint32_t)(s % a.C);
        const int t0 = tb * TPT;
        const uint32_t n = (uint32_t)si.n;

        double nx[TPT], ny[TPT], ng[TPT], ss[TPT];
        unsigned bits[TPT];
        int cnt[TPT];
        float mab[TPT];
        bool deg[TPT];
        Philox4 ph;
#pragma unroll
        for (int k = 0; k < TPT; ++k) {
            const int t = t0 + k;
            if ((k & 1) == 0 || TPT == 1)
                ph = philox4x32_10((uint32_t)t >> 1, c, (uint32_t)scan, (uint32_t)(scan >> 32),
                                   (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
            nx[k] = 0.0; ny[k] = 0.0; ng[k] = 0.0; ss[k] = 0.0;
            bits[k] = 0u; cnt[k] = 0; mab[k] = __int_as_float(0x7f800000); deg[k] = false;
            if (!skip && t < T) {
                int i, j;
                const bool odd = t & 1;
                pair_from_words(odd ? ph.r[2] : ph.r[0], odd ? ph.r[3] : ph.r[1], n, i, j);
                const FastLine f = fast_pair_model(P[i], P[j]);
                nx[k] = f.nx; ny[k] = f.ny; ng[k] = f.negc; deg[k] = f.degenerate;
            }
        }

#define LSLAM_PHE_STEP(PT)                                                       \
    {                                                                            \
        const double2 p_ = (PT);                                                 \
        _Pragma("unroll") for (int k_ = 0; k_ < TPT; ++k_) {                     \
            const double s_ = fma(p_.y, ny[k_], fma(p_.x, nx[k_], ng[k_]));      \
            ss[k_] = fma(s_, s_, ss[k_]);                                        \
            const int hi_ = __double2hiint(fabs(s_) - a.thr);                    \
            bits[k_] = __funnelshift_l((unsigned)hi_, bits[k_], 1);              \
            mab[k_] = fminf(mab[k_], fabsf(__int_as_float(hi_)));                \
        }                                                                        \
    }
        const int n_eff = skip ? 0 : (int)n;
        int k = 0;
        for (; k + 32 <= n_eff; k += 32) {
#pragma unroll 8
            for (int kk = 0; kk < 32; ++kk) LSLAM_PHE_STEP(P[k + kk]);
#pragma unroll
            for (int k_ = 0; k_ < TPT; ++k_) cnt[k_] += __popc(bits[k_]);
        }
        const int rem = n_eff - k;
        if (rem > 0) {
            for (int kk = 0; kk < rem; ++kk) LSLAM_PHE_STEP(P[k + kk]);
#pragma unroll
            for (int k_ = 0; k_ < TPT; ++k_) cnt[k_] += __popc(bits[k_] & ((1u << rem) - 1u));
        }
#undef LSLAM_PHE_STEP
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) {
            const int t = t0 + k_;
            if (t < T) {
                const bool flag = deg[k_] || cb.force_literal || __float_as_int(mab[k_]) < cb.band_hi;
                cnt_s[q * T + t] = skip ? 0 : (flag ? kCntFlagged : cnt[k_]);
                ssf_s[q * T + t] = ss[k_];
            }
        }
    }
    __syncthreads();

    // ================= pass 2: warp per chunk - exact selection, refit, outputs =================
    uint32_t *mcur = wmask + (size_t)warp * 2 * a.W;
    uint32_t *mbest = mcur + a.W;
    for (int q = warp; q < nsl; q += nwarps) {
        const SlotInfo si = info[q];
        const long long s = slot0 + q;
        const ChunkBand cb = chunk_band(si.scale_hi);
        if (!si.valid) { write_empty_slot(a, s, 0, lane); continue; }
        if (cb.bad) { write_empty_slot(a, s, kStatusNonFinite, lane); continue; }
        const int n = si.n;
        const double2 *P = pts + si.off;
        const uint64_t scan = a.scan_base + (uint64_t)(s / a.C);
        const uint32_t c = (uint32_t)(s % a.C);
        int *cq = cnt_s + q * T;
        double *sq = ssf_s + q * T;
        unsigned n_relit = 0, n_cand = 0, n_fault = 0;

        int mx = 0, best_t = -1, best_i = -1, best_j = -1;
        double best_ss = 0.0;
        for (int attempt = 0; attempt < 2; ++attempt) {
            // (1) literal re-score of flagged trials
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                unsigned m = __ballot_sync(LSLAM_FULL_MASK, t < T && cq[t] == kCntFlagged);
                while (m) {
                    const int tt = t0 + __ffs((int)m) - 1;
                    m &= m - 1;
                    int i, j, cl;
                    double sl;
                    sample_pair(a.seed, scan, c, (uint32_t)tt, (uint32_t)n, i, j);
                    warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, nullptr);
                    if (lane == 0) { cq[tt] = cl | kCntExact; sq[tt] = sl; }
                    ++n_relit;
                }
            }
            __syncwarp();
            // (2) maximal inlier count
            mx = 0;
            for (int t = lane; t < T; t += 32) { const int v = cq[t] & ~kCntExact; mx = mx > v ? mx : v; }
            mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
            if (mx < 2) break;     // upstream's final estimate() needs 2 points
            // (3) smallest upper bound of the literal residual sum among the tied trials
            double minhi = __longlong_as_double(0x7ff0000000000000LL);
            for (int t = lane; t < T; t += 32) {
                const int cv = cq[t];
                if ((cv & ~kCntExact) == mx) {
                    const double v = sq[t];
                    const double e = (cv & kCntExact) ? 0.0 : ssum_bound(v, cb.eabs, n);
                    minhi = fmin(minhi, v + e);
                }
            }
            minhi = warp_min(minhi);
            // (4) literal evaluation of the trials that can still be the minimum, in trial order
            best_t = -1;
            best_ss = __longlong_as_double(0x7ff0000000000000LL);
            bool fault = false;
            int first_tied = -1;
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                bool tied = false, cand = false;
                if (t < T) {
                    const int cv = cq[t];
                    if ((cv & ~kCntExact) == mx) {
                        tied = true;
                        const double v = sq[t];
                        const double e = (cv & kCntExact) ? 0.0 : ssum_bound(v, cb.eabs, n);
                        cand = (v - e) <= minhi;
                    }
                }
                const unsigned mt = __ballot_sync(LSLAM_FULL_MASK, tied);
                if (first_tied < 0 && mt) first_tied = t0 + __ffs((int)mt) - 1;
                unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
                while (m) {
                    const int tt = t0 + __ffs((int)m) - 1;
                    m &= m - 1;
                    int i, j, cl;
                    double sl;
                    sample_pair(a.seed, scan, c, (uint32_t)tt, (uint32_t)n, i, j);
                    warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, mcur);
                    ++n_cand;
                    if (cl != mx) fault = true;
                    if (sl < best_ss) {
                        best_ss = sl; best_t = tt; best_i = i; best_j = j;
                        uint32_t *tmp = mcur; mcur = mbest; mbest = tmp;
                    }
                }
            }
            if (!fault && best_t < 0 && first_tied >= 0) {
                // every candidate had a non-finite sum: the sequential rule keeps the first trial
                // that reached the maximal count
                int cl;
                sample_pair(a.seed, scan, c, (uint32_t)first_tied, (uint32_t)n, best_i, best_j);
                warp_literal_score(P, n, lit_pair_model(P[best_i], P[best_j]), a.thr, lane, cl, best_ss, mbest);
                best_t = first_tied;
                if (cl != mx) fault = true;
            }
            if (!fault) break;
            // A fast count disagreed with the literal one: the error bound was violated (never
            // observed; counted in counters[2]).  Re-score EVERY trial literally and redo.
            ++n_fault;
            __syncwarp();
            for (int t = lane; t < T; t += 32) cq[t] = kCntFlagged;
            __syncwarp();
        }
        if (lane == 0) {
            if (n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);
            if (n_cand > 1) atomicAdd(a.counters + 1, (unsigned long long)(n_cand - 1));
            if (n_fault) atomicAdd(a.counters + 2, (unsigned long long)n_fault);
        }
        if (mx < 2 || best_t < 0) { write_empty_slot(a, s, kStatusNoModel, lane); continue; }
        __syncwarp();

        // (5) refit on the winner's inliers: LineModelND.estimate(data[inliers])
        double ox, oy, dx, dy;
        if (mx == 2) {
            int i0 = -1, i1 = -1;
            if (lane == 0) {
                for (int w = 0; w * 32 < n; ++w) {
                    unsigned mw = mbest[w];
                    while (mw) {
                        const int k = w * 32 + __ffs((int)mw) - 1;
                        mw &= mw - 1;
                        if (i0 < 0) i0 = k; else if (i1 < 0) i1 = k;
                    }
                }
            }
            i0 = __shfl_sync(LSLAM_FULL_MASK, i0, 0);
            i1 = __shfl_sync(LSLAM_FULL_MASK, i1, 0);
            const LitLine m = lit_pair_model(P[i0], P[i1]);
            ox = m.ox; oy = m.oy; dx = m.dx; dy = m.dy;
            if (dx < 0.0 || (dx == 0.0 && dy < 0.0)) { dx = -dx; dy = -dy; }
        } else {
            double sx = 0.0, sy = 0.0;
            for (int k = lane; k < n; k += 32)
                if ((mbest[k >> 5] >> (k & 31)) & 1u) { const double2 p = P[k]; sx += p.x; sy += p.y; }
            sx = warp_sum(sx); sy = warp_sum(sy);
            ox = sx / (double)mx; oy = sy / (double)mx;
            double sxx = 0.0, sxy = 0.0, syy = 0.0;
            for (int k = lane; k < n; k += 32)
                if ((mbest[k >> 5] >> (k & 31)) & 1u) {
                    const double2 p = P[k];
                    const double cx = p.x - ox, cy = p.y - oy;
                    sxx = fma(cx, cx, sxx); sxy = fma(cx, cy, sxy); syy = fma(cy, cy, syy);
                }
            sxx = warp_sum(sxx); sxy = warp_sum(sxy); syy = warp_sum(syy);
            principal_axis(sxx, sxy, syy, dx, dy);
        }
        int last = -1;
        for (int w = lane; w * 32 < n; w += 32) {
            const unsigned mw = mbest[w];
            if (mw) { const int k = w * 32 + 31 - __clz((int)mw); last = last > k ? last : k; }
        }
        last = __reduce_max_sync(LSLAM_FULL_MASK, last);
        if (lane == 0) {
            // ransac_functions.py:26-30 (no guard for dir_x == 0, like the reference)
            const double sl = __ddiv_rn(dy, dx);
            const double ic = __dsub_rn(oy, __dmul_rn(sl, ox));
            const double tipx = P[last].x;
            const double tipy = __dadd_rn(__dmul_rn(tipx, sl), ic);
            double *L = a.line + s * 8;
            L[0] = sl; L[1] = ic; L[2] = ox; L[3] = oy; L[4] = dx; L[5] = dy; L[6] = tipx; L[7] = tipy;
            a.ninl[s] = mx;
            a.win_trial[s] = best_t;
            a.win_pair[2 * s] = best_i;
            a.win_pair[2 * s + 1] = best_j;
            a.ssum[s] = best_ss;
        }
        for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = (w * 32 < n) ? mbest[w] : 0u;
        __syncwarp();
    }
}

} // namespace lslam
