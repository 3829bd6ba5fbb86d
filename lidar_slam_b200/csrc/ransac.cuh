// ransac.cuh - K1/K2: batched 2-point RANSAC line extraction + inlier refit (one launch).
//
// Replaces, per chunk, /root/reference/ransac_functions.py:23-30 (skimage.measure.ransac with
// LineModelND, then slope/intercept/tip).  Semantics = oracle/ransac_line.py.
//
// One CTA owns G consecutive chunk slots, staged once in shared memory, and runs phases that
// each use the parallel axis that fits them (block barriers in between):
//   score   one THREAD per (chunk, TPT trials): hypotheses from Philox (one block per two
//           trials), then a walk over the chunk's points, read by every lane from the same
//           shared-memory address (broadcast LDS.128).  Per point-hypothesis: 2 DFMA (signed
//           distance, normal form) + 1 DFMA (sum of squares) + 1 DADD (|s| - thr); the sign bit
//           is funnel-shifted into a 32-point bitmask (POPC every 32 points) and the high word
//           of |s| - thr feeds an FMNMX3 that tracks how close any point came to the threshold.
//           No shuffles, no reductions, no divergence in the hot loop.
//   select  one WARP per chunk: trials that came within the re-check band of the threshold (or
//           had coincident samples) are re-scored with the literal formula; the maximal inlier
//           count and the trials whose fast residual sum could still be the minimum (rigorous
//           linear bound, lslam_core.cuh) become the candidate list (usually one trial).
//   models  one LANE per (chunk, candidate): Philox pair + literal 2-point model (sqrt, 2 div).
//   decide  one WARP per chunk: literal score of each candidate in trial order with the
//           oracle's summation tree; first best wins like the sequential accept rule.
//   refit   four LANES per chunk: one pass over the winner's inliers for the 2x2 scatter, then
//           the principal axis (closed form of the SVD), slope, intercept, tip and the stores.
// The early exit of the sequential loop (sum of squares <= 0) never changes the winner, so all
// trials are evaluated.
#pragma once
#include "lslam_core.cuh"
#include "async_copy.cuh"

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
    int ws_ns, ws_np, ws_nb;        // warp-specialised kernel: scorer warps, post warps, ring depth
    const int32_t *pairs;           // [B,C,trials,2] explicit sample pairs, or nullptr = the Philox sampler
    double inv_thr;                 // 1 / thr (tile kernel: keeps a division out of the per-slot set-up)
};

constexpr int kMaxCand = 8;         // candidates with precomputed models; more -> generic loop
constexpr int kCntFlagged = -1;
constexpr int kCntExact = 0x40000000;
constexpr int kStatusNoModel = -3;
constexpr int kStatusNonFinite = -5;

// the sample pair of trial t of global chunk slot gslot: the caller's explicit pair (indices clamped into the chunk),
// or the Philox sampler (lslam_core.cuh sample_pair)
__device__ __forceinline__ void explicit_pair(const RansacArgs &a, long long gslot, uint32_t t, uint32_t n, int &i, int &j)
{
    const int2 v = __ldg(reinterpret_cast<const int2 *>(a.pairs) + gslot * a.trials + t);
    const int hi = (int)n - 1;
    i = v.x < 0 ? 0 : (v.x > hi ? hi : v.x);
    j = v.y < 0 ? 0 : (v.y > hi ? hi : v.y);
}
__device__ __forceinline__ void trial_pair(const RansacArgs &a, long long gslot, uint64_t scan, uint32_t c, uint32_t t, uint32_t n,
                                           int &i, int &j)
{
    if (a.pairs != nullptr) explicit_pair(a, gslot, t, n, i, j);
    else sample_pair(a.seed, scan, c, t, n, i, j);
}

// per-slot state in shared memory
struct SlotState {
    unsigned long long scan;        // global scan index (Philox counter words 2,3)
    long long gslot;                // global chunk slot index (explicit sample pairs)
    double best_ss;                 // literal residual sum of the winner
    int off, n, valid, c;           // first point (index into the staged points), count, flag, chunk idx
    int scale_hi;                   // high word of the largest |coordinate|
    int flag;                       // some trial needs the literal re-score
    int mx, ncand, minhi_q;         // maximal inlier count, number of candidates, quantised bound
    int best_t, best_i, best_j;     // winner trial and its sample pair
    int status;                     // 0 ok, else the value written to ninl
    int cand_t[kMaxCand], cand_i[kMaxCand], cand_j[kMaxCand];
};

__host__ __device__ inline size_t align16(size_t b) { return (b + 15) & ~(size_t)15; }

__host__ __device__ inline size_t ransac_smem_bytes(int G, int chunk, int T, int W, int nwarps)
{
    size_t b = (size_t)G * chunk * 16;                       // points: double2
    b += (size_t)G * T * 8;                                  // fast residual sums
    b += align16((size_t)G * T * 4);                         // counts
    b += align16((size_t)G * sizeof(SlotState));
    b += (size_t)G * kMaxCand * sizeof(LitLine);             // candidate models
    b += align16((size_t)G * W * 4);                         // winner masks
    b += align16((size_t)nwarps * W * 4);                    // per-warp scratch mask
    b += 32;                                                 // CTA-wide scale + CTA geometry
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

// order-preserving 32-bit keys of non-negative doubles (high word), rounded down / up
// (the high word of a double <= 0 is negative as an int, or 0 for +0: an integer max instead of an fp64 compare + select;
//  a NaN maps to a huge key - every caller scores such a trial literally)
__device__ __forceinline__ int key_down(double v) { const int h = __double2hiint(v); return h > 0 ? h : 0; }
__device__ __forceinline__ int key_up(double v) { const int h = __double2hiint(v); return (h > 0 ? h : 0) + 1; }

// candidate test of trial t for the tie-break: tied on the maximal count and its lower bound
// does not exceed the smallest upper bound
__device__ __forceinline__ bool trial_is_candidate(int cv, double v, int mx, int minhi_q, const SsumBound &sb)
{
    if ((cv & ~kCntExact) != mx) return false;
    const double e = (cv & kCntExact) ? 0.0 : ssum_err(sb, v);
    return key_down(v - e) <= minhi_q;
}

// ---- CTA-level view of shared memory ---------------------------------------------------
struct RansacSmem {
    double2 *pts;       // staged points
    double *ssf;
    int *cnt;
    SlotState *slot;
    LitLine *cmodel;
    uint32_t *smask;    // [G][W] winner masks
    uint32_t *wtmp;     // [nwarps][W] scratch masks
    int *cta_scale;
    int *geo;           // b0, c0, number of staged points
};

__device__ __forceinline__ RansacSmem carve_smem(unsigned char *raw, int G, int chunk, int T, int W, int nwarps)
{
    RansacSmem s;
    unsigned char *p = raw;
    s.pts = reinterpret_cast<double2 *>(p);   p += (size_t)G * chunk * 16;
    s.ssf = reinterpret_cast<double *>(p);    p += (size_t)G * T * 8;
    s.cnt = reinterpret_cast<int *>(p);       p += align16((size_t)G * T * 4);
    s.slot = reinterpret_cast<SlotState *>(p); p += align16((size_t)G * sizeof(SlotState));
    s.cmodel = reinterpret_cast<LitLine *>(p); p += (size_t)G * kMaxCand * sizeof(LitLine);
    s.smask = reinterpret_cast<uint32_t *>(p); p += align16((size_t)G * W * 4);
    s.wtmp = reinterpret_cast<uint32_t *>(p);  p += align16((size_t)nwarps * W * 4);
    s.cta_scale = reinterpret_cast<int *>(p);
    s.geo = s.cta_scale + 4;
    return s;
}

// Exact, slow path for one chunk: literal score of EVERY trial and the sequential accept rule.
// Used only when the fast pass was caught disagreeing with the literal count (never observed;
// counted in counters[2]).  Leaves the winner in the slot state and its mask in smask.
template <class Slot>
__device__ void slot_exact_fallback(const RansacArgs &a, Slot *st, const double2 *P, uint32_t *smask_q,
                                    uint32_t *wtmp_w, int lane)
{
    const int n = st->n, T = a.trials;
    int best_n = 0, best_t = -1, best_i = -1, best_j = -1;
    double best_ss = __longlong_as_double(0x7ff0000000000000LL);
    for (int t = 0; t < T; ++t) {
        int i, j, cl;
        double sl;
        trial_pair(a, st->gslot, st->scan, (uint32_t)st->c, (uint32_t)t, (uint32_t)n, i, j);
        warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, wtmp_w);
        __syncwarp();
        if (cl > best_n || (cl == best_n && sl < best_ss)) {
            best_n = cl; best_ss = sl; best_t = t; best_i = i; best_j = j;
            for (int w = lane; w * 32 < n; w += 32) smask_q[w] = wtmp_w[w];
        }
        __syncwarp();
    }
    if (lane == 0) {
        st->mx = best_n; st->best_t = best_t; st->best_i = best_i; st->best_j = best_j; st->best_ss = best_ss;
        st->status = (best_t < 0 || best_n < 2) ? kStatusNoModel : 0;
    }
    __syncwarp();
}

template <int TPT, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) ransac_lines_kernel(const RansacArgs a)
{
    LSLAM_SMEM_DECL;
    const int T = a.trials, G = a.G, chunk = a.chunk, W = a.W;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const RansacSmem sm = carve_smem(smem_raw, G, chunk, T, W, nwarps);
    const long long slot0 = (long long)blockIdx.x * G;
    const long long rem_slots = a.nslots - slot0;
    const int nsl = rem_slots < (long long)G ? (int)rem_slots : G;

    // ---- CTA geometry: the 64-bit divisions are done by one thread and shared (a 64-bit division is
    //      ~100 instructions; doing it in every warp cost 3-4 % of the kernel) ----
    if (tid == 0) {
        *sm.cta_scale = 0;
        const int b0_ = (int)(slot0 / a.C), c0_ = (int)(slot0 - (long long)b0_ * a.C);
        const int lastc = c0_ + nsl - 1;                      // last slot, counted from scan b0_
        const int bl_ = b0_ + lastc / a.C, cl_ = lastc % a.C;
        const int cend = (cl_ + 1) * chunk < a.N ? (cl_ + 1) * chunk : a.N;
        sm.geo[0] = b0_; sm.geo[1] = c0_;
        sm.geo[2] = (int)((long long)(bl_ - b0_) * a.N + cend - (long long)c0_ * chunk);   // points to stage
    }
    __syncthreads();
    const int b0 = sm.geo[0], c0 = sm.geo[1], nstage = sm.geo[2];
    const long long gp0 = (long long)b0 * a.N + (long long)c0 * chunk;   // first staged point
#define LSLAM_SLOT_PTS(st_) (sm.pts + (st_)->off)
    // residual-sum error bound of a slot
    auto slot_bound = [&](const SlotState *st_, int n_) -> SsumBound {
        return ssum_bound_coeffs(chunk_band(st_->scale_hi).eabs, n_, a.thr);
    };

    // ---- stage the CTA's points (contiguous in global memory): 16 B per thread, coalesced;
    //      track the largest |coordinate| of the CTA on the way ----
    {
        const double2 *src = a.xy + gp0;
        int mh = 0;
        for (int k = tid; k < nstage; k += blockDim.x) {
            const double2 p = __ldg(src + k);
            sm.pts[k] = p;
            const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
            mh = mh > hx ? mh : hx;
            mh = mh > hy ? mh : hy;
        }
        mh = __reduce_max_sync(LSLAM_FULL_MASK, mh);
        if (lane == 0 && mh > 0) atomicMax(sm.cta_scale, mh);
    }
    __syncthreads();
    // ---- per-slot geometry (one lane per slot) ----
    const bool cta_scale_bad = chunk_band(*sm.cta_scale).bad;
    for (int q = tid; q < nsl; q += blockDim.x) {
        const int cq_ = c0 + q;                               // 32-bit: counted from scan b0
        const int b = b0 + cq_ / a.C, c = cq_ % a.C;
        int nb = a.npts ? a.npts[b] : a.N;
        nb = nb < 0 ? 0 : (nb > a.N ? a.N : nb);
        int n = nb - c * chunk;
        n = n < 0 ? 0 : (n > chunk ? chunk : n);
        SlotState *st = sm.slot + q;
        st->scan = a.scan_base + (uint64_t)b;
        st->gslot = slot0 + q;
        st->best_ss = 0.0;
        st->off = (int)((long long)b * a.N + (long long)c * chunk - gp0);
        st->n = n;
        // functions.py:64-76: full chunks, or a remainder of MORE than 2 points
        st->valid = (n >= 2) && (n == chunk || n > 2);
        st->c = c;
        st->scale_hi = *sm.cta_scale;
        st->flag = 0; st->mx = 0; st->ncand = 0; st->minhi_q = 0;
        st->best_t = -1; st->best_i = -1; st->best_j = -1;
        st->status = 0;
    }
    __syncthreads();
    if (cta_scale_bad) {
        // some coordinate of the CTA is non-finite (or garbage past npts): find out per slot
        for (int q = warp; q < nsl; q += nwarps) {
            SlotState *st = sm.slot + q;
            int mh = 0;
            for (int k = lane; k < st->n; k += 32) {
                const double2 p = LSLAM_SLOT_PTS(st)[k];
                const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
                mh = mh > hx ? mh : hx;
                mh = mh > hy ? mh : hy;
            }
            mh = __reduce_max_sync(LSLAM_FULL_MASK, mh);
            if (lane == 0) st->scale_hi = mh;
        }
        __syncthreads();
    }

    // ================= score: thread per (chunk, TPT trials), loop over the chunk's points =========
    const int items = nsl * a.ipc;
    for (int item = tid; item < items; item += blockDim.x) {
        const int q = item / a.ipc, tb = item - q * a.ipc;
        SlotState *st = sm.slot + q;
        const ChunkBand cb = chunk_band(st->scale_hi);
        const bool skip = !st->valid || cb.bad;
        const double2 *P = LSLAM_SLOT_PTS(st);
        const uint64_t scan = st->scan;
        const uint32_t c = (uint32_t)st->c;
        const int t0 = tb * TPT;
        const uint32_t n = (uint32_t)st->n;

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
                if (a.pairs != nullptr) explicit_pair(a, st->gslot, (uint32_t)t, n, i, j);
                else pair_from_words(odd ? ph.r[2] : ph.r[0], odd ? ph.r[3] : ph.r[1], n, i, j);
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
            int kk = 0;
            for (; kk + 4 <= rem; kk += 4) {
                LSLAM_PHE_STEP(P[k + kk]);
                LSLAM_PHE_STEP(P[k + kk + 1]);
                LSLAM_PHE_STEP(P[k + kk + 2]);
                LSLAM_PHE_STEP(P[k + kk + 3]);
            }
            for (; kk < rem; ++kk) LSLAM_PHE_STEP(P[k + kk]);
#pragma unroll
            for (int k_ = 0; k_ < TPT; ++k_) cnt[k_] += __popc(bits[k_] & ((1u << rem) - 1u));
        }
#undef LSLAM_PHE_STEP
        bool any_flag = false;
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) {
            const int t = t0 + k_;
            if (t < T) {
                // literal re-score wanted: coincident samples, denormal-scale chunk, a point inside
                // the band around the threshold, or a non-finite fast sum
                const bool flag = deg[k_] || cb.force_literal || __float_as_int(mab[k_]) < cb.band_hi ||
                                  !(ss[k_] <= 1.7976931348623157e308);
                sm.cnt[q * T + t] = skip ? 0 : (flag ? kCntFlagged : cnt[k_]);
                sm.ssf[q * T + t] = ss[k_];
                any_flag |= flag && !skip;
            }
        }
        if (any_flag) st->flag = 1;
    }
    __syncthreads();

    // ================= select: warp per chunk - flagged re-score, max count, candidate list ========
    for (int q = warp; q < nsl; q += nwarps) {
        SlotState *st = sm.slot + q;
        const ChunkBand cb = chunk_band(st->scale_hi);
        if (!st->valid || cb.bad) {
            if (lane == 0) st->status = !st->valid ? 0x7fffffff : kStatusNonFinite;   // 0x7fffffff: unused slot
            continue;
        }
        const int n = st->n;
        const double2 *P = LSLAM_SLOT_PTS(st);
        int *cq = sm.cnt + q * T;
        double *sq = sm.ssf + q * T;
        if (st->flag) {
            unsigned n_relit = 0;
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                unsigned m = __ballot_sync(LSLAM_FULL_MASK, t < T && cq[t] == kCntFlagged);
                while (m) {
                    const int tt = t0 + __ffs((int)m) - 1;
                    m &= m - 1;
                    int i, j, cl;
                    double sl;
                    trial_pair(a, st->gslot, st->scan, (uint32_t)st->c, (uint32_t)tt, (uint32_t)n, i, j);
                    warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, nullptr);
                    if (lane == 0) { cq[tt] = cl | kCntExact; sq[tt] = sl; }
                    ++n_relit;
                }
            }
            if (lane == 0 && n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);
            __syncwarp();
        }
        int mx = 0, minhi_q = 0x7fffffff, ncand = 0;
        if (T <= 128) {
            // the chunk's (count, fast sum) pairs live in registers: 4 per lane, trial t = lane + 32 r
            int cv[4];
            double sv[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int t = lane + 32 * r;
                cv[r] = t < T ? cq[t] : -2;                  // -2: never the maximum, never a candidate
                sv[r] = t < T ? sq[t] : 0.0;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) { const int v = cv[r] < 0 ? 0 : (cv[r] & ~kCntExact); mx = mx > v ? mx : v; }
            mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
            if (mx < 2) {                   // upstream's final estimate() needs 2 points
                if (lane == 0) { st->mx = mx; st->status = kStatusNoModel; }
                continue;
            }
            const SsumBound sb = slot_bound(st, n);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (cv[r] >= 0 && (cv[r] & ~kCntExact) == mx) {
                    const double e = (cv[r] & kCntExact) ? 0.0 : ssum_err(sb, sv[r]);
                    const int k = key_up(sv[r] + e);
                    minhi_q = minhi_q < k ? minhi_q : k;
                }
            }
            minhi_q = __reduce_min_sync(LSLAM_FULL_MASK, minhi_q);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (32 * r < T) {                            // uniform
                    const bool cand = cv[r] >= 0 && trial_is_candidate(cv[r], sv[r], mx, minhi_q, sb);
                    const unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
                    if (cand) {
                        const int pos = ncand + __popc(m & ((1u << lane) - 1u));
                        if (pos < kMaxCand) st->cand_t[pos] = lane + 32 * r;
                    }
                    ncand += __popc(m);
                }
            }
        } else {
        // maximal inlier count
        for (int t = lane; t < T; t += 32) { const int v = cq[t] & ~kCntExact; mx = mx > v ? mx : v; }
        mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
        if (mx < 2) {                       // upstream's final estimate() needs 2 points
            if (lane == 0) { st->mx = mx; st->status = kStatusNoModel; }
            continue;
        }
        // smallest upper bound of the literal residual sum among the tied trials (quantised up)
        const SsumBound sb = slot_bound(st, n);
        for (int t = lane; t < T; t += 32) {
            const int cv = cq[t];
            if ((cv & ~kCntExact) == mx) {
                const double v = sq[t];
                const double e = (cv & kCntExact) ? 0.0 : ssum_err(sb, v);
                const int k = key_up(v + e);
                minhi_q = minhi_q < k ? minhi_q : k;
            }
        }
        minhi_q = __reduce_min_sync(LSLAM_FULL_MASK, minhi_q);
        // candidate list in trial order
        for (int t0 = 0; t0 < T; t0 += 32) {
            const int t = t0 + lane;
            const bool cand = t < T && trial_is_candidate(cq[t], sq[t], mx, minhi_q, sb);
            const unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
            if (cand) {
                const int pos = ncand + __popc(m & ((1u << lane) - 1u));
                if (pos < kMaxCand) st->cand_t[pos] = t;
            }
            ncand += __popc(m);
        }
        }
        if (lane == 0) { st->mx = mx; st->ncand = ncand; st->minhi_q = minhi_q; }
    }
    __syncthreads();

    // ================= models: lane per (chunk, candidate) - Philox pair + literal 2-point model ====
    for (int e = tid; e < nsl * kMaxCand; e += blockDim.x) {
        const int q = e / kMaxCand, ci = e - q * kMaxCand;
        SlotState *st = sm.slot + q;
        if (st->status == 0 && ci < st->ncand) {
            const double2 *P = LSLAM_SLOT_PTS(st);
            int i, j;
            trial_pair(a, st->gslot, st->scan, (uint32_t)st->c, (uint32_t)st->cand_t[ci], (uint32_t)st->n, i, j);
            st->cand_i[ci] = i;
            st->cand_j[ci] = j;
            // a single unflagged candidate with no residual sum requested is decided on the fast model
            if (!(st->ncand == 1 && a.ssum == nullptr && !(sm.cnt[q * T + st->cand_t[0]] & kCntExact)))
                sm.cmodel[e] = lit_pair_model(P[i], P[j]);
        }
    }
    __syncthreads();

    // ================= decide: warp per chunk - literal tie-break in trial order ====================
    for (int q = warp; q < nsl; q += nwarps) {
        SlotState *st = sm.slot + q;
        if (st->status != 0) continue;
        const int n = st->n, mx = st->mx, ncand = st->ncand;
        const double2 *P = LSLAM_SLOT_PTS(st);
        uint32_t *wt = sm.wtmp + (size_t)warp * W;
        uint32_t *mq = sm.smask + (size_t)q * W;
        bool fault = false;
        int best_t = -1, best_i = -1, best_j = -1;
        double best_ss = __longlong_as_double(0x7ff0000000000000LL);
        if (ncand == 1 && a.ssum == nullptr && !(sm.cnt[q * T + st->cand_t[0]] & kCntExact)) {
            // One candidate, the caller does not want its residual sum: the winner is known and was
            // not flagged, i.e. no point lies inside the re-check band, so its literal inlier mask
            // is the fast one.  Rebuild the mask from the fast model (2 DFMA per point) instead of
            // the literal score; the count is still checked against the selected maximum.
            const int i = st->cand_i[0], j = st->cand_j[0];
            const FastLine f = fast_pair_model(P[i], P[j]);
            int c = 0;
            for (int k0 = 0; k0 < n; k0 += 32) {
                const int k = k0 + lane;
                bool in = false;
                if (k < n) {
                    const double2 p = P[k];
                    in = fabs(fma(p.y, f.ny, fma(p.x, f.nx, f.negc))) < a.thr;
                }
                const unsigned b = __ballot_sync(LSLAM_FULL_MASK, in);
                c += __popc(b);
                if (lane == 0) mq[k0 >> 5] = b;
            }
            __syncwarp();
            if (c != mx) fault = true;
            best_t = st->cand_t[0]; best_i = i; best_j = j;
            best_ss = __longlong_as_double(0x7ff8000000000000LL);
        } else if (ncand <= kMaxCand) {
            for (int ci = 0; ci < ncand; ++ci) {
                int cl;
                double sl;
                warp_literal_score(P, n, sm.cmodel[q * kMaxCand + ci], a.thr, lane, cl, sl, ncand == 1 ? mq : wt);
                __syncwarp();
                if (cl != mx) fault = true;
                if (sl < best_ss) {
                    best_ss = sl; best_t = st->cand_t[ci]; best_i = st->cand_i[ci]; best_j = st->cand_j[ci];
                    if (ncand > 1) for (int w = lane; w * 32 < n; w += 32) mq[w] = wt[w];
                }
                __syncwarp();
            }
        } else {
            // many candidates (e.g. noise-free collinear data): sweep the trials again
            int *cq = sm.cnt + q * T;
            double *sq = sm.ssf + q * T;
            const SsumBound sb = slot_bound(st, n);
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                const bool cand = t < T && trial_is_candidate(cq[t], sq[t], mx, st->minhi_q, sb);
                unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
                while (m) {
                    const int tt = t0 + __ffs((int)m) - 1;
                    m &= m - 1;
                    int i, j, cl;
                    double sl;
                    trial_pair(a, st->gslot, st->scan, (uint32_t)st->c, (uint32_t)tt, (uint32_t)n, i, j);
                    warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, wt);
                    __syncwarp();
                    if (cl != mx) fault = true;
                    if (sl < best_ss) {
                        best_ss = sl; best_t = tt; best_i = i; best_j = j;
                        for (int w = lane; w * 32 < n; w += 32) mq[w] = wt[w];
                    }
                    __syncwarp();
                }
            }
        }
        if (lane == 0) {
            st->best_t = best_t; st->best_i = best_i; st->best_j = best_j; st->best_ss = best_ss;
            if (ncand > 1) atomicAdd(a.counters + 1, (unsigned long long)(ncand - 1));
        }
        __syncwarp();
        if (fault || best_t < 0) {
            // a fast count disagreed with the literal one (error bound violated - never observed):
            // redo this chunk with the exact sequential rule
            if (lane == 0) atomicAdd(a.counters + 2, 1ULL);
            slot_exact_fallback(a, st, P, mq, wt, lane);
        }
    }
    __syncthreads();

    // ================= refit + outputs: four lanes per chunk ========================================
    // LineModelND.estimate(data[inliers]) as the principal axis of the 2x2 scatter, accumulated in
    // one pass relative to the winner's first sample point (an inlier, so the shift is benign).
    const int lanes_total = nsl * 4;
    for (int base = warp * 32; base < lanes_total; base += nwarps * 32) {
        const int e = base + lane;
        const bool act = e < lanes_total;
        const int q = act ? (e >> 2) : 0, r = e & 3;
        SlotState *st = sm.slot + q;
        const bool ok = act && st->status == 0;
        const double2 *P = LSLAM_SLOT_PTS(st);
        const uint32_t *mq = sm.smask + (size_t)q * W;
        const int n = st->n;
        double sx = 0.0, sy = 0.0, sxx = 0.0, sxy = 0.0, syy = 0.0;
        double2 p0 = make_double2(0.0, 0.0);
        if (ok) {
            p0 = P[st->best_i];
            for (int k = r; k < n; k += 4) {
                if ((mq[k >> 5] >> (k & 31)) & 1u) {
                    const double2 p = P[k];
                    const double dx = p.x - p0.x, dy = p.y - p0.y;
                    sx += dx; sy += dy;
                    sxx = fma(dx, dx, sxx); sxy = fma(dx, dy, sxy); syy = fma(dy, dy, syy);
                }
            }
        }
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
            sx += __shfl_xor_sync(LSLAM_FULL_MASK, sx, o);
            sy += __shfl_xor_sync(LSLAM_FULL_MASK, sy, o);
            sxx += __shfl_xor_sync(LSLAM_FULL_MASK, sxx, o);
            sxy += __shfl_xor_sync(LSLAM_FULL_MASK, sxy, o);
            syy += __shfl_xor_sync(LSLAM_FULL_MASK, syy, o);
        }
        if (act && r == 0) {
            const long long s = slot0 + q;
            double *L = a.line + s * 8;
            if (ok) {
                const double m = (double)st->mx;
                const double mdx = sx / m, mdy = sy / m;
                const double ox = p0.x + mdx, oy = p0.y + mdy;
                double dx, dy;
                principal_axis(fma(-sx, mdx, sxx), fma(-sx, mdy, sxy), fma(-sy, mdy, syy), dx, dy);
                int last = 0;
                for (int w = (n - 1) >> 5; w >= 0; --w) {
                    const unsigned mw = mq[w];
                    if (mw) { last = w * 32 + 31 - __clz((int)mw); break; }
                }
                // ransac_functions.py:26-30 (no guard for dir_x == 0, like the reference)
                const double sl = __ddiv_rn(dy, dx);
                const double ic = __dsub_rn(oy, __dmul_rn(sl, ox));
                const double tipx = P[last].x;
                const double tipy = __dadd_rn(__dmul_rn(tipx, sl), ic);
                L[0] = sl; L[1] = ic; L[2] = ox; L[3] = oy; L[4] = dx; L[5] = dy; L[6] = tipx; L[7] = tipy;
                a.ninl[s] = st->mx;
                a.win_trial[s] = st->best_t;
                a.win_pair[2 * s] = st->best_i;
                a.win_pair[2 * s + 1] = st->best_j;
                if (a.ssum != nullptr) a.ssum[s] = st->best_ss;
            } else {
                const double qnan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
                for (int k = 0; k < 8; ++k) L[k] = qnan;
                a.ninl[s] = st->status == 0x7fffffff ? 0 : st->status;
                a.win_trial[s] = -1;
                a.win_pair[2 * s] = -1;
                a.win_pair[2 * s + 1] = -1;
                if (a.ssum != nullptr) a.ssum[s] = qnan;
            }
        }
    }
    // inlier masks: consecutive slots are consecutive in global memory -> coalesced stores
    for (int i = tid; i < nsl * W; i += blockDim.x) {
        const int q = i / W, w = i - q * W;
        const SlotState *st = sm.slot + q;
        a.mask[slot0 * W + i] = (st->status == 0 && w * 32 < st->n) ? sm.smask[i] : 0u;
    }
#undef LSLAM_SLOT_PTS
}

} // namespace lslam
