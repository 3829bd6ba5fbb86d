// ransac_ws.cuh - K1/K2, warp-specialised persistent form (`ransac.kernel = 2`).
//
// Same semantics and the same arithmetic building blocks as ransac.cuh (the phase-structured
// kernel, `ransac.kernel = 1`); what changes is how the work is scheduled on an SM.  Measured on
// B200 (profiles/r01a_ransac_c2_ncu_summary.txt): with every warp of a CTA walking through
// stage -> score -> select -> decide -> refit between block barriers, and the co-resident CTAs
// doing so in lock step, the fp64 pipe - the unit that bounds this kernel - idles whenever the CTAs
// are in their latency-bound phases (52 % busy).  Here ONE persistent CTA per SM runs four roles
// concurrently on a ring of NB tile buffers in shared memory:
//   loader    (1 warp)     per tile: slot table (geometry of the G chunk slots), TMA bulk copy of the
//                          tile's points (cp.async.bulk, SASS UBLKCP, completion on an mbarrier), then
//                          per chunk: centred fp32 copies of the points, second moments, bound coefficients and
//                          the fp32 re-check band (the tile kernel's arithmetic, ransac_tile.cuh)
//   scorers   (NS warps)   the hot loop only: thread per (chunk, TPT trials), fp32 screening of every
//                          point-hypothesis, residual sum from the chunk moments; rounds of 32 work items are
//                          claimed from a shared-memory counter, so a warp flows into the next tile as soon as
//                          one runs dry; a work item's flagged trials are re-scored literally by the whole warp
//   post      (NP warps)   warp per chunk: candidate selection, literal tie-break, refit sums -> a small
//                          per-slot record
//   epilogue  (1 warp)     LANE per chunk slot: the scalar tail (divisions, principal axis, slope,
//                          intercept, tip) and the result stores, for all slots of a tile at once
// Tiles flow loader -> scorers -> post -> epilogue -> loader through five mbarriers per ring slot
// (full / ready / scored / postdone / free); there is no block-wide barrier after start-up.
#pragma once
#include "async_copy.cuh"
#include "ransac_tile.cuh"     // the tile kernel's screening arithmetic: tile_screen32, tile_moment_sum / _err, tile_sqrt_up

namespace lslam {

struct WsCtl {                      // per ring buffer
    Mbar full, ready, scored, postdone, freeb;   // ready: the loader has centred the points and filled the slot table
    int next_round;
    int pad;
};

struct WsSlotInfo {                 // written by the loader warp
    unsigned long long scan;        // global scan index (Philox counter words 2, 3)
    int off, n, c, valid;           // first point within the staged tile, points, chunk index, flag
    int scale_hi, force_lit;        // high word of the largest |coordinate| of the chunk; every trial scored literally
    // screening data of the tile kernel's arithmetic (ransac_tile.cuh): centre, moments of the centred points, bound of
    // | sum r^2 - M | = alpha + beta M (alpha includes the moment form's own error at the chunk's radius), fp32 band
    double c0x, c0y, Sx, Sy, Sxx, Sxy, Syy, nn, alpha, beta;
    float bandf, kscale;            // tile_screen32: the band, and 2 / (the power of two >= band) - the loop's scaling
};

struct WsRec {                      // post -> epilogue
    double sx, sy, sxx, sxy, syy;   // sums over the inliers relative to p0
    double p0x, p0y, tipx, best_ss;
    int mx, best_t, best_i, best_j; // inlier count, winner trial and pair
    int status, pad0, pad1, pad2;   // 0 ok; else the value for ninl (0 = unused slot)
};

struct WsPostScratch {              // per post warp (followed by 2*W mask words)
    int cand_t[kMaxCand];
};

struct WsLayout {
    size_t ctl, bufs, buf_stride, o_ptsf, o_ssf, o_cm, o_pij, o_info, o_rec, post, post_stride, total;
};

__host__ __device__ inline size_t align128(size_t b) { return (b + 127) & ~(size_t)127; }

__host__ __device__ inline WsLayout ws_layout(int G, int chunk, int T, int W, int NB, int NP)
{
    WsLayout L;
    L.ctl = 0;
    L.bufs = align128((size_t)NB * sizeof(WsCtl));
    L.o_ptsf = align128((size_t)G * chunk * 16);                   // fp64 points first (one bulk copy), then their centred fp32 twins
    L.o_ssf = L.o_ptsf + align128((size_t)G * ((chunk + 1) & ~1) * 8);
    L.o_cm = L.o_ssf + align128((size_t)G * T * 8);
    L.o_pij = L.o_cm + align128((size_t)G * T * 4);
    L.o_info = L.o_pij + align128((size_t)G * T * 4);
    L.o_rec = L.o_info + align128((size_t)G * sizeof(WsSlotInfo));
    L.buf_stride = L.o_rec + align128((size_t)G * sizeof(WsRec));
    L.post = L.bufs + (size_t)NB * L.buf_stride;
    L.post_stride = align128(sizeof(WsPostScratch) + (size_t)2 * W * 4);
    L.total = L.post + (size_t)NP * L.post_stride;
    return L;
}

// wait with back-off: the roles that are not the bottleneck wait a lot, and a bare spin would
// compete with the scorers for issue slots
template <unsigned NS_SLEEP>
__device__ __forceinline__ void ws_wait(Mbar *bar, unsigned parity)
{
#ifndef LSLAM_EMU
    if (mbar_try_wait(bar, parity)) return;
    for (unsigned tries = 0; !mbar_try_wait(bar, parity); ++tries) {
        __nanosleep(NS_SLEEP);                 // measured: a 64 ns sleep still spent 33 % of all issued instructions here
        if (tries > (1u << 24)) __trap();      // a protocol bug must surface as an error, not a hang
    }
#else
    mbar_wait(bar, parity);
#endif
}

// packed per-trial word written by the scorers: inlier count (low 16 bits) and the high half of the
// float whose bits are the high word of min_k | |s_k| - thr |  (0 = "re-score me literally")
__device__ __forceinline__ uint32_t ws_pack(int cnt, float mab) { return ((uint32_t)__float_as_int(mab) & 0x7fff0000u) | (uint32_t)cnt; }

// ---- scorer: one work item = TPT consecutive trials of one chunk --------------------------------
// The tile kernel's arithmetic (ransac_tile.cuh): the signed distance in fp32 on the chunk's centred points, line and
// threshold scaled by the slot's kscale so that a funnel shift by two takes the inlier bit (sign) and the far-from-the-
// threshold bit (exponent bit 30) of every point: 2 FFMA + FADD + SHF per point-hypothesis.  A trial with a point inside
// the slot's band is re-scored literally; the fast residual sum comes from the chunk's moments, once per trial.
// Called by ALL lanes of a scorer warp (has = the lane has a work item): flagged trials are re-scored literally by the
// whole warp on the spot and marked exact (packed word 0x7f00....), so the post warps only see finished scores.
constexpr uint32_t kWsExact = 0x7f000000u;
template <int TPT>
__device__ __forceinline__ void ws_score_item(const RansacArgs &a, int item, bool has, int lane, const WsSlotInfo *info,
                                              const double2 *pts, const float2 *ptsf, double *ssf, uint32_t *cm, uint32_t *pij)
{
    static_assert(TPT % 2 == 0, "a Philox block serves two trials");
    const int T = a.trials;
    const int q = has ? item / a.ipc : 0, tb = has ? item - q * a.ipc : 0;
    const WsSlotInfo *g = info + q;
    const double2 *P = pts + g->off;
    const int t0 = tb * TPT;
    const uint32_t n = (uint32_t)g->n;
    const bool valid = has && g->valid != 0, force_lit = g->force_lit != 0;
    const unsigned long long scan = g->scan;

    float nx[TPT], ny[TPT], ng[TPT];
    unsigned bits[TPT];
    int cnt[TPT];
    unsigned far[TPT];                                            // AND of the scored words (the far bits matter)
    unsigned deg = 0;
    const float ks = g->kscale;
#pragma unroll
    for (int k = 0; k < TPT; k += 2) {
        const Philox4 ph = philox4x32_10((uint32_t)(t0 + k) >> 1, (uint32_t)g->c, (uint32_t)scan, (uint32_t)(scan >> 32),
                                         (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int t = t0 + k + h;
            nx[k + h] = 0.f; ny[k + h] = 0.f; ng[k + h] = 0.f;
            bits[k + h] = 0u; cnt[k + h] = 0; far[k + h] = 0xffffffffu;
            if (valid && t < T) {
                int i, j;
                pair_from_words(h ? ph.r[2] : ph.r[0], h ? ph.r[3] : ph.r[1], n, i, j);
                pij[q * T + t] = (uint32_t)i | ((uint32_t)j << 16);
                const FastLine f = fast_pair_model(P[i], P[j]);
                const double cp = fma(f.nx, g->c0x, fma(f.ny, g->c0y, f.negc));
                const double a1 = fma(f.nx, g->Sxx, f.ny * g->Sxy), a2 = fma(f.nx, g->Sxy, f.ny * g->Syy);
                const double M = fma(cp, fma(g->nn, cp, 2.0 * fma(f.nx, g->Sx, f.ny * g->Sy)), fma(f.nx, a1, f.ny * a2));
                ssf[q * T + t] = M;
                if (f.degenerate || force_lit || !(fabs(M) <= 1.7976931348623157e308)) deg |= 1u << (k + h);
                nx[k + h] = (float)f.nx * ks; ny[k + h] = (float)f.ny * ks; ng[k + h] = (float)cp * ks;   // exact scalings
            }
        }
    }
    const float thf = (float)a.thr * ks;
    const float2 *Pf = ptsf + (size_t)q * ((a.chunk + 1) & ~1);
    const float4 *Pf4 = reinterpret_cast<const float4 *>(Pf);
    // f = ks (|s_f| - thr_f): sign bit = inlier, exponent bit 30 set <=> the unscaled | |s_f| - thr_f | >= band' >= bandf
    // (tile_screen32).  16 points per word: inlier bits in the odd positions, far bits in the even ones.
#define LSLAM_WS_STEP2(XA, YA, XB, YB)                                                            \
    {                                                                                             \
        _Pragma("unroll") for (int k_ = 0; k_ < TPT; ++k_) {                                      \
            const float fa_ = fabsf(fmaf((YA), ny[k_], fmaf((XA), nx[k_], ng[k_]))) - thf;        \
            const float fb_ = fabsf(fmaf((YB), ny[k_], fmaf((XB), nx[k_], ng[k_]))) - thf;        \
            bits[k_] = __funnelshift_l(__float_as_uint(fa_), bits[k_], 2);                        \
            bits[k_] = __funnelshift_l(__float_as_uint(fb_), bits[k_], 2);                        \
        }                                                                                         \
    }
    const int n_eff = (valid && !force_lit) ? (int)n : 0;
    int k = 0;
    for (; k + 16 <= n_eff; k += 16) {
#pragma unroll 4
        for (int kk = 0; kk < 16; kk += 2) { const float4 pp = Pf4[(k + kk) >> 1]; LSLAM_WS_STEP2(pp.x, pp.y, pp.z, pp.w); }
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) { cnt[k_] += __popc(bits[k_] & 0xaaaaaaaau); far[k_] &= bits[k_]; }
    }
    const int rem = n_eff - k;
    if (rem > 0) {
        // the tail in pairs; an odd last point is paired with itself (shifted in twice, one shift is undone)
#pragma unroll 2
        for (int kk = 0; kk < rem; kk += 2) {
            float4 pp;
            if (kk + 1 < rem) pp = Pf4[(k + kk) >> 1];
            else { const float2 p1 = Pf[k + kk]; pp = make_float4(p1.x, p1.y, p1.x, p1.y); }
            LSLAM_WS_STEP2(pp.x, pp.y, pp.z, pp.w);
        }
        const unsigned keep = 0xffffffffu >> (32 - 2 * rem);                   // rem < 16
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) {
            const unsigned w_ = (bits[k_] >> (2 * (rem & 1))) & keep;
            cnt[k_] += __popc(w_ & 0xaaaaaaaau);
            far[k_] &= w_ | ~keep;
        }
    }
#undef LSLAM_WS_STEP2
    unsigned flagm = 0;
#pragma unroll
    for (int k_ = 0; k_ < TPT; ++k_) {
        const int t = t0 + k_;
        if (valid && t < T) {
            const bool flagged = ((deg >> k_) & 1u) || (far[k_] & 0x55555555u) != 0x55555555u;
            if (flagged) flagm |= 1u << k_;
            cm[q * T + t] = ws_pack(cnt[k_], __int_as_float(0x7f800000));
        } else if (has && t < T) {
            cm[q * T + t] = ws_pack(0, 0.0f); ssf[q * T + t] = 0.0; pij[q * T + t] = 0u;
        }
    }
    // flagged trials: scored again by the whole warp, on the spot - with the fp64 fast line first (its own re-check band is
    // 2^-36 of the coordinate scale: the count is then the literal one and the moment-based sum stays), literally if a
    // point is inside that band as well or the trial was marked for it at the set-up (degenerate pair, slot out of range)
    const ChunkBand cbq = chunk_band(g->scale_hi);
    if (cbq.bad) flagm = 0;                                       // non-finite chunk: the post warp reports it
    __syncwarp();                                                 // the pairs written above are read by the other lanes below
    unsigned fm = __ballot_sync(LSLAM_FULL_MASK, flagm != 0u);
    unsigned n_relit = 0;
    while (fm) {
        const int src = __ffs((int)fm) - 1;
        fm &= fm - 1;
        const int fq = __shfl_sync(LSLAM_FULL_MASK, q, src), ft0 = __shfl_sync(LSLAM_FULL_MASK, t0, src);
        unsigned fl = __shfl_sync(LSLAM_FULL_MASK, flagm, src);
        const unsigned fdeg = __shfl_sync(LSLAM_FULL_MASK, deg, src);
        const int fband = __shfl_sync(LSLAM_FULL_MASK, cbq.band_hi, src);
        const WsSlotInfo *fg = info + fq;
        const double2 *FP = pts + fg->off;
        while (fl) {
            const int kf = __ffs((int)fl) - 1;
            fl &= fl - 1;
            const int tt = ft0 + kf;
            int cl = 0;
            double sl = 0.0;
            const uint32_t pw_ = pij[fq * T + tt];
            const double2 pi_ = FP[pw_ & 0xffffu], pj_ = FP[pw_ >> 16];
            bool lit = (fdeg >> kf) & 1u;
            if (!lit) warp_fast64_score_strided(FP, fg->n, fast_pair_model(pi_, pj_), a.thr, fband, lane, cl, lit, nullptr, 0);
            if (lit) {
                warp_literal_score(FP, fg->n, lit_pair_model(pi_, pj_), a.thr, lane, cl, sl, nullptr);
                ++n_relit;
            }
            if (lane == src) {
                if (lit) { cm[fq * T + tt] = kWsExact | (uint32_t)cl; ssf[fq * T + tt] = sl; }
                else cm[fq * T + tt] = ws_pack(cl, __int_as_float(0x7f800000));
            }
        }
    }
    if (lane == 0 && n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);
}

// ---- post: from the scores of one chunk to its record, inside one warp --------------------------
__device__ __forceinline__ void ws_rec_status(WsRec *rec, int status, const RansacArgs &a, long long s, int lane)
{
    if (lane == 0) rec->status = status;
    for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = 0u;
}

// literal model of trial t from the pair the scorer stored
__device__ __forceinline__ LitLine ws_trial_model(const double2 *P, const uint32_t *pq, int t, int &i, int &j)
{
    const uint32_t p = pq[t];
    i = (int)(p & 0xffffu);
    j = (int)(p >> 16);
    return lit_pair_model(P[i], P[j]);
}

__device__ __forceinline__ void ws_post_chunk(const RansacArgs &a, long long s, const WsSlotInfo &g, const double2 *pts,
                                              double *sq, uint32_t *cmq, const uint32_t *pq, WsRec *rec,
                                              WsPostScratch *sc, uint32_t *smask, uint32_t *wtmp, int lane)
{
    const int T = a.trials;
    if (!g.valid) { ws_rec_status(rec, 0x7fffffff, a, s, lane); return; }    // unused slot
    const int n = g.n;
    const double2 *P = pts + g.off;
    ChunkBand cb = chunk_band(g.scale_hi);
    if (cb.bad) { ws_rec_status(rec, kStatusNonFinite, a, s, lane); return; }
    cb.force_literal = cb.force_literal || g.force_lit != 0;
    SsumBound sb;                                                 // | sum r^2 - M | <= alpha + beta M (the loader's slot table)
    sb.alpha = g.alpha; sb.beta = g.beta; sb.kappa = 0.0;
    int *cq = reinterpret_cast<int *>(cmq);
    int mx = 0, minhi_q = 0x7fffffff, ncand = 0;
    unsigned n_relit = 0;
    if (T <= 128) {
        // ---- the chunk's trials live in registers: lane holds trials lane, lane+32, lane+64, lane+96 ----
        int cv[4];
        double v[4];
        unsigned fl[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int t = lane + 32 * k;
            const uint32_t w = t < T ? cmq[t] : 0x7fff0000u;
            v[k] = t < T ? sq[t] : 0.0;
            cv[k] = t < T ? (int)(w & 0xffffu) : -1;
            if (t < T && (w & 0x7fff0000u) == kWsExact) cv[k] |= kCntExact;      // re-scored literally by its scorer warp
            fl[k] = __ballot_sync(LSLAM_FULL_MASK, t < T && (w & 0x7fff0000u) != kWsExact &&
                                                       (cb.force_literal || (int)(w & 0x7fff0000u) < cb.band_hi));
        }
        if (fl[0] | fl[1] | fl[2] | fl[3]) {                    // literal re-scores (0.16 % of the trials)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                unsigned m = fl[k];
                while (m) {
                    const int src = __ffs((int)m) - 1, tt = src + 32 * k;
                    m &= m - 1;
                    int i, j, cl;
                    double sl;
                    const LitLine mdl = ws_trial_model(P, pq, tt, i, j);
                    warp_literal_score(P, n, mdl, a.thr, lane, cl, sl, nullptr);
                    if (src == lane) { cv[k] = cl | kCntExact; v[k] = sl; }
                    ++n_relit;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { const int c = cv[k] < 0 ? 0 : (cv[k] & ~kCntExact); mx = mx > c ? mx : c; }
        mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
        if (mx >= 2) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (cv[k] >= 0 && (cv[k] & ~kCntExact) == mx) {
                    const double e = (cv[k] & kCntExact) ? 0.0 : fma(sb.beta, v[k], sb.alpha);
                    const int key = key_up(v[k] + e);
                    minhi_q = minhi_q < key ? minhi_q : key;
                }
            }
            minhi_q = __reduce_min_sync(LSLAM_FULL_MASK, minhi_q);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int t = lane + 32 * k;
                const bool cand = cv[k] >= 0 && trial_is_candidate(cv[k], v[k], mx, minhi_q, sb);
                const unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
                if (cand) {
                    const int pos = ncand + __popc(m & ((1u << lane) - 1u));
                    if (pos < kMaxCand) sc->cand_t[pos] = t;
                }
                ncand += __popc(m);
            }
            if (ncand > kMaxCand) {                              // unpacked form for the rare sweep below
#pragma unroll
                for (int k = 0; k < 4; ++k) { const int t = lane + 32 * k; if (t < T) { cq[t] = cv[k]; sq[t] = v[k]; } }
            }
        }
    } else {
        // ---- generic trial count: the same steps as loops over shared memory ----
        for (int t0 = 0; t0 < T; t0 += 32) {
            const int t = t0 + lane;
            const uint32_t w = t < T ? cmq[t] : 0x7fff0000u;
            int cv = (int)(w & 0xffffu);
            if (t < T && (w & 0x7fff0000u) == kWsExact) cv |= kCntExact;            // re-scored literally by its scorer warp
            const bool flag = t < T && !(cv & kCntExact) && (cb.force_literal || (int)(w & 0x7fff0000u) < cb.band_hi);
            unsigned m = __ballot_sync(LSLAM_FULL_MASK, flag);
            while (m) {
                const int tt = t0 + __ffs((int)m) - 1;
                m &= m - 1;
                int i, j, cl;
                double sl;
                const LitLine mdl = ws_trial_model(P, pq, tt, i, j);
                warp_literal_score(P, n, mdl, a.thr, lane, cl, sl, nullptr);
                if (tt == t) { cv = cl | kCntExact; sq[t] = sl; }
                ++n_relit;
            }
            if (t < T) cq[t] = cv;
            const int c = cv & ~kCntExact;
            mx = mx > c ? mx : c;
        }
        mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
        __syncwarp();
        if (mx >= 2) {
            for (int t = lane; t < T; t += 32) {
                const int cv = cq[t];
                if ((cv & ~kCntExact) == mx) {
                    const double v = sq[t];
                    const double e = (cv & kCntExact) ? 0.0 : fma(sb.beta, v, sb.alpha);
                    const int key = key_up(v + e);
                    minhi_q = minhi_q < key ? minhi_q : key;
                }
            }
            minhi_q = __reduce_min_sync(LSLAM_FULL_MASK, minhi_q);
            for (int t0 = 0; t0 < T; t0 += 32) {
                const int t = t0 + lane;
                const bool cand = t < T && trial_is_candidate(cq[t], sq[t], mx, minhi_q, sb);
                const unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
                if (cand) {
                    const int pos = ncand + __popc(m & ((1u << lane) - 1u));
                    if (pos < kMaxCand) sc->cand_t[pos] = t;
                }
                ncand += __popc(m);
            }
        }
    }
    if (lane == 0 && n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);
    if (mx < 2) { ws_rec_status(rec, kStatusNoModel, a, s, lane); return; }  // upstream's final estimate() needs 2 points
    __syncwarp();
    // ---- literal tie-break in trial order (first best wins, like the sequential accept rule) ----
    bool fault = false;
    int best_t = -1, best_i = -1, best_j = -1;
    double best_ss = __longlong_as_double(0x7ff0000000000000LL);
    if (ncand <= kMaxCand) {
        for (int ci = 0; ci < ncand; ++ci) {
            int i, j, cl;
            double sl;
            const int tt = sc->cand_t[ci];
            const LitLine mdl = ws_trial_model(P, pq, tt, i, j);      // every lane, redundantly: no hand-off
            warp_literal_score(P, n, mdl, a.thr, lane, cl, sl, ncand == 1 ? smask : wtmp);
            __syncwarp();
            if (cl != mx) fault = true;
            if (sl < best_ss) {
                best_ss = sl; best_t = tt; best_i = i; best_j = j;
                if (ncand > 1) for (int w = lane; w * 32 < n; w += 32) smask[w] = wtmp[w];
            }
            __syncwarp();
        }
    } else {
        for (int t0 = 0; t0 < T; t0 += 32) {                     // many candidates (e.g. noise-free collinear data)
            const int t = t0 + lane;
            const bool cand = t < T && trial_is_candidate(cq[t], sq[t], mx, minhi_q, sb);
            unsigned m = __ballot_sync(LSLAM_FULL_MASK, cand);
            while (m) {
                const int tt = t0 + __ffs((int)m) - 1;
                m &= m - 1;
                int i, j, cl;
                double sl;
                const LitLine mdl = ws_trial_model(P, pq, tt, i, j);
                warp_literal_score(P, n, mdl, a.thr, lane, cl, sl, wtmp);
                __syncwarp();
                if (cl != mx) fault = true;
                if (sl < best_ss) {
                    best_ss = sl; best_t = tt; best_i = i; best_j = j;
                    for (int w = lane; w * 32 < n; w += 32) smask[w] = wtmp[w];
                }
                __syncwarp();
            }
        }
    }
    if (lane == 0 && ncand > 1) atomicAdd(a.counters + 1, (unsigned long long)(ncand - 1));
    if (fault || best_t < 0) {
        // a fast count disagreed with the literal one (error bound violated - never observed):
        // redo this chunk with the exact sequential rule
        if (lane == 0) atomicAdd(a.counters + 2, 1ULL);
        best_t = -1; best_ss = __longlong_as_double(0x7ff0000000000000LL);
        int best_n = 0;
        for (int t = 0; t < T; ++t) {
            int i, j, cl;
            double sl;
            const LitLine mdl = ws_trial_model(P, pq, t, i, j);
            warp_literal_score(P, n, mdl, a.thr, lane, cl, sl, wtmp);
            __syncwarp();
            if (cl > best_n || (cl == best_n && sl < best_ss)) {
                best_n = cl; best_ss = sl; best_t = t; best_i = i; best_j = j;
                for (int w = lane; w * 32 < n; w += 32) smask[w] = wtmp[w];
            }
            __syncwarp();
        }
        mx = best_n;
        if (best_t < 0 || best_n < 2) { ws_rec_status(rec, kStatusNoModel, a, s, lane); return; }
    }
    // ---- refit sums over the inliers, relative to the winner's first sample point ----
    const double2 p0 = P[best_i];
    double sx = 0.0, sy = 0.0, sxx = 0.0, sxy = 0.0, syy = 0.0;
    for (int k = lane; k < n; k += 32) {
        if ((smask[k >> 5] >> (k & 31)) & 1u) {
            const double2 p = P[k];
            const double dx = p.x - p0.x, dy = p.y - p0.y;
            sx += dx; sy += dy;
            sxx = fma(dx, dx, sxx); sxy = fma(dx, dy, sxy); syy = fma(dy, dy, syy);
        }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        sx += __shfl_xor_sync(LSLAM_FULL_MASK, sx, o);
        sy += __shfl_xor_sync(LSLAM_FULL_MASK, sy, o);
        sxx += __shfl_xor_sync(LSLAM_FULL_MASK, sxx, o);
        sxy += __shfl_xor_sync(LSLAM_FULL_MASK, sxy, o);
        syy += __shfl_xor_sync(LSLAM_FULL_MASK, syy, o);
    }
    if (lane == 0) {
        int last = 0;
        for (int w = (n - 1) >> 5; w >= 0; --w) {
            const unsigned mw = smask[w];
            if (mw) { last = w * 32 + 31 - __clz((int)mw); break; }
        }
        rec->sx = sx; rec->sy = sy; rec->sxx = sxx; rec->sxy = sxy; rec->syy = syy;
        rec->p0x = p0.x; rec->p0y = p0.y; rec->tipx = P[last].x; rec->best_ss = best_ss;
        rec->mx = mx; rec->best_t = best_t; rec->best_i = best_i; rec->best_j = best_j;
        rec->status = 0;
    }
    for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = (w * 32 < n) ? smask[w] : 0u;
    __syncwarp();
}

// ---- epilogue: one LANE per chunk slot - LineModelND.estimate tail + ransac_functions.py:26-30 ----
__device__ __forceinline__ void ws_epilogue_slot(const RansacArgs &a, long long s, const WsRec &r)
{
    double *L = a.line + s * 8;
    if (r.status == 0) {
        const double m = (double)r.mx;
        const double mdx = r.sx / m, mdy = r.sy / m;
        const double ox = r.p0x + mdx, oy = r.p0y + mdy;
        double dx, dy;
        principal_axis(fma(-r.sx, mdx, r.sxx), fma(-r.sx, mdy, r.sxy), fma(-r.sy, mdy, r.syy), dx, dy);
        // no guard for dir_x == 0, like the reference
        const double sl = __ddiv_rn(dy, dx);
        const double ic = __dsub_rn(oy, __dmul_rn(sl, ox));
        const double tipy = __dadd_rn(__dmul_rn(r.tipx, sl), ic);
        L[0] = sl; L[1] = ic; L[2] = ox; L[3] = oy; L[4] = dx; L[5] = dy; L[6] = r.tipx; L[7] = tipy;
        a.ninl[s] = r.mx;
        a.win_trial[s] = r.best_t;
        a.win_pair[2 * s] = r.best_i;
        a.win_pair[2 * s + 1] = r.best_j;
        if (a.ssum != nullptr) a.ssum[s] = r.best_ss;
    } else {
        const double qnan = __longlong_as_double(0x7ff8000000000000LL);
#pragma unroll
        for (int k = 0; k < 8; ++k) L[k] = qnan;
        a.ninl[s] = r.status == 0x7fffffff ? 0 : r.status;
        a.win_trial[s] = -1;
        a.win_pair[2 * s] = -1;
        a.win_pair[2 * s + 1] = -1;
        if (a.ssum != nullptr) a.ssum[s] = qnan;
    }
}

// ---- the kernel: one persistent CTA per SM, (NS + NP + 2) warps -----------------------------------
// a.G = chunk slots per tile, a.ws_ns / a.ws_np / a.ws_nb = scorer warps, post warps, ring depth.
template <int TPT, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) ransac_ws_kernel(const RansacArgs a)
{
    LSLAM_SMEM_DECL;
    const int T = a.trials, G = a.G, NS = a.ws_ns, NP = a.ws_np, NB = a.ws_nb;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const WsLayout L = ws_layout(G, a.chunk, T, a.W, NB, NP);
    WsCtl *ctl = reinterpret_cast<WsCtl *>(smem_raw + L.ctl);
    const long long ntiles = (a.nslots + G - 1) / G;

    if (tid == 0) {
        for (int b = 0; b < NB; ++b) {
            mbar_init(&ctl[b].full, 1);
            mbar_init(&ctl[b].ready, 1);
            mbar_init(&ctl[b].scored, (unsigned)NS + 1u);
            mbar_init(&ctl[b].postdone, (unsigned)NP);
            mbar_init(&ctl[b].freeb, 1);
            ctl[b].next_round = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();

    long long k = 0;                                   // this CTA's k-th tile goes to ring slot k % NB
    if (warp == NS + NP) {
        // ================= loader =================
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            const long long use = k / NB;
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            WsSlotInfo *info = reinterpret_cast<WsSlotInfo *>(buf + L.o_info);
            const double2 *pts = reinterpret_cast<const double2 *>(buf);
            if (use > 0) ws_wait<400>(&ctl[b].freeb, (unsigned)((use - 1) & 1));
            const long long slot0 = tile * G;
            const long long remain = a.nslots - slot0;
            const int nsl = remain < (long long)G ? (int)remain : G;
            const long long b0 = slot0 / a.C;
            const int c0 = (int)(slot0 - b0 * a.C);
            const long long gp0 = b0 * a.N + (long long)c0 * a.chunk;       // first staged point
            for (int q = lane; q < nsl; q += 32) {                          // slot table (functions.py:64-76)
                const int cc = c0 + q;
                const int db = cc / a.C, c = cc - db * a.C;
                const long long bb = b0 + db;
                int nb = a.npts ? a.npts[bb] : a.N;
                nb = nb < 0 ? 0 : (nb > a.N ? a.N : nb);
                int n = nb - c * a.chunk;
                n = n < 0 ? 0 : (n > a.chunk ? a.chunk : n);
                WsSlotInfo g;
                g.scan = a.scan_base + (unsigned long long)bb;
                g.off = (int)(bb * a.N + (long long)c * a.chunk - gp0);
                g.n = n; g.c = c;
                g.valid = (n >= 2) && (n == a.chunk || n > 2);
                g.scale_hi = 0; g.force_lit = 0;
                g.c0x = 0.0; g.c0y = 0.0; g.Sx = 0.0; g.Sy = 0.0; g.Sxx = 0.0; g.Sxy = 0.0; g.Syy = 0.0;
                g.nn = (double)n; g.alpha = 0.0; g.beta = 0.0; g.bandf = 0.0f; g.kscale = 0.0f;
                info[q] = g;
            }
            __syncwarp();
            if (lane == 0) {
                const long long sl = slot0 + nsl - 1;
                const long long bl = sl / a.C, cl = sl - bl * a.C;
                const long long cend = (cl + 1) * a.chunk < a.N ? (cl + 1) * a.chunk : a.N;
                const unsigned bytes = (unsigned)((bl * a.N + cend - gp0) * 16);
                ctl[b].next_round = 0;
                mbar_expect_tx(&ctl[b].full, bytes);
                bulk_g2s(buf, a.xy + gp0, bytes, &ctl[b].full);
            }
            ws_wait<200>(&ctl[b].full, (unsigned)(use & 1));
            // per chunk slot, by the whole warp: coordinate scale, the points centred on the chunk's middle point and rounded
            // to fp32 (what the scorers' loop reads), their moments and radius -> the slot's screening data
            float2 *ptsf = reinterpret_cast<float2 *>(buf + L.o_ptsf);
            const int chunk_pad = (a.chunk + 1) & ~1;
            for (int q = 0; q < nsl; ++q) {
                const int off = info[q].off, n = info[q].n;
                const double2 cc = n > 0 ? pts[off + (n >> 1)] : make_double2(0.0, 0.0);
                float2 *dstf = ptsf + (size_t)q * chunk_pad;
                int mh = 0;
                double sx = 0.0, sy = 0.0, sxx = 0.0, sxy = 0.0, syy = 0.0, r2 = 0.0;
                for (int i = lane; i < n; i += 32) {
                    const double2 p = pts[off + i];
                    const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
                    mh = mh > hx ? mh : hx;
                    mh = mh > hy ? mh : hy;
                    const double qx = p.x - cc.x, qy = p.y - cc.y;
                    sx += qx; sy += qy;
                    sxx = fma(qx, qx, sxx); sxy = fma(qx, qy, sxy); syy = fma(qy, qy, syy);
                    dstf[i] = make_float2((float)qx, (float)qy);
                    r2 = fmax(r2, fma(qx, qx, qy * qy));
                }
                mh = __reduce_max_sync(LSLAM_FULL_MASK, mh);
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    sx += __shfl_xor_sync(LSLAM_FULL_MASK, sx, o);
                    sy += __shfl_xor_sync(LSLAM_FULL_MASK, sy, o);
                    sxx += __shfl_xor_sync(LSLAM_FULL_MASK, sxx, o);
                    sxy += __shfl_xor_sync(LSLAM_FULL_MASK, sxy, o);
                    syy += __shfl_xor_sync(LSLAM_FULL_MASK, syy, o);
                    r2 = fmax(r2, __shfl_xor_sync(LSLAM_FULL_MASK, r2, o));
                }
                if (lane == 0) {
                    WsSlotInfo *g = info + q;
                    const ChunkBand cb = chunk_band(mh);
                    const double R = (double)tile_sqrt_up(r2);                     // upper bound of the radius (inf / 0: out of range)
                    const TileScreen32 scr = tile_screen32(R, a.thr, cb.eabs * 0x1p36);
                    const SsumBound sb = ssum_bound_coeffs(cb.eabs, n, a.thr);
                    // the moment form's own error at the largest |cp| a line through two points of the chunk can have (the radius)
                    const double e2 = tile_moment_err(R, (double)n, sxx + syy);
                    g->scale_hi = mh;
                    g->force_lit = (cb.force_literal || scr.force_literal || !(e2 <= 1.7976931348623157e308)) ? 1 : 0;
                    g->c0x = cc.x; g->c0y = cc.y; g->Sx = sx; g->Sy = sy; g->Sxx = sxx; g->Sxy = sxy; g->Syy = syy;
                    g->alpha = fma(1.0 + sb.beta, e2, sb.alpha); g->beta = sb.beta; g->bandf = scr.bandf; g->kscale = scr.kscale;
                }
            }
            __syncwarp();
            if (lane == 0) { mbar_arrive(&ctl[b].ready); }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl[b].scored);
        }
    } else if (warp == NS + NP + 1) {
        // ================= epilogue =================
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            const WsRec *rec = reinterpret_cast<const WsRec *>(buf + L.o_rec);
            const long long slot0 = tile * G;
            const long long remain = a.nslots - slot0;
            const int nsl = remain < (long long)G ? (int)remain : G;
            ws_wait<800>(&ctl[b].postdone, (unsigned)((k / NB) & 1));
            for (int q = lane; q < nsl; q += 32) ws_epilogue_slot(a, slot0 + q, rec[q]);
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl[b].freeb);
        }
    } else if (warp < NS) {
        // ================= scorers =================
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            const long long remain = a.nslots - tile * G;
            const int nsl = remain < (long long)G ? (int)remain : G;
            const int items = nsl * a.ipc, nrounds = (items + 31) >> 5;
            ws_wait<100>(&ctl[b].ready, (unsigned)((k / NB) & 1));
            for (;;) {
                int r = 0;
                if (lane == 0) r = atomicAdd(&ctl[b].next_round, 1);
                r = __shfl_sync(LSLAM_FULL_MASK, r, 0);
                if (r >= nrounds) break;
                const int item = r * 32 + lane;
                ws_score_item<TPT>(a, item, item < items, lane, reinterpret_cast<const WsSlotInfo *>(buf + L.o_info),
                                       reinterpret_cast<const double2 *>(buf), reinterpret_cast<const float2 *>(buf + L.o_ptsf),
                                       reinterpret_cast<double *>(buf + L.o_ssf),
                                       reinterpret_cast<uint32_t *>(buf + L.o_cm), reinterpret_cast<uint32_t *>(buf + L.o_pij));
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl[b].scored);
        }
    } else {
        // ================= post =================
        const int pw = warp - NS;
        unsigned char *ps = smem_raw + L.post + (size_t)pw * L.post_stride;
        WsPostScratch *sc = reinterpret_cast<WsPostScratch *>(ps);
        uint32_t *smask = reinterpret_cast<uint32_t *>(ps + sizeof(WsPostScratch));
        uint32_t *wtmp = smask + a.W;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            const WsSlotInfo *info = reinterpret_cast<const WsSlotInfo *>(buf + L.o_info);
            WsRec *rec = reinterpret_cast<WsRec *>(buf + L.o_rec);
            const long long slot0 = tile * G;
            const long long remain = a.nslots - slot0;
            const int nsl = remain < (long long)G ? (int)remain : G;
            ws_wait<400>(&ctl[b].scored, (unsigned)((k / NB) & 1));
            for (int q = pw; q < nsl; q += NP)
                ws_post_chunk(a, slot0 + q, info[q], reinterpret_cast<const double2 *>(buf),
                              reinterpret_cast<double *>(buf + L.o_ssf) + (size_t)q * T,
                              reinterpret_cast<uint32_t *>(buf + L.o_cm) + (size_t)q * T,
                              reinterpret_cast<const uint32_t *>(buf + L.o_pij) + (size_t)q * T, rec + q, sc, smask,
                              wtmp, lane);
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl[b].postdone);
        }
    }
}

} // namespace lslam
