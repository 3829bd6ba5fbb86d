// ransac_ws.cuh - K1/K2, warp-specialised persistent form (the default RANSAC kernel).
//
// Same semantics and the same arithmetic building blocks as ransac.cuh (the phase-structured
// kernel, kept as `ransac.kernel = 1`); what changes is how the work is scheduled on an SM.
// Measured on B200 (profiles/r01a_ransac_c2_ncu_summary.txt): with every warp of a CTA walking
// through stage -> score -> select -> decide -> refit between block barriers, and the co-resident
// CTAs doing so in lock step, the fp64 pipe - the unit that bounds this kernel - idles whenever the
// CTAs are in their latency-bound phases (52 % busy).  Here ONE persistent CTA per SM runs three
// roles concurrently on a ring of NB tile buffers in shared memory:
//   producer  (1 thread)   TMA bulk copy (cp.async.bulk, SASS UBLKCP) of the next tile's points,
//                          completion on an mbarrier; runs NB tiles ahead of the scorers
//   scorers   (NS warps)   the hot loop only: thread per (chunk, TPT trials), 3 DFMA + 1 DADD per
//                          point-hypothesis; rounds of 32 work items are claimed with a shared-memory
//                          counter, so a warp moves on to the next tile as soon as one runs dry
//   post      (NP warps)   per chunk, start to finish inside one warp: re-check band, literal
//                          re-scores, candidate selection, tie-break, refit, outputs
// Tiles flow producer -> scorers -> post -> producer through three mbarriers per buffer
// (full / scored / free); there is no block-wide barrier after start-up.
#pragma once
#include "async_copy.cuh"
#include "ransac.cuh"

namespace lslam {

struct WsCtl {                      // per ring buffer
    Mbar full, scored, freeb;
    int next_round;
    int pad;
};

struct WsPostScratch {              // per post warp (followed by 2*W mask words)
    int cand_t[kMaxCand], cand_i[kMaxCand], cand_j[kMaxCand];
    LitLine cmodel[kMaxCand];
};

struct WsLayout {
    size_t ctl, bufs, buf_stride, o_ssf, o_cm, post, post_stride, total;
};

__host__ __device__ inline size_t align128(size_t b) { return (b + 127) & ~(size_t)127; }

__host__ __device__ inline WsLayout ws_layout(int G, int chunk, int T, int W, int NB, int NP)
{
    WsLayout L;
    L.ctl = 0;
    L.bufs = align128((size_t)NB * sizeof(WsCtl));
    L.o_ssf = align128((size_t)G * chunk * 16);
    L.o_cm = L.o_ssf + align128((size_t)G * T * 8);
    L.buf_stride = L.o_cm + align128((size_t)G * T * 4);
    L.post = L.bufs + (size_t)NB * L.buf_stride;
    L.post_stride = align128(sizeof(WsPostScratch) + (size_t)2 * W * 4);
    L.total = L.post + (size_t)NP * L.post_stride;
    return L;
}

// packed per-trial word written by the scorers: inlier count (low 16 bits) and the high half of the
// float whose bits are the high word of min_k | |s_k| - thr |  (0 = "re-score me literally")
__device__ __forceinline__ uint32_t ws_pack(int cnt, float mab) { return ((uint32_t)__float_as_int(mab) & 0x7fff0000u) | (uint32_t)cnt; }

// geometry of chunk slot s: scan b, chunk c, point count n, validity (functions.py:64-76)
struct WsSlot { int b, c, n; bool valid; };
__device__ __forceinline__ WsSlot ws_slot(const RansacArgs &a, long long s)
{
    WsSlot g;
    g.b = (int)(s / a.C);
    g.c = (int)(s - (long long)g.b * a.C);
    int nb = a.npts ? a.npts[g.b] : a.N;
    nb = nb < 0 ? 0 : (nb > a.N ? a.N : nb);
    int n = nb - g.c * a.chunk;
    g.n = n < 0 ? 0 : (n > a.chunk ? a.chunk : n);
    g.valid = (g.n >= 2) && (g.n == a.chunk || g.n > 2);
    return g;
}

// ---- scorer: one work item = TPT consecutive trials of one chunk --------------------------------
template <int TPT>
__device__ __forceinline__ void ws_score_item(const RansacArgs &a, long long slot0, long long gp0, int item,
                                              const double2 *pts, double *ssf, uint32_t *cm)
{
    const int T = a.trials;
    const int q = item / a.ipc, tb = item - q * a.ipc;
    const WsSlot g = ws_slot(a, slot0 + q);
    const double2 *P = pts + ((long long)g.b * a.N + (long long)g.c * a.chunk - gp0);
    const uint64_t scan = a.scan_base + (uint64_t)g.b;
    const int t0 = tb * TPT;
    const uint32_t n = (uint32_t)g.n;

    double nx[TPT], ny[TPT], ng[TPT], ss[TPT];
    unsigned bits[TPT];
    int cnt[TPT];
    float mab[TPT];
    Philox4 ph;
#pragma unroll
    for (int k = 0; k < TPT; ++k) {
        const int t = t0 + k;
        if ((k & 1) == 0 || TPT == 1)
            ph = philox4x32_10((uint32_t)t >> 1, (uint32_t)g.c, (uint32_t)scan, (uint32_t)(scan >> 32),
                               (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
        nx[k] = 0.0; ny[k] = 0.0; ng[k] = 0.0; ss[k] = 0.0;
        bits[k] = 0u; cnt[k] = 0; mab[k] = __int_as_float(0x7f800000);
        if (g.valid && t < T) {
            int i, j;
            const bool odd = t & 1;
            pair_from_words(odd ? ph.r[2] : ph.r[0], odd ? ph.r[3] : ph.r[1], n, i, j);
            const FastLine f = fast_pair_model(P[i], P[j]);
            nx[k] = f.nx; ny[k] = f.ny; ng[k] = f.negc;
            if (f.degenerate) mab[k] = 0.0f;                     // coincident samples: literal re-score
        }
    }
#define LSLAM_WS_STEP(PT)                                                        \
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
    const int n_eff = g.valid ? (int)n : 0;
    int k = 0;
    for (; k + 32 <= n_eff; k += 32) {
#pragma unroll 8
        for (int kk = 0; kk < 32; ++kk) LSLAM_WS_STEP(P[k + kk]);
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) cnt[k_] += __popc(bits[k_]);
    }
    const int rem = n_eff - k;
    if (rem > 0) {
        int kk = 0;
        for (; kk + 4 <= rem; kk += 4) {
            LSLAM_WS_STEP(P[k + kk]);
            LSLAM_WS_STEP(P[k + kk + 1]);
            LSLAM_WS_STEP(P[k + kk + 2]);
            LSLAM_WS_STEP(P[k + kk + 3]);
        }
        for (; kk < rem; ++kk) LSLAM_WS_STEP(P[k + kk]);
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) cnt[k_] += __popc(bits[k_] & ((1u << rem) - 1u));
    }
#undef LSLAM_WS_STEP
#pragma unroll
    for (int k_ = 0; k_ < TPT; ++k_) {
        const int t = t0 + k_;
        if (t < T) {
            const bool finite = ss[k_] <= 1.7976931348623157e308;
            ssf[q * T + t] = ss[k_];
            cm[q * T + t] = ws_pack(cnt[k_], finite ? mab[k_] : 0.0f);
        }
    }
}

// ---- post: everything after the scores of one chunk, inside one warp ---------------------------
__device__ __forceinline__ void ws_write_empty(const RansacArgs &a, long long s, int status, int lane)
{
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    if (lane < 8) a.line[s * 8 + lane] = qnan;
    if (lane == 0) {
        a.ninl[s] = status;
        a.win_trial[s] = -1;
        a.win_pair[2 * s] = -1;
        a.win_pair[2 * s + 1] = -1;
        a.ssum[s] = qnan;
    }
    for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = 0u;
}

__device__ __forceinline__ void ws_post_chunk(const RansacArgs &a, long long s, long long gp0, const double2 *pts,
                                              double *sq, uint32_t *cmq, WsPostScratch *sc, uint32_t *smask,
                                              uint32_t *wtmp, int lane)
{
    const int T = a.trials;
    const WsSlot g = ws_slot(a, s);
    if (!g.valid) { ws_write_empty(a, s, 0, lane); return; }                 // unused slot
    const int n = g.n;
    const double2 *P = pts + ((long long)g.b * a.N + (long long)g.c * a.chunk - gp0);
    const uint64_t scan = a.scan_base + (uint64_t)g.b;
    // ---- scale of the chunk -> re-check band ----
    int mh = 0;
    for (int k = lane; k < n; k += 32) {
        const double2 p = P[k];
        const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
        mh = mh > hx ? mh : hx;
        mh = mh > hy ? mh : hy;
    }
    mh = __reduce_max_sync(LSLAM_FULL_MASK, mh);
    const ChunkBand cb = chunk_band(mh);
    if (cb.bad) { ws_write_empty(a, s, kStatusNonFinite, lane); return; }
    int *cq = reinterpret_cast<int *>(cmq);
    // ---- flagged trials: literal re-score; unpack the counts; maximal count ----
    int mx = 0;
    unsigned n_relit = 0;
    for (int t0 = 0; t0 < T; t0 += 32) {
        const int t = t0 + lane;
        const uint32_t w = t < T ? cmq[t] : 0x7fff0000u;
        int cv = (int)(w & 0xffffu);
        const bool flag = t < T && (cb.force_literal || (int)(w & 0x7fff0000u) < cb.band_hi);
        unsigned m = __ballot_sync(LSLAM_FULL_MASK, flag);
        while (m) {
            const int tt = t0 + __ffs((int)m) - 1;
            m &= m - 1;
            int i, j, cl;
            double sl;
            sample_pair(a.seed, scan, (uint32_t)g.c, (uint32_t)tt, (uint32_t)n, i, j);
            warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, nullptr);
            if (tt == t) { cv = cl | kCntExact; sq[t] = sl; }
            ++n_relit;
        }
        if (t < T) cq[t] = cv;
        const int v = cv & ~kCntExact;
        mx = mx > v ? mx : v;
    }
    mx = __reduce_max_sync(LSLAM_FULL_MASK, mx);
    if (lane == 0 && n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);
    __syncwarp();
    if (mx < 2) { ws_write_empty(a, s, kStatusNoModel, lane); return; }     // upstream's final estimate() needs 2 points
    // ---- smallest upper bound of the literal residual sum among the tied trials ----
    const SsumBound sb = ssum_bound_coeffs(cb.eabs, n, a.thr);
    int minhi_q = 0x7fffffff;
    for (int t = lane; t < T; t += 32) {
        const int cv = cq[t];
        if ((cv & ~kCntExact) == mx) {
            const double v = sq[t];
            const double e = (cv & kCntExact) ? 0.0 : fma(sb.beta, v, sb.alpha);
            const int k = key_up(v + e);
            minhi_q = minhi_q < k ? minhi_q : k;
        }
    }
    minhi_q = __reduce_min_sync(LSLAM_FULL_MASK, minhi_q);
    // ---- candidate list in trial order ----
    int ncand = 0;
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
    __syncwarp();
    // ---- candidate models: one lane each ----
    if (lane < ncand && lane < kMaxCand) {
        int i, j;
        sample_pair(a.seed, scan, (uint32_t)g.c, (uint32_t)sc->cand_t[lane], (uint32_t)n, i, j);
        sc->cand_i[lane] = i;
        sc->cand_j[lane] = j;
        sc->cmodel[lane] = lit_pair_model(P[i], P[j]);
    }
    __syncwarp();
    // ---- literal tie-break in trial order (first best wins, like the sequential accept rule) ----
    bool fault = false;
    int best_t = -1, best_i = -1, best_j = -1;
    double best_ss = __longlong_as_double(0x7ff0000000000000LL);
    if (ncand <= kMaxCand) {
        for (int ci = 0; ci < ncand; ++ci) {
            int cl;
            double sl;
            warp_literal_score(P, n, sc->cmodel[ci], a.thr, lane, cl, sl, ncand == 1 ? smask : wtmp);
            __syncwarp();
            if (cl != mx) fault = true;
            if (sl < best_ss) {
                best_ss = sl; best_t = sc->cand_t[ci]; best_i = sc->cand_i[ci]; best_j = sc->cand_j[ci];
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
                sample_pair(a.seed, scan, (uint32_t)g.c, (uint32_t)tt, (uint32_t)n, i, j);
                warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, wtmp);
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
            sample_pair(a.seed, scan, (uint32_t)g.c, (uint32_t)t, (uint32_t)n, i, j);
            warp_literal_score(P, n, lit_pair_model(P[i], P[j]), a.thr, lane, cl, sl, wtmp);
            __syncwarp();
            if (cl > best_n || (cl == best_n && sl < best_ss)) {
                best_n = cl; best_ss = sl; best_t = t; best_i = i; best_j = j;
                for (int w = lane; w * 32 < n; w += 32) smask[w] = wtmp[w];
            }
            __syncwarp();
        }
        mx = best_n;
        if (best_t < 0 || best_n < 2) { ws_write_empty(a, s, kStatusNoModel, lane); return; }
    }
    // ---- refit on the inliers (principal axis of the 2x2 scatter, relative to the first sample) ----
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
        const double m = (double)mx;
        const double mdx = sx / m, mdy = sy / m;
        const double ox = p0.x + mdx, oy = p0.y + mdy;
        double dx, dy;
        principal_axis(fma(-sx, mdx, sxx), fma(-sx, mdy, sxy), fma(-sy, mdy, syy), dx, dy);
        int last = 0;
        for (int w = (n - 1) >> 5; w >= 0; --w) {
            const unsigned mw = smask[w];
            if (mw) { last = w * 32 + 31 - __clz((int)mw); break; }
        }
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
    for (int w = lane; w < a.W; w += 32) a.mask[s * a.W + w] = (w * 32 < n) ? smask[w] : 0u;
    __syncwarp();
}

// ---- the kernel: one persistent CTA per SM, (NS + NP + 1) warps -----------------------------------
// a.G = chunk slots per tile, a.ws_ns / a.ws_np / a.ws_nb = scorer warps, post warps, ring depth.
template <int TPT>
__global__ void __launch_bounds__(640, 1) ransac_ws_kernel(const RansacArgs a)
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
            mbar_init(&ctl[b].scored, (unsigned)NS);
            mbar_init(&ctl[b].freeb, (unsigned)NP);
            ctl[b].next_round = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();

    // first staged point and byte count of tile `tile`
    auto tile_span = [&](long long tile, long long &slot0, int &nsl, long long &gp0, unsigned &bytes) {
        slot0 = tile * G;
        const long long rem = a.nslots - slot0;
        nsl = rem < (long long)G ? (int)rem : G;
        const long long b0 = slot0 / a.C, c0 = slot0 - b0 * a.C;
        gp0 = b0 * a.N + c0 * a.chunk;
        const long long sl = slot0 + nsl - 1;
        const long long bl = sl / a.C, cl = sl - bl * a.C;
        const long long cend = (cl + 1) * a.chunk < a.N ? (cl + 1) * a.chunk : a.N;
        bytes = (unsigned)((bl * a.N + cend - gp0) * 16);
    };

    if (warp == NS + NP) {
        // ================= producer =================
        if (lane == 0) {
            long long k = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
                const int b = (int)(k % NB);
                const long long j = k / NB;
                if (j > 0) mbar_wait(&ctl[b].freeb, (unsigned)((j - 1) & 1));
                long long slot0, gp0;
                int nsl;
                unsigned bytes;
                tile_span(tile, slot0, nsl, gp0, bytes);
                ctl[b].next_round = 0;
                mbar_expect_tx(&ctl[b].full, bytes);
                bulk_g2s(smem_raw + L.bufs + (size_t)b * L.buf_stride, a.xy + gp0, bytes, &ctl[b].full);
            }
        }
    } else if (warp < NS) {
        // ================= scorers =================
        long long k = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            long long slot0, gp0;
            int nsl;
            unsigned bytes;
            tile_span(tile, slot0, nsl, gp0, bytes);
            const int items = nsl * a.ipc, nrounds = (items + 31) >> 5;
            mbar_wait(&ctl[b].full, (unsigned)((k / NB) & 1));
            for (;;) {
                int r = 0;
                if (lane == 0) r = atomicAdd(&ctl[b].next_round, 1);
                r = __shfl_sync(LSLAM_FULL_MASK, r, 0);
                if (r >= nrounds) break;
                const int item = r * 32 + lane;
                if (item < items)
                    ws_score_item<TPT>(a, slot0, gp0, item, reinterpret_cast<const double2 *>(buf),
                                       reinterpret_cast<double *>(buf + L.o_ssf), reinterpret_cast<uint32_t *>(buf + L.o_cm));
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
        long long k = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++k) {
            const int b = (int)(k % NB);
            unsigned char *buf = smem_raw + L.bufs + (size_t)b * L.buf_stride;
            long long slot0, gp0;
            int nsl;
            unsigned bytes;
            tile_span(tile, slot0, nsl, gp0, bytes);
            mbar_wait(&ctl[b].scored, (unsigned)((k / NB) & 1));
            for (int q = pw; q < nsl; q += NP)
                ws_post_chunk(a, slot0 + q, gp0, reinterpret_cast<const double2 *>(buf),
                              reinterpret_cast<double *>(buf + L.o_ssf) + (size_t)q * T,
                              reinterpret_cast<uint32_t *>(buf + L.o_cm) + (size_t)q * T, sc, smask, wtmp, lane);
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl[b].freeb);
        }
    }
}

} // namespace lslam
