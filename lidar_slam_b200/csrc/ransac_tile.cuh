// ransac_tile.cuh - K3: batched 2-point RANSAC for SMALL chunks (the reference's own 100 points x 100 trials),
// one launch.  Replaces, per chunk, /root/reference/ransac_functions.py:23-30 (skimage.measure.ransac with
// LineModelND, then slope/intercept/tip).  Semantics = oracle/ransac_line.py; every integer output and the residual
// sum are bit-exact, the refit is held to 1e-5.
//
// Round-2 redesign of ransac.cuh for the shape where per-chunk overhead, not the scoring loop, was the loss
// (profiles/r03a_ransac_c2_ncu_summary.txt: seven block barriers, hot loop only 59.5 % of the instructions):
//
//   * hot loop = 2 FFMA + FADD + 1 SHF per point-hypothesis, all fp32 (round 1: 3 DFMA + DADD + SHF + FMNMX/2): the
//     signed distance is evaluated in fp32 on the chunk's points centred on its middle point.  |s_f - s| <= 7 u R
//     (u = 2^-24, R = the chunk's radius about that point; derivation below at tile_screen32), so a point with
//     | |s_f| - thr_f | inside delta + u thr cannot be decided in fp32: its trial is flagged and re-scored in fp64; every
//     output stays bit-exact.  Line and threshold are scaled by a power of two such that "outside that band" is one
//     exponent bit of the difference: a funnel shift by TWO collects the inlier (sign) and the far-from-threshold bit
//     of each point - no running minimum of the distance to the threshold;
//   * the fast residual sum (only used to rank trials tied on the count, inside a rigorous error bound) is NOT
//     accumulated per point: sum_k s_k^2 is a quadratic form in the line coefficients whose matrix is the
//     chunk's second-moment matrix about one of its own points (tile_moment_sum, bound in tile_moment_err);
//   * every trial's inlier bits are kept in shared memory (one STS per 32 points), so the winner's mask is read,
//     not rebuilt;
//   * one work item (TPT trials of one chunk) per thread, per-trial results stay in registers; the chunk-wide
//     maximum (count, then smallest upper bound of the residual sum) is a max-scan over the slot's lanes and one 64-bit
//     shared-memory atomicMax by its last lane; trials with a point too close to the threshold are re-scored
//     warp-cooperatively on the spot, with the fp64 fast line first and the literal formula only if that is too close as well;
//   * three block barriers per tile (stage+moments | score | candidates | decide+refit+stores by lane groups).
#pragma once
#include "ransac.cuh"

namespace lslam {

constexpr int kTileLG = 8;             // lanes that stage / refit one chunk slot (8, 16 or 32): the narrow phases are latency-
                                       // bound, so more lanes per slot (fewer points per lane, more warps busy) shorten them

struct TileSlot {
    unsigned long long scan;           // global scan index (Philox counter words 2, 3)
    long long gslot;                   // global chunk slot index (explicit sample pairs)
    unsigned long long key;            // max over trials of (count << 32) | (0x7fffffff - key_up(upper bound of sum r^2))
    double c0x, c0y;                   // centre of the moments: the chunk's middle point
    double Sx, Sy, Sxx, Sxy, Syy;      // moments of q = fl(p - c0)
    double best_ss;
    double alpha, beta;                // bound on |sum r^2 - sum s^2| = alpha + beta * sum s^2 (ssum_bound_coeffs)
    double A, nn;                      // Sxx + Syy and n as a double
    double kb;                         // (1 + beta) x the factor of tile_moment_err: the moment form's own error inside the bound
    float bandf, kscale;               // fp32 screening: band of | |s_f| - thr_f | that needs a re-score; 2 / band' (tile_screen32)
    int off, n, valid, c;
    int scale_hi;                      // high word of the largest |coordinate| of the slot
    int force_lit;                     // every trial is scored literally (threshold / scale outside the fast path's range)
    int mx, ncand, status;
    int band_hi;                       // high word of the fp64 re-check band 2^(e2-36) (chunk_band)
    int best_t, best_i, best_j;
    int cand_exact;                    // bit p: candidate p was re-scored literally (its count / mask are the literal ones)
    int cand_t[kMaxCand], cand_i[kMaxCand], cand_j[kMaxCand];
};

// Shared memory of a CTA: two regions of equal size.  Region p holds a tile's points (fp64, then the centred fp32 copies),
// region p ^ 1 its per-trial scratch (tie-break keys, inlier words of every trial); the winner's inlier words are copied out
// of the scratch into `smask` when the candidates are collected.  The phase functions below take the parity p so that a CTA
// can walk several tiles with the roles swapping (stage tile i + 1 into the dead scratch of tile i while tile i is refitted);
// the kernel as shipped runs one tile per CTA with p = 0 - the persistent form measured slower.
struct TileSmem {
    unsigned char *region0;    // region p at region0 + p * rbytes (no arrays here: an index known at run time would put them in local memory)
    size_t rbytes;
    uint32_t *smask;    // [G][W] the winner's inlier words
    uint32_t *wtmp;     // [nwarps][W]
    unsigned char *slot0;      // slot state of the tile whose points are in region p: at slot0 + p * sbytes
    size_t sbytes;
    __device__ __forceinline__ unsigned char *region(int p) const { return region0 + (size_t)p * rbytes; }
    __device__ __forceinline__ TileSlot *slot(int p) const { return reinterpret_cast<TileSlot *>(slot0 + (size_t)p * sbytes); }
    size_t ptsf_off, mw_off;   // byte offsets inside a region: points | fp32 copies,  keys | trial masks
};

__host__ __device__ inline size_t tile_region_bytes(int G, int chunk, int ipc, int tpt, int W)
{
    size_t pts = (size_t)G * chunk * 16;
    pts += (size_t)G * ((chunk + 1) & ~1) * 8;
    const size_t scr = (size_t)G * tpt * ipc * 8 + (size_t)G * W * tpt * ipc * 4;
    return align16(pts > scr ? pts : scr);
}

__host__ __device__ inline size_t tile_smem_bytes(int G, int chunk, int ipc, int tpt, int W, int nwarps)
{
    size_t b = 2 * tile_region_bytes(G, chunk, ipc, tpt, W);
    b += align16((size_t)G * W * 4);
    b += align16((size_t)nwarps * W * 4);
    b += 2 * align16((size_t)G * sizeof(TileSlot));
    return b;
}

__device__ __forceinline__ TileSmem tile_carve(unsigned char *raw, int G, int chunk, int ipc, int tpt, int W, int nwarps)
{
    TileSmem s;
    unsigned char *p = raw;
    const size_t rb = tile_region_bytes(G, chunk, ipc, tpt, W);
    s.region0 = p; s.rbytes = rb; p += 2 * rb;
    s.ptsf_off = (size_t)G * chunk * 16;
    s.mw_off = (size_t)G * tpt * ipc * 8;
    s.smask = reinterpret_cast<uint32_t *>(p); p += align16((size_t)G * W * 4);
    s.wtmp = reinterpret_cast<uint32_t *>(p);  p += align16((size_t)nwarps * W * 4);
    s.slot0 = p; s.sbytes = align16((size_t)G * sizeof(TileSlot));
    return s;
}

// ---- fp32 screening ---------------------------------------------------------------------------------------------
// Points centred on a point c0 of the chunk and rounded, q = fl32(p - c0) (difference formed in fp64); per trial the
// unit normal (nx, ny) and cp = nx c0x + ny c0y + ng (the line's offset relative to c0, |cp| <= R because the line
// passes through two points of the chunk) rounded from their fp64 values; s_f = fma(q_y, n_y, fma(q_x, n_x, cp)).  With
// R = max |p - c0|_2 over the chunk and u = 2^-24:
//   q rounding <= u R,  n rounding <= u R,  cp rounding <= u R,  inner fma <= 2 u R,  outer fma <= u |s_f| <= 2 u R
// so |s_f - s| <= 7 u R, and the fp64 quantities (normal from a Newton-refined rsqrt, cp, the literal residual itself)
// add O(2^-40 R):  delta = 4.25e-7 R (7 u = 4.172e-7) bounds | |s_f| - r_literal |.  The test runs on f = |s_f| - thr_f,
// thr_f = fl32(thr) (|thr_f - thr| <= u thr; the subtraction is exact near zero and otherwise off by a factor 1 + u):
//   |f| >= bandf = (delta + u thr)(1 + 2^-20)  ==>  sign(f) is the literal decision r < thr.
// The fp64 side (the Newton-refined normal, cp, and the literal residual itself, whose origin fl((p_i + p_j) / 2) is rounded at
// the chunk's coordinate scale S) adds at most ~2^-51 S; the slack between delta and 7 u R is 2^-26.9 R, so the screening is
// only used while S <= 2^22 R (a cluster of points 4 million radii away from the origin is scored literally).
// The tile kernel does not compare |f| with bandf point by point: it scales the line and the threshold by kscale = 2 / band',
// band' = the power of two >= bandf, so that the scaled f = kscale (|s_f| - thr_f) - bit for bit the unscaled one times a power
// of two - has |f| >= band' exactly when its exponent bit 30 is set.  One funnel shift by TWO then takes both the sign (inlier)
// and that bit (far from the threshold) of every point.  band' <= 1 is required (kscale >= 2: scaling up cannot underflow).
struct TileScreen32 { float bandf; float kscale; bool force_literal; };
__device__ __forceinline__ float tile_sqrt_up(double r2)
{
#ifndef LSLAM_EMU
    return __fsqrt_ru(__double2float_ru(r2));
#else
    const float f = __double2float_ru(r2);
    float s = sqrtf(f);
    while ((double)s * (double)s < (double)f) s = nextafterf(s, INFINITY);
    return s;
#endif
}
__device__ __forceinline__ TileScreen32 tile_screen32(double R, double thr, double S)
{
    TileScreen32 b;
    const double band = (4.25e-7 * R + 0x1p-24 * thr) * (1.0 + 0x1p-20);
    // outside these ranges every trial of the chunk is scored literally: fp32 range of the products, a threshold the
    // band is not small against, thr <= 0 or NaN, R not finite, a radius tiny against the coordinates
    b.force_literal = !(R >= 0x1p-40 && R <= 0x1p40 && thr >= 0x1p-40 && thr <= 0x1p40 && 4.0 * band <= thr && S <= 0x1p22 * R);
    b.bandf = b.force_literal ? 0.0f : __double2float_ru(band);
    b.kscale = 0.0f;
    if (!b.force_literal) {
        const int bits = __float_as_int(b.bandf);                                  // positive and normal here
        const int e = ((bits >> 23) & 0xff) + ((bits & 0x7fffff) ? 1 : 0);          // biased exponent of band'
        if (e > 127) b.force_literal = true;                                        // band' > 1: a chunk kilometres wide
        else b.kscale = __int_as_float((254 - e + 1) << 23);                        // 2 / band' = 2^(1 - (e - 127))
    }
    return b;
}

// the line of a trial relative to the slot's centre c0:  s(p) = n . (p - c0) + cp,  cp = n . (c0 - p_i).  The unit normal
// comes from the fp64 difference p_j - p_i (Newton-refined rsqrt, relative error < 2^-42); cp is formed from the differences
// c0 - p_i (no cancellation at the coordinate scale).  Rounded once to fp32 they are the hot loop's line: the "n rounding"
// and "cp rounding" terms u R of the bound above.  degenerate: coincident samples or an absurd scale -> literal score.
struct TileLine { double nx, ny, cp; bool degenerate; };
__device__ __forceinline__ TileLine tile_line(double2 pi, double2 pj, double c0x, double c0y)
{
    const double dx = pj.x - pi.x, dy = pj.y - pi.y;
    const double n2 = fma(dx, dx, dy * dy);
    TileLine l;
    l.degenerate = !(n2 > 0x1p-100 && n2 < 0x1p100);     // also keeps (float)n2 in range
    const double n2s = l.degenerate ? 1.0 : n2;
    double y = (double)rsqrtf((float)n2s);
    y = y * fma(-0.5 * n2s * y, y, 1.5);
    l.nx = -dy * y; l.ny = dx * y;
    l.cp = fma(l.nx, c0x - pi.x, l.ny * (c0y - pi.y));
    return l;
}

// ---- residual sum from moments -------------------------------------------------------------------------------
// For the line (nx, ny, ng) and the centred points q_k = fl(p_k - c0):  sigma_k = nx qx_k + ny qy_k + cp with
// cp = fl(nx c0x + ny c0y + ng) differs from the hot loop's s_k by < 2^-50 S (roundings of q, cp and the two FMAs),
// far inside the per-point bound eabs = 2^-36 S that ssum_bound_coeffs() already allows between |s_k| and the
// literal residual.  sum_k sigma_k^2 = nx^2 Sxx + 2 nx ny Sxy + ny^2 Syy + 2 cp (nx Sx + ny Sy) + n cp^2 exactly; the
// evaluation below (moments summed with FMAs in any order, ~12 roundings here) is off by at most
//   (2.5 n + 10) eps (A + n cp^2),  A = Sxx + Syy,  eps = 2^-53
// (each moment carries <= n eps of its absolute sum; Cauchy-Schwarz and AM-GM fold the mixed terms into A + n cp^2).
// tile_moment_err() returns twice that with n + 32 for n.
__device__ __forceinline__ double tile_moment_sum(double nx, double ny, double cp, double nn, const TileSlot *st)
{
    const double a1 = fma(nx, st->Sxx, ny * st->Sxy), a2 = fma(nx, st->Sxy, ny * st->Syy);
    const double quad = fma(nx, a1, ny * a2), lin = fma(nx, st->Sx, ny * st->Sy);
    return fma(cp, fma(nn, cp, 2.0 * lin), quad);
}
__device__ __forceinline__ double tile_moment_err_factor(double nn) { return (5.0 * (nn + 32.0) + 20.0) * 0x1p-53 * 1.01; }
__device__ __forceinline__ double tile_moment_err(double cp, double nn, double A)
{
    return tile_moment_err_factor(nn) * fma(nn * cp, cp, A);
}

// score of one hypothesis by a whole warp with the fp64 fast line (lslam_core.cuh FastLine: |s| is within 2^-40 S of the
// literal residual, the re-check band 2^(e2-36) is sixteen times that): the count and the mask are the literal ones unless
// `near` comes back set (a point inside the band, or a non-finite distance) - then the caller scores literally
__device__ __forceinline__ void warp_fast64_score_strided(const double2 *P, int n, const FastLine m, double thr, int band_hi, int lane,
                                                          int &cnt, bool &near, uint32_t *mask_out, int stride)
{
    int c = 0;
    bool nr = false;
    for (int k0 = 0; k0 < n; k0 += 32) {
        const int k = k0 + lane;
        bool in = false;
        if (k < n) {
            const double2 p = P[k];
            const double u = fabs(fma(p.y, m.ny, fma(p.x, m.nx, m.negc))) - thr;
            in = u < 0.0;
            nr = nr || (__double2hiint(u) & 0x7fffffff) < band_hi || !(fabs(u) <= 1.7976931348623157e308);
        }
        const unsigned b = __ballot_sync(LSLAM_FULL_MASK, in);
        c += __popc(b);
        if (mask_out != nullptr && lane == 0) mask_out[(k0 >> 5) * stride] = __brev(b);   // the per-trial words are kept bit-reversed (tile_score)
    }
    cnt = c;
    near = __any_sync(LSLAM_FULL_MASK, nr);
}

// literal score of one hypothesis by a whole warp, mask words stored with a stride
__device__ __forceinline__ void warp_literal_score_strided(const double2 *P, int n, const LitLine m, double thr, int lane,
                                                           int &cnt, double &ssum, uint32_t *mask_out, int stride)
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
        if (mask_out != nullptr && lane == 0) mask_out[(k0 >> 5) * stride] = __brev(b);   // bit-reversed like the fast words
    }
    cnt = c;
    ssum = warp_tree32(part);
}

// ================= stage + moments of one tile: LG lanes per slot, warps [w0, w0 + nw) ====================
template <int TPT>
__device__ __forceinline__ void tile_stage(const RansacArgs &a, int tile, int par, int w0, int nw)
{
    LSLAM_SMEM_DECL;                                              // the CTA's dynamic shared memory, known to be shared memory here
    constexpr int LG = kTileLG;
    constexpr int SPW = 32 / LG;
    const int G = a.G, chunk = a.chunk, W = a.W, ipc = a.ipc;
    const int chunk_pad = (chunk + 1) & ~1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const TileSmem sm = tile_carve(smem_raw, G, chunk, ipc, TPT, W, nwarps);
    (void)chunk_pad; (void)tid; (void)nwarps;

    const long long slot0 = (long long)tile * G;
    const long long rem_slots = a.nslots - slot0;
    const int nsl = rem_slots < (long long)G ? (int)rem_slots : G;
    double2 *pts = reinterpret_cast<double2 *>(sm.region(par));
    float2 *ptsf = reinterpret_cast<float2 *>(sm.region(par) + sm.ptsf_off);
    const int wrel = warp - w0;
    const int grp = wrel * SPW + lane / LG, r = lane % LG, ngrp = nw * SPW;
    // CTA geometry (32-bit division whenever the slot index fits)
    long long b0;
    int c0;
    if (slot0 <= 0x7fffffffLL) {
        const unsigned s = (unsigned)slot0, bq = s / (unsigned)a.C;
        b0 = bq; c0 = (int)(s - bq * (unsigned)a.C);
    } else {
        b0 = slot0 / a.C; c0 = (int)(slot0 - b0 * a.C);
    }
    const long long gp0 = b0 * a.N + (long long)c0 * chunk;       // first staged point
    for (int qb = 0; qb < nsl; qb += ngrp) {                  // uniform trip count: the shuffles below are warp-wide
        if (qb + wrel * SPW >= nsl) break;                    // (warp-uniform) none of this warp's groups has a slot
        const int q = qb + grp;
        const bool act = q < nsl;
        int n = 0, off = 0, c = 0;
        long long b = b0;
        if (act) {
            const int cq = c0 + q, bb = cq / a.C;
            c = cq - bb * a.C; b = b0 + bb;
            int nb = a.npts ? a.npts[b] : a.N;
            nb = nb < 0 ? 0 : (nb > a.N ? a.N : nb);
            n = nb - c * chunk;
            n = n < 0 ? 0 : (n > chunk ? chunk : n);
            off = (int)((long long)bb * a.N + (long long)(c - c0) * chunk);
        }
        const double2 *src = a.xy + gp0 + off;
        double2 *dst = pts + off;
        // asynchronous copies (no register staging: all of a lane's loads are in flight at once), then one pass
        // over the staged points for the slot's scale and its moments about its middle point
        for (int k = r; k < n; k += LG) cp_async16(dst + k, src + k);
        cp_async_wait_all();
        __syncwarp();
        // functions.py:64-76: full chunks, or a remainder of MORE than 2 points
        const bool valid = (n >= 2) && (n == chunk || n > 2);
        const double2 cc = n > 0 ? dst[n >> 1] : make_double2(0.0, 0.0);
        float2 *dstf = ptsf + (size_t)(act ? q : 0) * chunk_pad;
        int mh = 0;
        double sx = 0.0, sy = 0.0, sxx = 0.0, sxy = 0.0, syy = 0.0, r2 = 0.0;
        for (int k = r; k < n; k += LG) {
            const double2 p = dst[k];
            const int hx = __double2hiint(p.x) & 0x7fffffff, hy = __double2hiint(p.y) & 0x7fffffff;
            mh = mh > hx ? mh : hx;
            mh = mh > hy ? mh : hy;
            const double qx = p.x - cc.x, qy = p.y - cc.y;
            sx += qx; sy += qy;
            sxx = fma(qx, qx, sxx); sxy = fma(qx, qy, sxy); syy = fma(qy, qy, syy);
            dstf[k] = make_float2((float)qx, (float)qy);
            r2 = fmax(r2, fma(qx, qx, qy * qy));
        }
#pragma unroll
        for (int o = LG / 2; o >= 1; o >>= 1) { const int v = __shfl_xor_sync(LSLAM_FULL_MASK, mh, o); mh = mh > v ? mh : v; }
        const ChunkBand cb = chunk_band(mh);
#pragma unroll
        for (int o = LG / 2; o >= 1; o >>= 1) {
            sx += __shfl_xor_sync(LSLAM_FULL_MASK, sx, o);
            sy += __shfl_xor_sync(LSLAM_FULL_MASK, sy, o);
            sxx += __shfl_xor_sync(LSLAM_FULL_MASK, sxx, o);
            sxy += __shfl_xor_sync(LSLAM_FULL_MASK, sxy, o);
            syy += __shfl_xor_sync(LSLAM_FULL_MASK, syy, o);
            r2 = fmax(r2, __shfl_xor_sync(LSLAM_FULL_MASK, r2, o));
        }
        if (act && r == 0) {
            TileSlot *st = sm.slot(par) + q;
            st->scan = a.scan_base + (uint64_t)b;
            st->gslot = slot0 + q;
            st->key = 0ull;
            st->c0x = cc.x; st->c0y = cc.y;
            st->Sx = sx; st->Sy = sy; st->Sxx = sxx; st->Sxy = sxy; st->Syy = syy;
            st->best_ss = 0.0;
            const SsumBound sbq = ssum_bound_coeffs_inv(cb.eabs, n, a.thr, a.inv_thr);
            st->alpha = sbq.alpha; st->beta = sbq.beta; st->A = sxx + syy; st->nn = (double)n;
            st->kb = (1.0 + sbq.beta) * tile_moment_err_factor((double)n);
            st->off = off; st->n = n; st->valid = valid ? 1 : 0; st->c = c;
            st->scale_hi = mh;
            // an upper bound of the radius is enough: fp32 square root rounded up (overflow -> inf, underflow -> 0:
            // both outside the screening range, the slot is then scored literally)
            const TileScreen32 sc = tile_screen32((double)tile_sqrt_up(r2), a.thr, cb.eabs * 0x1p36);
            st->bandf = sc.bandf; st->kscale = sc.kscale;
            st->force_lit = (cb.force_literal || sc.force_literal) ? 1 : 0;
            st->mx = 0; st->ncand = 0; st->band_hi = cb.band_hi;
            st->status = !valid ? 0x7fffffff : (cb.bad ? kStatusNonFinite : 0);   // 0x7fffffff: unused slot
            st->best_t = -1; st->best_i = -1; st->best_j = -1;
            st->cand_exact = 0;
        }
    }
}

// ================= decide (only where the literal formula is needed) + refit + stores of one tile: LG lanes per
// slot, warps [0, nw).  Reads the tile's points, its slots and the winner masks - not its scratch region. ======
template <int TPT>
__device__ __forceinline__ void tile_refit(const RansacArgs &a, int tile, int par, int nw)
{
    LSLAM_SMEM_DECL;
    constexpr int LG = kTileLG;
    constexpr int SPW = 32 / LG;
    const int G = a.G, chunk = a.chunk, W = a.W, ipc = a.ipc;
    const int chunk_pad = (chunk + 1) & ~1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const TileSmem sm = tile_carve(smem_raw, G, chunk, ipc, TPT, W, nwarps);
    (void)chunk_pad; (void)tid; (void)nwarps;

    const long long slot0 = (long long)tile * G;
    const long long rem_slots = a.nslots - slot0;
    const int nsl = rem_slots < (long long)G ? (int)rem_slots : G;
    const double2 *pts = reinterpret_cast<const double2 *>(sm.region(par));
    TileSlot *slots = sm.slot(par);
    for (int qw = warp * SPW; qw < nsl; qw += nw * SPW) {
        // ---- the common case, all slots of the warp at once: one candidate and no residual sum wanted - it is the winner ----
        unsigned hard;                                            // bit i: slot qw + i needs the literal tie-break
        {
            const int qd = qw + lane / LG;
            bool need = false;
            if (lane % LG == 0 && qd < nsl) {
                TileSlot *s1 = slots + qd;
                if (s1->status == 0) {
                    if (s1->ncand == 1 && a.ssum == nullptr) {
                        s1->best_t = s1->cand_t[0]; s1->best_i = s1->cand_i[0]; s1->best_j = s1->cand_j[0];
                        s1->best_ss = __longlong_as_double(0x7ff8000000000000LL);
                    } else need = true;
                }
            }
            hard = __ballot_sync(LSLAM_FULL_MASK, need);
        }
        // ---- literal tie-break, a warp per slot: several candidates, or the caller wants the residual sum ----
        while (hard) {
            const int qq = qw + (__ffs((int)hard) - 1) / LG;
            hard &= hard - 1;
            TileSlot *s2 = slots + qq;
            const int ncand = s2->ncand;
            const double2 *P2 = pts + s2->off;
            uint32_t *wt = sm.wtmp + (size_t)warp * W;
            uint32_t *mq = sm.smask + qq * W;
            bool fault = false;
            int best_t = -1, best_i = -1, best_j = -1;
            double best_ss = __longlong_as_double(0x7ff0000000000000LL);
            if (ncand <= kMaxCand) {
                // candidates were appended in no particular order: smallest sum wins, the earlier trial on a tie -
                // the sequential accept rule (first best in trial order)
                for (int ci = 0; ci < ncand; ++ci) {
                    int cl;
                    double sl;
                    const int ti = s2->cand_t[ci], ii = s2->cand_i[ci], jj = s2->cand_j[ci];
                    warp_literal_score(P2, s2->n, lit_pair_model(P2[ii], P2[jj]), a.thr, lane, cl, sl, wt);
                    __syncwarp();
                    if (cl != s2->mx) fault = true;
                    if (sl < best_ss || (sl == best_ss && ti < best_t)) {
                        best_ss = sl; best_t = ti; best_i = ii; best_j = jj;
                        for (int w = lane; w * 32 < s2->n; w += 32) mq[w] = wt[w];
                    }
                    __syncwarp();
                }
                if (lane == 0 && ncand > 1) atomicAdd(a.counters + 1, (unsigned long long)(ncand - 1));
            }
            if (lane == 0) { s2->best_t = best_t; s2->best_i = best_i; s2->best_j = best_j; s2->best_ss = best_ss; }
            __syncwarp();
            if (fault || best_t < 0) {
                // more candidates than the list holds (e.g. noise-free collinear data), or a fast count that disagrees
                // with the literal one (error bound violated - never observed, counted): the exact sequential rule
                if (lane == 0 && fault) atomicAdd(a.counters + 2, 1ULL);
                slot_exact_fallback(a, s2, P2, mq, wt, lane);
            }
        }
        __syncwarp();

        // ---- refit + outputs: LineModelND.estimate(data[inliers]) as the principal axis of the 2x2 scatter,
        //      accumulated relative to the winner's first sample point (an inlier, so the shift is benign) ----
        const int qr = qw + lane / LG, r = lane % LG;
        const bool act = qr < nsl;
        const TileSlot *s3 = slots + (act ? qr : qw);
        const bool ok = act && s3->status == 0;
        const double2 *P3 = pts + s3->off;
        const uint32_t *mrow = sm.smask + (act ? qr : qw) * W;
        const int n3 = s3->n;
        double sx = 0.0, sy = 0.0, sxx = 0.0, sxy = 0.0, syy = 0.0;
        double2 p0 = make_double2(0.0, 0.0);
        // independent of the sums below, so started first: 1 / count, and the tip (the LAST inlier in data order)
        double rm = 0.0, tipx = 0.0;
        if (ok) {
            rm = 1.0 / (double)s3->mx;
            int last = 0;
            for (int w = (n3 - 1) >> 5; w >= 0; --w) {
                const unsigned mw_ = mrow[w];
                if (mw_) { last = w * 32 + 31 - __clz((int)mw_); break; }
            }
            tipx = P3[last].x;
            p0 = P3[s3->best_i];
            // loads are unconditional and independent of the mask test (they pipeline); an outlier contributes zeros
#pragma unroll 4
            for (int k = r; k < n3; k += LG) {
                const double2 p = P3[k];
                const bool in = (mrow[k >> 5] >> (k & 31)) & 1u;
                const double dx = in ? p.x - p0.x : 0.0, dy = in ? p.y - p0.y : 0.0;
                sx += dx; sy += dy;
                sxx = fma(dx, dx, sxx); sxy = fma(dx, dy, sxy); syy = fma(dy, dy, syy);
            }
        }
#pragma unroll
        for (int o = LG / 2; o >= 1; o >>= 1) {
            sx += __shfl_xor_sync(LSLAM_FULL_MASK, sx, o);
            sy += __shfl_xor_sync(LSLAM_FULL_MASK, sy, o);
            sxx += __shfl_xor_sync(LSLAM_FULL_MASK, sxx, o);
            sxy += __shfl_xor_sync(LSLAM_FULL_MASK, sxy, o);
            syy += __shfl_xor_sync(LSLAM_FULL_MASK, syy, o);
        }
        if (act) {
            const long long s = slot0 + qr;
            const double qnan = __longlong_as_double(0x7ff8000000000000LL);
            double out = qnan;                                   // lane r stores element r of the line record
            if (ok) {
                const double mdx = sx * rm, mdy = sy * rm;
                const double ox = p0.x + mdx, oy = p0.y + mdy;
                double dx, dy;
                principal_axis_rsqrt(fma(-sx, mdx, sxx), fma(-sx, mdy, sxy), fma(-sy, mdy, syy), dx, dy);
                // ransac_functions.py:26-30 (no guard for dir_x == 0, like the reference)
                const double sl = __ddiv_rn(dy, dx);
                const double ic = __dsub_rn(oy, __dmul_rn(sl, ox));
                const double tipy = __dadd_rn(__dmul_rn(tipx, sl), ic);
                out = (r & 7) == 0 ? sl : (r & 7) == 1 ? ic : (r & 7) == 2 ? ox : (r & 7) == 3 ? oy : (r & 7) == 4 ? dx : (r & 7) == 5 ? dy : (r & 7) == 6 ? tipx : tipy;
            }
            static_assert(LG >= 8, "one lane per element of the line record");
            constexpr int R0 = LG >= 16 ? 8 : 0;                 // the scalar outputs: on lanes 8.. when the group has them
            if (r < 8) a.line[s * 8 + r] = out;
            if (r == R0) a.ninl[s] = ok ? s3->mx : (s3->status == 0x7fffffff ? 0 : s3->status);
            if (r == R0 + 1) a.win_trial[s] = ok ? s3->best_t : -1;
            if (r == R0 + 2) a.win_pair[2 * s] = ok ? s3->best_i : -1;
            if (r == R0 + 3) a.win_pair[2 * s + 1] = ok ? s3->best_j : -1;
            if (r == R0 + 4 && a.ssum != nullptr) a.ssum[s] = ok ? s3->best_ss : qnan;
        }
        // inlier masks of this warp's slots: consecutive slots are consecutive in global memory
        const int nq = (nsl - qw) < SPW ? (nsl - qw) : SPW;
        for (int i = lane; i < nq * W; i += 32) {
            const int qi = i / W, w = i - qi * W;
            const TileSlot *s4 = slots + qw + qi;
            a.mask[(slot0 + qw) * W + i] = (s4->status == 0 && w * 32 < s4->n) ? sm.smask[(qw + qi) * W + w] : 0u;
        }
    }
}

// score + candidates of one tile: one work item (TPT trials of one chunk) per thread.  Inlined into the kernel, with the
// kernel's own parameter (constant-bank operands).
template <int TPT>
__device__ __forceinline__ void tile_score(const RansacArgs &a, int tile, int par)
{
    LSLAM_SMEM_DECL;
    const int T = a.trials, G = a.G, chunk = a.chunk, W = a.W, ipc = a.ipc;
    const int chunk_pad = (chunk + 1) & ~1;
    const int tid = threadIdx.x;
    const int lane = tid & 31, nwarps = blockDim.x >> 5;
    const TileSmem sm = tile_carve(smem_raw, G, chunk, ipc, TPT, W, nwarps);
    const long long slot0 = (long long)tile * G;
    const long long rem_slots = a.nslots - slot0;
    const int nsl = rem_slots < (long long)G ? (int)rem_slots : G;
    double2 *const pts = reinterpret_cast<double2 *>(sm.region(par));
    float2 *const ptsf = reinterpret_cast<float2 *>(sm.region(par) + sm.ptsf_off);
    int2 *const keys = reinterpret_cast<int2 *>(sm.region(par ^ 1));                    // [G][TPT][ipc]
    uint32_t *const mw = reinterpret_cast<uint32_t *>(sm.region(par ^ 1) + sm.mw_off);  // [G][W][TPT][ipc]

    // ================= score: one work item (TPT trials of one chunk) per thread =============================
    const int items = nsl * ipc;
    const bool has_item = tid < items;
    const int q = has_item ? tid / ipc : 0, tb = has_item ? tid - q * ipc : 0;
    TileSlot *st = sm.slot(par) + q;
    const int n = st->n, t0 = tb * TPT;
    const bool skip = !has_item || st->status != 0;
    const double2 *P = pts + st->off;
    const int row0 = (q * W * TPT) * ipc + tb;                    // trial k, word w: mw[row0 + (w * TPT + k) * ipc]
    int2 *krow = keys + (size_t)q * TPT * ipc + tb;               // trial k: krow[k * ipc]
    // (the per-trial inlier words in `mw` are kept BIT-REVERSED, as the funnel shifts of the loop leave them: only the
    //  winner's words are reversed, when they are copied to `smask` - 4 BREV per chunk instead of 16 per thread)
    int cnt[TPT];
    unsigned live = 0, exact = 0;                                 // bit k: trial t0 + k exists / was scored literally
    {
        const bool force_lit = st->force_lit != 0;
        const uint64_t scan = st->scan;
        // the hot loop's line in fp32 on the centred points, scaled by the slot's kscale (a power of two: tile_screen32)
        float nx[TPT], ny[TPT], ng[TPT];
        unsigned be[TPT], bo[TPT], far[TPT];                      // two bits per point: of the even / odd points of a word; AND of them
        const float ks = st->kscale;
        unsigned flagl = 0;                                       // bit k: trial t0 + k must be scored literally
        const double nn = st->nn, A = st->A, sb_alpha = st->alpha, sb_beta = st->beta, kb = st->kb, c0x = st->c0x, c0y = st->c0y;
#pragma unroll
        for (int k = 0; k < TPT; k += 2) {
            const Philox4 ph = philox4x32_10((uint32_t)(t0 + k) >> 1, (uint32_t)st->c, (uint32_t)scan, (uint32_t)(scan >> 32),
                                             (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                nx[k + h] = 0; ny[k + h] = 0; ng[k + h] = 0;
                be[k + h] = 0u; bo[k + h] = 0u; cnt[k + h] = 0; far[k + h] = 0xffffffffu;
                if (!skip && t0 + k + h < T) {
                    live |= 1u << (k + h);
                    int i, j;
                    if (a.pairs != nullptr) explicit_pair(a, slot0 + q, (uint32_t)(t0 + k + h), (uint32_t)n, i, j);
                    else pair_from_words(h ? ph.r[2] : ph.r[0], h ? ph.r[3] : ph.r[1], (uint32_t)n, i, j);
                    const TileLine f = tile_line(P[i], P[j], c0x, c0y);
                    const double fnx = f.nx, fny = f.ny, cp = f.cp;
                    const bool degen = f.degenerate;
                    nx[k + h] = (float)fnx * ks; ny[k + h] = (float)fny * ks; ng[k + h] = (float)cp * ks;   // exact scalings
                    // residual sum from the moments with its error bound -> the two keys the tie-break needs
                    // E = beta (M+ + e2) + alpha + e2 with e2 = tile_moment_err(cp): the per-slot factors are folded (kb)
                    const double M = tile_moment_sum(fnx, fny, cp, nn, st);
                    const double E = fma(kb, fma(nn * cp, cp, A), fma(sb_beta, M > 0.0 ? M : 0.0, sb_alpha));
                    const double up = M + E;
                    krow[(k + h) * ipc] = make_int2(key_down(M - E), key_up(up));
                    // literal score wanted: coincident samples, a chunk outside the fast path's range, a non-finite sum /
                    // bound (exponent field of M + E all ones; E >= 0, so an infinite or NaN M or E always shows there).
                    // (A point too close to the threshold - known after the loop - gets the fp64 fast line first.)
                    if (degen || force_lit || (__double2hiint(up) & 0x7ff00000) == 0x7ff00000) flagl |= 1u << (k + h);
                }
            }
        }
        const int n_eff = (skip || force_lit) ? 0 : n;
        {
            // f = ks (|s_f| - thr_f), bit for bit (ks is a power of two).  Its sign bit = inlier; its exponent bit 30 is set
            // exactly when |f| >= 2, i.e. when the unscaled | |s_f| - thr_f | >= band' >= bandf - the point is far enough
            // from the threshold for the fp32 decision to be the literal one.  A funnel shift by two takes both bits; the
            // even and the odd points of a 32-point word go to two registers, so that the inlier bits of the first (odd bit
            // positions) and of the second shifted down by one (even positions) merge into the word of all 32 with ONE
            // LOP3, bit 31 - j = point j.  (The loop is bound by issue cycles: the running minimum of |f| this replaces
            // was 18 % of them - tools/probe/mix_probe5.cu.)
            const float thf = (float)a.thr * ks;
            const float2 *Pf = ptsf + (size_t)q * chunk_pad;
            const float4 *Pf4 = reinterpret_cast<const float4 *>(Pf);     // two points per LDS.128 (chunk_pad is even)
#define LSLAM_TILE_STEP2(XA, YA, XB, YB)                                                          \
    {                                                                                             \
        _Pragma("unroll") for (int k_ = 0; k_ < TPT; ++k_) {                                      \
            const float fa_ = fabsf(fmaf((YA), ny[k_], fmaf((XA), nx[k_], ng[k_]))) - thf;        \
            const float fb_ = fabsf(fmaf((YB), ny[k_], fmaf((XB), nx[k_], ng[k_]))) - thf;        \
            be[k_] = __funnelshift_l(__float_as_uint(fa_), be[k_], 2);                            \
            bo[k_] = __funnelshift_l(__float_as_uint(fb_), bo[k_], 2);                            \
        }                                                                                         \
    }
            int k = 0, w = 0;
            for (; k + 32 <= n_eff; k += 32, ++w) {
#pragma unroll
                for (int kk = 0; kk < 32; kk += 2) { const float4 pp = Pf4[(k + kk) >> 1]; LSLAM_TILE_STEP2(pp.x, pp.y, pp.z, pp.w); }
#pragma unroll
                for (int k_ = 0; k_ < TPT; ++k_) {
                    const unsigned m = (be[k_] & 0xaaaaaaaau) | ((bo[k_] >> 1) & 0x55555555u);   // bit 31 - j = point 32 w + j
                    far[k_] &= be[k_] & bo[k_];
                    cnt[k_] += __popc(m);
                    mw[row0 + (w * TPT + k_) * ipc] = m;
                }
            }
            const int rem = n_eff - k;
            if (rem > 0) {
                // the tail in pairs; an odd last point is paired with itself (its copy is masked out of the word)
                for (int kk = 0; kk < rem; kk += 2) {
                    float4 pp;
                    if (kk + 1 < rem) pp = Pf4[(k + kk) >> 1];
                    else { const float2 p1 = Pf[k + kk]; pp = make_float4(p1.x, p1.y, p1.x, p1.y); }
                    LSLAM_TILE_STEP2(pp.x, pp.y, pp.z, pp.w);
                }
                const int sh = 32 - ((rem + 1) & ~1);                              // 32 - 2 x pairs: 0 .. 30
                const unsigned keep = 0xffffffffu << (32 - rem), low = ~(0xffffffffu << sh);
#pragma unroll
                for (int k_ = 0; k_ < TPT; ++k_) {
                    const unsigned e_ = be[k_] << sh, o_ = bo[k_] << sh;
                    const unsigned m = ((e_ & 0xaaaaaaaau) | ((o_ >> 1) & 0x55555555u)) & keep;
                    far[k_] &= (e_ & o_) | low;
                    cnt[k_] += __popc(m);
                    mw[row0 + (w * TPT + k_) * ipc] = m;
                }
            }
#undef LSLAM_TILE_STEP2
        }
        // a point too close to the threshold for the fp32 distance to decide (some "far" bit clear): the trial is re-scored
        // with the fp64 fast line
        unsigned flagb = 0;
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_)
            if ((far[k_] & 0x55555555u) != 0x55555555u) flagb |= 1u << k_;
        flagl &= live;
        flagb &= live & ~flagl;

        // flagged trials: scored again by the whole warp, on the spot - with the fp64 fast line (its own re-check band is
        // 2^-36 of the coordinate scale: the count and the mask are then the literal ones), or with the literal formula
        // (trials flagged for it at the set-up, or a point inside the fp64 band as well); the new mask replaces the fast one
        unsigned fm = __ballot_sync(LSLAM_FULL_MASK, (flagl | flagb) != 0u);
        unsigned n_relit = 0;
        while (fm) {
            const int src = __ffs((int)fm) - 1;
            fm &= fm - 1;
            const int fq = __shfl_sync(LSLAM_FULL_MASK, q, src), ftb = __shfl_sync(LSLAM_FULL_MASK, tb, src);
            const unsigned fll = __shfl_sync(LSLAM_FULL_MASK, flagl, src);
            unsigned fl = fll | __shfl_sync(LSLAM_FULL_MASK, flagb, src);
            const TileSlot *fs = sm.slot(par) + fq;
            const double2 *FP = pts + fs->off;
            const int frow0 = (fq * W * TPT) * ipc + ftb;
            while (fl) {
                const int kf = __ffs((int)fl) - 1;
                fl &= fl - 1;
                int i, j, cl = 0;
                double sl = 0.0;
                trial_pair(a, slot0 + fq, fs->scan, (uint32_t)fs->c, (uint32_t)(ftb * TPT + kf), (uint32_t)fs->n, i, j);
                bool lit = (fll >> kf) & 1u;
                if (!lit)
                    warp_fast64_score_strided(FP, fs->n, fast_pair_model(FP[i], FP[j]), a.thr, fs->band_hi, lane, cl, lit,
                                              mw + frow0 + kf * ipc, TPT * ipc);
                if (lit) {
                    warp_literal_score_strided(FP, fs->n, lit_pair_model(FP[i], FP[j]), a.thr, lane, cl, sl, mw + frow0 + kf * ipc,
                                               TPT * ipc);
                    ++n_relit;
                }
                if (lane == src) {
                    if (lit) {
                        krow[kf * ipc] = make_int2(key_down(sl), key_up(sl));
                        exact |= 1u << kf;
                    }
#pragma unroll
                    for (int k_ = 0; k_ < TPT; ++k_)
                        if (k_ == kf) cnt[k_] = cl;
                }
            }
        }
        if (lane == 0 && n_relit) atomicAdd(a.counters + 0, (unsigned long long)n_relit);

        // chunk-wide maximum: count, then the smallest upper bound of the literal residual sum
        unsigned long long best = 0ull;
#pragma unroll
        for (int k_ = 0; k_ < TPT; ++k_) {
            if ((live >> k_) & 1u) {
                const unsigned long long key = ((unsigned long long)(unsigned)cnt[k_] << 32) | (unsigned)(0x7fffffff - krow[k_ * ipc].y);
                best = best > key ? best : key;
            }
        }
        // the work items of a slot are consecutive threads: an inclusive max-scan over the slot's lanes of this warp leaves the
        // slot's maximum in its last lane, which alone touches the slot's key (a 64-bit shared-memory atomicMax is a
        // compare-and-swap loop: 25 lanes contending for one word took four rounds of it)
        const int seg_lo = q * ipc - (tid - lane), seg0 = has_item ? (seg_lo > 0 ? seg_lo : 0) : lane;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(LSLAM_FULL_MASK, best, d);
            if (lane - d >= seg0 && o > best) best = o;
        }
        if (has_item && best != 0ull && (tb == ipc - 1 || lane == 31)) atomicMax(&st->key, best);
    }
    __syncthreads();

    // ================= candidates: trials tied on the maximal count whose lower bound reaches the smallest upper bound
    if (!skip) {
        const unsigned long long key = st->key;
        const int mx = (int)(key >> 32), minhi = 0x7fffffff - (int)(unsigned)key;
        if (tb == 0) {
            st->mx = mx;
            if (mx < 2) st->status = kStatusNoModel;              // upstream's final estimate() needs 2 points
        }
        if (mx >= 2) {
            unsigned cm = 0;
#pragma unroll
            for (int k_ = 0; k_ < TPT; ++k_)
                if (((live >> k_) & 1u) && cnt[k_] == mx && krow[k_ * ipc].x <= minhi) cm |= 1u << k_;
            while (cm) {                                          // rare: about one trial per chunk
                const int kc = __ffs((int)cm) - 1;
                cm &= cm - 1;
                const int pos = atomicAdd(&st->ncand, 1);
                if (pos < kMaxCand) {
                    int i, j;
                    trial_pair(a, slot0 + q, st->scan, (uint32_t)st->c, (uint32_t)(t0 + kc), (uint32_t)n, i, j);
                    st->cand_t[pos] = t0 + kc; st->cand_i[pos] = i; st->cand_j[pos] = j;
                    if ((exact >> kc) & 1u) atomicOr(&st->cand_exact, 1 << pos);
                    // the first candidate's inlier words leave the scratch region; with several candidates the literal
                    // tie-break writes the winner's
                    if (pos == 0)
                        for (int w2 = 0; w2 * 32 < n; ++w2) sm.smask[q * W + w2] = __brev(mw[row0 + (w2 * TPT + kc) * ipc]);
                }
            }
        }
    }
    __syncthreads();
}

template <int TPT, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) ransac_tile_kernel(const RansacArgs a)
{
    static_assert(TPT % 2 == 0, "a Philox block serves two trials");
    // One tile per CTA.  Co-resident CTAs fall out of step and hide each other's narrow phases (a persistent CTA staging
    // its next tile while it refits the current one measured SLOWER on B200: 5.1 against 4.3 ms on C2 - the CTAs of an SM
    // then run in lockstep and idle together)
    const int tile = blockIdx.x, nwarps = blockDim.x >> 5;
    tile_stage<TPT>(a, tile, 0, 0, nwarps);
    __syncthreads();
    tile_score<TPT>(a, tile, 0);                             // score | barrier | candidates
    __syncthreads();
    tile_refit<TPT>(a, tile, 0, nwarps);
}

} // namespace lslam
