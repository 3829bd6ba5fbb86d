// lslam_core.cuh - arithmetic building blocks shared by the kernels.
//
// Two kinds of arithmetic live here:
//   * "literal" ops follow the operation order of the upstream algorithm the reference
//     calls (skimage LineModelND, restated in oracle/ransac_line.py) with every product and
//     sum rounded on its own (__dadd_rn/__dmul_rn/... are never contracted into FMAs), so
//     inlier masks, counts and residual sums are bit-identical to the CPU oracle;
//   * "fast" ops use FMAs and a normal-form line; they only ever decide things that are
//     provably the same under the literal formula (see FastLine / chunk_band below).
#pragma once
#include <math.h>
#include <stdint.h>

#ifndef LSLAM_EMU
#include <cuda_runtime.h>
#endif

#define LSLAM_FULL_MASK 0xffffffffu

namespace lslam {

// ---- Philox4x32-10 (Salmon et al., SC'11); contract in oracle/philox.py -------------------
struct Philox4 { uint32_t r[4]; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int rnd = 0; rnd < 10; ++rnd) {
        if (rnd) { k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    }
    Philox4 o; o.r[0] = c0; o.r[1] = c1; o.r[2] = c2; o.r[3] = c3;
    return o;
}

// (u, v) -> index pair i != j in [0, n)
__device__ __forceinline__ void pair_from_words(uint32_t u, uint32_t v, uint32_t n, int &i, int &j)
{
    const uint32_t a = __umulhi(u, n);
    uint32_t b = __umulhi(v, n - 1u);
    if (b >= a) ++b;
    i = (int)a; j = (int)b;
}

__device__ __forceinline__ void sample_pair(uint64_t seed, uint64_t scan, uint32_t chunk, uint32_t trial,
                                            uint32_t n, int &i, int &j)
{
    const Philox4 p = philox4x32_10(trial >> 1, chunk, (uint32_t)scan, (uint32_t)(scan >> 32),
                                    (uint32_t)seed, (uint32_t)(seed >> 32));
    const bool odd = trial & 1u;
    pair_from_words(odd ? p.r[2] : p.r[0], odd ? p.r[3] : p.r[1], n, i, j);
}

// ---- literal LineModelND (oracle/ransac_line.py: line_from_pair, line_residuals) ----------
struct LitLine { double ox, oy, dx, dy; };

__device__ __forceinline__ LitLine lit_pair_model(double2 pi, double2 pj)
{
    LitLine m;
    m.ox = __dmul_rn(__dadd_rn(pi.x, pj.x), 0.5);           // (p_i + p_j) / 2
    m.oy = __dmul_rn(__dadd_rn(pi.y, pj.y), 0.5);
    const double cix = __dsub_rn(pi.x, m.ox), ciy = __dsub_rn(pi.y, m.oy);
    const double cjx = __dsub_rn(pj.x, m.ox), cjy = __dsub_rn(pj.y, m.oy);
    double dx = __dsub_rn(cjx, cix), dy = __dsub_rn(cjy, ciy);
    const double n = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    if (n != 0.0) { dx = __ddiv_rn(dx, n); dy = __ddiv_rn(dy, n); }
    m.dx = dx; m.dy = dy;
    return m;
}

__device__ __forceinline__ double lit_residual(const LitLine &m, double2 p)
{
    const double vx = __dsub_rn(p.x, m.ox), vy = __dsub_rn(p.y, m.oy);
    const double t = __dadd_rn(__dmul_rn(vx, m.dx), __dmul_rn(vy, m.dy));
    const double rx = __dsub_rn(vx, __dmul_rn(t, m.dx)), ry = __dsub_rn(vy, __dmul_rn(t, m.dy));
    return __dsqrt_rn(__dadd_rn(__dmul_rn(rx, rx), __dmul_rn(ry, ry)));
}

// ---- fast normal-form line ----------------------------------------------------------------
// signed distance s(p) = nx*px + ny*py + negc, evaluated with two FMAs.  (nx, ny) is the unit
// normal of the line through the two sample points, computed from the same literal centred
// difference d as lit_pair_model but normalised with a Newton-refined reciprocal square root
// (relative error ~1e-16).  |s| differs from the literal residual by at most ~40 eps * S where
// S bounds the chunk's |coordinates| (derivation in DESIGN.md "Exactness of the fast path");
// chunk_band() turns that into the re-check band, three orders of magnitude wider.
struct FastLine { double nx, ny, negc; bool degenerate; };

__device__ __forceinline__ FastLine fast_pair_model(double2 pi, double2 pj)
{
    const double ox = (pi.x + pj.x) * 0.5, oy = (pi.y + pj.y) * 0.5;
    const double dx = (pj.x - ox) - (pi.x - ox), dy = (pj.y - oy) - (pi.y - oy);
    const double n2 = fma(dx, dx, dy * dy);
    FastLine f;
    // outside this range (coincident samples, n2 == 0, or absurd scales) the trial is scored
    // with the literal formula instead
    f.degenerate = !(n2 > 0x1p-100 && n2 < 0x1p100);   // also keeps (float)n2 in range
    const double n2s = f.degenerate ? 1.0 : n2;
    double y = (double)rsqrtf((float)n2s);
    const double h = 0.5 * n2s;
    y = y * fma(-h * y, y, 1.5);      // one Newton step: relative error < 2^-42, far inside the band
    f.nx = -dy * y;
    f.ny = dx * y;
    f.negc = -fma(f.nx, ox, f.ny * oy);
    return f;
}

// Per-chunk scale -> re-check band.  e2 = exponent such that every |coordinate| < 2^e2.
//   band  = 2^(e2-36): a trial with any | |s| - thr | < band is re-scored literally
//   eabs  = band as a double, used in the bound on the fast residual sum
// Returned band_hi is the high word of the double `band` (a power of two, low word 0), so
// |u| < band  <=>  (hi(u) & 0x7fffffff) < band_hi.
struct ChunkBand { int band_hi; double eabs; bool force_literal; bool bad; };

__device__ __forceinline__ ChunkBand chunk_band(int max_abs_hi /* max over coords of hi(|c|) */)
{
    ChunkBand b;
    const int ex = (max_abs_hi >> 20) & 0x7ff;               // biased exponent of the largest |coord|
    b.bad = ex >= 1023 + 400;                                // Inf/NaN or absurdly large
    b.force_literal = ex <= 1023 - 400;                      // all-zero / denormal-scale chunk
    int e2 = ex - 1023 + 1;
    if (b.bad || b.force_literal) e2 = 0;
    b.band_hi = (e2 - 36 + 1023) << 20;
    b.eabs = __hiloint2double(b.band_hi, 0);
    return b;
}

// Upper bound on |ssum_literal - ssum_fast| for a chunk of n points:
//   per point | r^2 - s^2 | <= eabs * (2|s| + eabs);  sum |s| <= sqrt(n * ssf)  (Cauchy-Schwarz)
//   2 sqrt(n * ssf) <= n*c + ssf/c for any c > 0 (AM-GM); c = the threshold clamped to
//   [eabs, chunk scale] keeps the bound tight where it matters (residuals of the order of thr)
//   plus 2^-38 * ssf for the differing summation orders and roundings (n <= 8192).
// The bound is linear in ssf, E(ssf) = alpha + beta * ssf, so a sweep over the trials needs no sqrt.
struct SsumBound { double alpha, beta, kappa; };   // E(ssf) = alpha + beta * ssf (kappa: unused, 0)
__device__ __forceinline__ double ssum_err(const SsumBound &b, double v)
{
    double e = fma(b.beta, v, b.alpha);
    if (b.kappa != 0.0) e = fma(b.kappa, sqrt(v) * 1.0000001, e);
    return e;
}
__device__ __forceinline__ SsumBound ssum_bound_coeffs(double eabs, int n, double thr)
{
    const double nn = (double)n;
    const double scale = eabs * 0x1p36;                  // chunk scale 2^e2
    double c = thr;
    if (!(c > eabs)) c = eabs;
    if (c > scale) c = scale;
    SsumBound b;
    b.alpha = 1.0001 * eabs * nn * (c + eabs);
    b.beta = 1.0001 * (eabs / c + 0x1p-38);
    b.kappa = 0.0;
    return b;
}

// the same with 1 / thr supplied (its rounding, 2^-53 relative, disappears in the factor 1.0001)
__device__ __forceinline__ SsumBound ssum_bound_coeffs_inv(double eabs, int n, double thr, double inv_thr)
{
    const double scale = eabs * 0x1p36;
    if (!(thr > eabs) || thr > scale) return ssum_bound_coeffs(eabs, n, thr);
    const double nn = (double)n;
    SsumBound b;
    b.alpha = 1.0001 * eabs * nn * (thr + eabs);
    b.beta = 1.0001 * fma(eabs, inv_thr, 0x1p-38);
    b.kappa = 0.0;
    return b;
}

// ---- principal axis of a 2x2 scatter matrix (closed form of the SVD upstream uses) --------
// direction of the largest eigenvalue of [[sxx, sxy], [sxy, syy]], unit length, d_x >= 0.
__device__ __forceinline__ void principal_axis(double sxx, double sxy, double syy, double &dx, double &dy)
{
    const double del = 0.5 * (sxx - syy);
    const double h = sqrt(fma(del, del, sxy * sxy));
    double vx, vy;
    if (h == 0.0) { vx = 1.0; vy = 0.0; }
    else if (del >= 0.0) { vx = del + h; vy = sxy; }
    else { vx = sxy; vy = h - del; }
    const double nn = sqrt(fma(vx, vx, vy * vy));
    if (nn != 0.0) { vx /= nn; vy /= nn; } else { vx = 1.0; vy = 0.0; }
    if (vx < 0.0 || (vx == 0.0 && vy < 0.0)) { vx = -vx; vy = -vy; }
    dx = vx; dy = vy;
}

// the same with reciprocal square roots instead of sqrt + divisions (results agree to a few ulp; the refit is held to 1e-5)
__device__ __forceinline__ void principal_axis_rsqrt(double sxx, double sxy, double syy, double &dx, double &dy)
{
    const double del = 0.5 * (sxx - syy);
    const double h2 = fma(del, del, sxy * sxy);
    double vx = 1.0, vy = 0.0;
    if (h2 > 0.0) {
        const double h = h2 * rsqrt(h2);
        if (del >= 0.0) { vx = del + h; vy = sxy; }
        else { vx = sxy; vy = h - del; }
        const double n2 = fma(vx, vx, vy * vy);
        if (n2 > 0.0 && n2 < 1.7976931348623157e308) { const double inv = rsqrt(n2); vx *= inv; vy *= inv; }
        else principal_axis(sxx, sxy, syy, vx, vy);
    }
    if (vx < 0.0 || (vx == 0.0 && vy < 0.0)) { vx = -vx; vy = -vy; }
    dx = vx; dy = vy;
}

// ---- warp helpers ------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(LSLAM_FULL_MASK, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v = fmin(v, __shfl_xor_sync(LSLAM_FULL_MASK, v, o));
    return v;
}
// the fixed summation tree of oracle/ransac_line.py tree32_sum (lane partials -> xor butterfly)
__device__ __forceinline__ double warp_tree32(double part)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) part = __dadd_rn(part, __shfl_xor_sync(LSLAM_FULL_MASK, part, o));
    return part;
}

} // namespace lslam
