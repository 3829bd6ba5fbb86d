// assoc.cuh - K3: landmark data association, one lane group (8, 16 or 32 lanes) per landmark table (stream).
//
// Replaces the association loop of landmark_extraction (/root/reference/ransac_functions.py:34-54,
// is_equal from landmarking.py:66-77) plus the caller-side append and extraction counter of
// check_ransac (ransac_functions.py:74-78).  Semantics = oracle/assoc.py.
//
// The stored landmarks of a table are tested against the query by the lanes of the warp (32 per
// tile, table staged in shared memory); the reference's SEQUENTIAL side effects - every landmark
// visited before the hit loses a life, an expired one is removed, and the element that slides
// into a removed slot is skipped (mode 0, the reference's bug) - are then resolved on the ballot
// masks with a handful of integer ops per event, so the walk costs O(events), not O(M).
#pragma once
#include "lslam_core.cuh"
#include "async_copy.cuh"
#include <cmath>

namespace lslam {

struct AssocArgs {
    const double *query;       // [T, Q, qstride]
    const int32_t *q_valid;    // [T, Q] or null
    int qstride, layout;       // layout 0: a,b,px,py,ex,ey ; 1: ransac line record (a,b,ox,oy,dx,dy,tipx,tipy)
    int T, Q, M;
    double *table;             // [T, M, 6]
    int32_t *life, *ids;       // [T, M]
    int32_t *count, *next_id;  // [T]
    double tolA, tolB, tol;
    double tol_sq;             // largest double s with sqrt_rn(s) <= tol (assoc_gate_sq), -1 if there is none: the gate on squares
    int life_init, mode, append;
    int32_t *match, *match_id; // [T, Q]
    double *proj_ab;           // [T, Q, 2] or null
    double *hit_pos;           // [T, Q, 2] or null
    int stage_q;               // the table's Q queries (+ q_valid) are staged in shared memory with the table (they fit)
    int async16;               // table, life, ids and queries are copied by 16-byte cp.async (alignment / sizes checked on the host)
};

constexpr int kLifeRemoved = -1;

// shared memory of one table: entries [M][6], life [M], ids [M]; with stage_q also the queries [Q][qstride] and q_valid [Q]
__host__ __device__ inline size_t assoc_smem_per_table(int M, int Q, int qstride, int stage_q)
{
    size_t b = (((size_t)M * (6 * 8 + 2 * 4)) + 15) & ~(size_t)15;
    if (stage_q) b += (((size_t)Q * qstride * 8 + (size_t)Q * 4) + 15) & ~(size_t)15;
    return b;
}

// ||(vx, vy)|| with the rounding pinned in oracle/assoc.py: sqrt(fma(vy, vy, vx*vx))
__device__ __forceinline__ double norm2_pinned(double vx, double vy)
{
    return __dsqrt_rn(__fma_rn(vy, vy, __dmul_rn(vx, vx)));
}

// The gate ||v|| <= tol without the square root: the correctly rounded sqrt is monotonic, so sqrt_rn(s) <= tol  <=>  s <= S*
// with S* the largest double whose rounded root does not exceed tol - found on the host by stepping from tol * tol (the
// host's sqrt is the same correctly rounded operation).  NaN sums fail both forms; tol < 0 or NaN: no s passes (S* = -1).
inline double assoc_gate_sq(double tol)
{
    if (!(tol >= 0.0)) return -1.0;
    if (tol > 1.3407807929942596e154) return HUGE_VAL;          // tol * tol overflows: every finite (and infinite) sum passes
    double s = tol * tol;
    while (std::sqrt(s) > tol) s = std::nextafter(s, 0.0);
    for (;;) {
        const double up = std::nextafter(s, HUGE_VAL);
        if (!(std::sqrt(up) <= tol) || up == s) break;
        s = up;
    }
    return s;
}

// landmarking.py:66-77, self = stored entry e[6], other = query q[6].  WANT_D: also return the smaller of the two gate
// distances (the nearest mode ranks by it); without it the two square roots are not taken at all - the gates are
// compared on the squared sums against tol_sq, the same decision bit for bit (assoc_gate_sq).
template <bool WANT_D>
__device__ __forceinline__ bool landmark_is_equal(const double *e, const double *q, double tolA, double tolB,
                                                  double tol, double tol_sq, double &dmin)
{
    const double distA = fabs(__dsub_rn(e[0], q[0]));
    const double distB = fabs(__dsub_rn(e[1], q[1]));
    const double ax = __dsub_rn(e[4], q[2]), ay = __dsub_rn(e[5], q[3]);             // end_s - pos_q
    const double bx = __dsub_rn(e[2], q[4]), by = __dsub_rn(e[3], q[5]);             // pos_s - end_q
    if (WANT_D) {
        const double dEO = norm2_pinned(ax, ay), dOE = norm2_pinned(bx, by);
        dmin = dEO < dOE ? dEO : dOE;
        return distA <= tolA && distB <= tolB && (dEO <= tol || dOE <= tol);
    }
    const double sEO = __fma_rn(ay, ay, __dmul_rn(ax, ax)), sOE = __fma_rn(by, by, __dmul_rn(bx, bx));
    return distA <= tolA && distB <= tolB && (sEO <= tol_sq || sOE <= tol_sq);
}

// GS lanes (8, 16 or 32) serve one table; a warp serves 32/GS tables.  GS < 32 requires M <= GS (one
// tile per table), so that the reference's table sizes (LANDMARK_NUMBER = 8) do not leave three
// quarters of every warp idle.  Collectives are restricted to the group's lanes: groups of one warp
// run different tables and diverge freely.
template <int GS>
__global__ void __launch_bounds__(256, 4) associate_kernel(const AssocArgs a)
{
    LSLAM_SMEM_DECL;
    constexpr int NG = 32 / GS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int grp = lane / GS, g = lane % GS;
    const unsigned gfull = GS == 32 ? 0xffffffffu : ((1u << GS) - 1u);
    const unsigned gmask = gfull << (grp * GS);
    const long long tl = ((long long)blockIdx.x * nwarps + warp) * NG + grp;
    if (tl >= a.T) return;                      // the whole group leaves together
    const int t = (int)tl;
    const int M = a.M;
    unsigned char *base = smem_raw + (size_t)(warp * NG + grp) * assoc_smem_per_table(M, a.Q, a.qstride, a.stage_q);
    double *ent = reinterpret_cast<double *>(base);                 // [M][6]
    int *life = reinterpret_cast<int *>(ent + (size_t)M * 6);       // [M]
    int *ids = life + M;                                            // [M]
    double *qbuf = reinterpret_cast<double *>(base + ((((size_t)M * (6 * 8 + 2 * 4)) + 15) & ~(size_t)15));   // [Q][qstride] (stage_q)
    int *qv = reinterpret_cast<int *>(qbuf + (size_t)a.Q * a.qstride);                                       // [Q]
#define GBALLOT(pred) ((__ballot_sync(gmask, (pred)) >> (grp * GS)) & gfull)

    // Everything the walk reads comes in ONE global round trip: the copies below do not depend on the live count (all M slots
    // are copied; slots past the count are never looked at), and the queries of the table come with it instead of one
    // dependent load per query inside the loop.  With async16 the copies are cp.async (no register staging: all in flight).
    int cnt, next_id;
    {
        const double *src = a.table + (size_t)t * M * 6;
        const int32_t *lsrc = a.life + (size_t)t * M, *isrc = a.ids + (size_t)t * M;
        const double *qsrc = a.query + (size_t)t * a.Q * a.qstride;
        if (a.async16) {
            for (int k = g; k < M * 3; k += GS) cp_async16(ent + 2 * k, src + 2 * k);
            for (int k = g; k < (M >> 2); k += GS) { cp_async16(life + 4 * k, lsrc + 4 * k); cp_async16(ids + 4 * k, isrc + 4 * k); }
            if (a.stage_q)
                for (int k = g; k < ((a.Q * a.qstride) >> 1); k += GS) cp_async16(qbuf + 2 * k, qsrc + 2 * k);
        }
        cnt = a.count[t];
        next_id = a.next_id[t];
        if (a.stage_q && a.q_valid != nullptr)
            for (int k = g; k < a.Q; k += GS) qv[k] = a.q_valid[(size_t)t * a.Q + k];
        cnt = cnt < 0 ? 0 : (cnt > M ? M : cnt);
        if (a.async16) {
            cp_async_wait_all();
        } else {
            for (int k = g; k < cnt * 6; k += GS) ent[k] = src[k];
            for (int k = g; k < cnt; k += GS) { life[k] = lsrc[k]; ids[k] = isrc[k]; }
            if (a.stage_q)
                for (int k = g; k < a.Q * a.qstride; k += GS) qbuf[k] = qsrc[k];
        }
    }
    __syncwarp(gmask);

    for (int qi = 0; qi < a.Q; ++qi) {
        const size_t qo = (size_t)t * a.Q + qi;
        if (a.q_valid != nullptr && (a.stage_q ? qv[qi] : a.q_valid[qo]) <= 0) {
            if (g == 0) {
                a.match[qo] = -2; a.match_id[qo] = -1;
                const double qn = __longlong_as_double(0x7ff8000000000000LL);
                if (a.proj_ab) { a.proj_ab[2 * qo] = qn; a.proj_ab[2 * qo + 1] = qn; }
                if (a.hit_pos) { a.hit_pos[2 * qo] = qn; a.hit_pos[2 * qo + 1] = qn; }
            }
            continue;
        }
        const double *qs = a.stage_q ? qbuf + (size_t)qi * a.qstride : a.query + qo * a.qstride;
        double q[6];
        q[0] = qs[0]; q[1] = qs[1]; q[2] = qs[2]; q[3] = qs[3];
        q[4] = a.layout ? qs[6] : qs[4];
        q[5] = a.layout ? qs[7] : qs[5];
        const int qid = next_id++;

        int hit = -1, removed = 0;
        if (a.mode != 2) {
            // ---- first match in list order with the reference's side effects ----
            int p = 0;
            for (int tb = 0; tb < cnt && hit < 0; tb += GS) {
                const int m = tb + g;
                const bool in = m < cnt;
                double dmin;
                const bool eq = in && landmark_is_equal<false>(ent + (size_t)m * 6, q, a.tolA, a.tolB, a.tol, a.tol_sq, dmin);
                const int lf = in ? life[m] : 0;
                const bool die = in && !eq && lf <= 1;           // decrease_life() would return True
                const unsigned eqm = GBALLOT(eq);
                const unsigned diem = GBALLOT(die);
                const int tend = cnt - tb < GS ? cnt - tb : GS;  // valid positions in this tile
                unsigned vis = 0, rem = 0;
                while (p < tb + tend) {
                    const int pp = p - tb;
                    const unsigned ev = (eqm | diem) >> pp;
                    const unsigned upto_end = (tend == 32 ? 0xffffffffu : ((1u << tend) - 1u)) & ~((1u << pp) - 1u);
                    if (ev == 0) { vis |= upto_end; p = tb + tend; break; }
                    const int pe = pp + __ffs((int)ev) - 1;
                    vis |= upto_end & ((1u << pe) - 1u);         // [pp, pe) visited, no event
                    if ((eqm >> pe) & 1u) { hit = tb + pe; break; }
                    vis |= 1u << pe;
                    rem |= 1u << pe;
                    p = tb + pe + (a.mode == 0 ? 2 : 1);         // mode 0: the slot after a removal is skipped
                }
                if (in && ((vis >> g) & 1u)) life[m] = ((rem >> g) & 1u) ? kLifeRemoved : (lf > 0 ? lf - 1 : 0);
                removed += __popc(rem);
            }
        } else {
            // ---- nearest gated landmark; every other landmark counts as "not seen" ----
            double best = __longlong_as_double(0x7ff0000000000000LL);
            for (int tb = 0; tb < cnt; tb += GS) {
                const int m = tb + g;
                double dmin = best;
                const bool eq = m < cnt && landmark_is_equal<true>(ent + (size_t)m * 6, q, a.tolA, a.tolB, a.tol, a.tol_sq, dmin);
                double cand = eq ? dmin : __longlong_as_double(0x7ff0000000000000LL);
                int ci = eq ? m : 0x7fffffff;
#pragma unroll
                for (int o = GS / 2; o >= 1; o >>= 1) {
                    const double od = __shfl_xor_sync(gmask, cand, o);
                    const int oi = __shfl_xor_sync(gmask, ci, o);
                    if (od < cand || (od == cand && oi < ci)) { cand = od; ci = oi; }
                }
                if (ci != 0x7fffffff && cand < best) { best = cand; hit = ci; }
            }
            for (int tb = 0; tb < cnt; tb += GS) {
                const int m = tb + g;
                bool rm = false;
                if (m < cnt && m != hit) {
                    const int lf = life[m];
                    rm = lf <= 1;
                    life[m] = rm ? kLifeRemoved : lf - 1;
                }
                removed += __popc(GBALLOT(rm));
            }
        }
        __syncwarp(gmask);
        int hit_new = -1, hit_id = -1;
        double pa = q[0], pb = q[1];
        double hx = __longlong_as_double(0x7ff8000000000000LL), hy = hx;
        if (hit >= 0) {
            if (g == 0) life[hit] = a.life_init;                 // reset_life(); a, b, pos, end NOT updated
            // removals all precede the hit in first-match modes; count them for the nearest mode
            int before = removed;
            if (a.mode == 2) {
                before = 0;
                for (int tb = 0; tb < hit; tb += GS) {
                    const int m = tb + g;
                    before += __popc(GBALLOT(m < hit && life[m] == kLifeRemoved));
                }
            }
            hit_new = hit - before;
            hit_id = ids[hit];
            pa = ent[(size_t)hit * 6];
            pb = ent[(size_t)hit * 6 + 1];
            hx = ent[(size_t)hit * 6 + 2];
            hy = ent[(size_t)hit * 6 + 3];
        }
        __syncwarp(gmask);
        // ---- stable compaction of the removed landmarks (list.remove shifts the tail left) ----
        if (removed > 0) {
            int outp = 0;
            for (int tb = 0; tb < cnt; tb += GS) {
                const int m = tb + g;
                const bool keep = m < cnt && life[m] != kLifeRemoved;
                double e6[6];
                int lf = 0, id = 0;
                if (keep) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) e6[k] = ent[(size_t)m * 6 + k];
                    lf = life[m]; id = ids[m];
                }
                const unsigned km = GBALLOT(keep);
                __syncwarp(gmask);
                if (keep) {
                    const int d = outp + __popc(km & ((1u << g) - 1u));
#pragma unroll
                    for (int k = 0; k < 6; ++k) ent[(size_t)d * 6 + k] = e6[k];
                    life[d] = lf; ids[d] = id;
                }
                outp += __popc(km);
                __syncwarp(gmask);
            }
            cnt = outp;
        }
        int result = hit_new, rid = hit_id;
        if (hit < 0) {
            rid = qid;
            if (a.append && cnt < M) {
                if (g == 0) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) ent[(size_t)cnt * 6 + k] = q[k];
                    life[cnt] = a.life_init;
                    ids[cnt] = qid;
                }
                ++cnt;
                result = -1;
            } else {
                result = a.append ? -3 : -1;
            }
        }
        __syncwarp(gmask);
        if (g == 0) {
            a.match[qo] = result;
            a.match_id[qo] = rid;
            if (a.proj_ab) { a.proj_ab[2 * qo] = pa; a.proj_ab[2 * qo + 1] = pb; }
            if (a.hit_pos) { a.hit_pos[2 * qo] = hx; a.hit_pos[2 * qo + 1] = hy; }
        }
    }
    // ---- write the table back; slots past the live count are zeroed ----
    {
        double *dst = a.table + (size_t)t * M * 6;
        for (int k = g; k < M * 6; k += GS) dst[k] = k < cnt * 6 ? ent[k] : 0.0;
        for (int k = g; k < M; k += GS) {
            a.life[(size_t)t * M + k] = k < cnt ? life[k] : 0;
            a.ids[(size_t)t * M + k] = k < cnt ? ids[k] : 0;
        }
        if (g == 0) { a.count[t] = cnt; a.next_id[t] = next_id; }
    }
#undef GBALLOT
}

} // namespace lslam
