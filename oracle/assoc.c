/* C restatement of the landmark association loop (oracle side).
 *
 * TEST INFRASTRUCTURE (see oracle/__init__.py): the compiled twin of oracle/assoc.py (which is
 * pinned to the real /root/reference/landmarking.py through tests/golden/landmarking_golden.npz);
 * used as the fast CPU baseline of bench.py and checked against assoc.py in
 * tests/test_oracle_assoc.py.  Follows /root/reference/ransac_functions.py:34-54 (the loop, with
 * its remove-while-indexing skip), landmarking.py:48-77 (is_equal, decrease_life, reset_life) and
 * ransac_functions.py:74-78 (append on a miss, extraction counter).  Same argument meaning and
 * output conventions as lslam_associate (include/lslam.h).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static double norm2_pinned(double vx, double vy) { return sqrt(fma(vy, vy, vx * vx)); }

/* landmarking.py:66-77: self = stored row e, other = query q (a, b, px, py, ex, ey) */
static int is_equal(const double *e, const double *q, double tolA, double tolB, double tol, double *dmin)
{
    double distA = fabs(e[0] - q[0]), distB = fabs(e[1] - q[1]);
    double dEO = norm2_pinned(e[4] - q[2], e[5] - q[3]);
    double dOE = norm2_pinned(e[2] - q[4], e[3] - q[5]);
    *dmin = dEO < dOE ? dEO : dOE;
    return distA <= tolA && distB <= tolB && (dEO <= tol || dOE <= tol);
}

static void remove_row(double *rows, int32_t *life, int32_t *ids, int *cnt, int i)
{
    memmove(rows + 6 * i, rows + 6 * (i + 1), sizeof(double) * 6 * (size_t)(*cnt - i - 1));
    memmove(life + i, life + i + 1, sizeof(int32_t) * (size_t)(*cnt - i - 1));
    memmove(ids + i, ids + i + 1, sizeof(int32_t) * (size_t)(*cnt - i - 1));
    --*cnt;
}

static void stream_associate(const double *query, int qstride, int layout, const int32_t *q_valid, int Q, int M,
                             double *rows, int32_t *life, int32_t *ids, int32_t *count, int32_t *next_id,
                             double tolA, double tolB, double tol, int life_init, int mode, int append,
                             int32_t *match, int32_t *match_id, double *proj_ab, double *hit_pos)
{
    int cnt = *count < 0 ? 0 : (*count > M ? M : *count);
    int nid = *next_id;
    for (int qi = 0; qi < Q; ++qi) {
        if (q_valid && q_valid[qi] <= 0) {
            match[qi] = -2; match_id[qi] = -1;
            if (proj_ab) { proj_ab[2 * qi] = NAN; proj_ab[2 * qi + 1] = NAN; }
            if (hit_pos) { hit_pos[2 * qi] = NAN; hit_pos[2 * qi + 1] = NAN; }
            continue;
        }
        const double *qs = query + (size_t)qi * qstride;
        double q[6] = { qs[0], qs[1], qs[2], qs[3], layout ? qs[6] : qs[4], layout ? qs[7] : qs[5] };
        const int qid = nid++;
        int hit = -1;
        double dmin;
        if (mode != 2) {
            int i = 0, equal = 0;                               /* ransac_functions.py:34-43 */
            while (i < cnt && !equal) {
                equal = is_equal(rows + 6 * i, q, tolA, tolB, tol, &dmin);
                if (!equal) {
                    if (life[i] > 0) life[i] -= 1;              /* decrease_life(), landmarking.py:48-52 */
                    if (life[i] == 0) {
                        remove_row(rows, life, ids, &cnt, i);
                        if (mode != 0) --i;                     /* mode 0 keeps the reference's skip */
                    }
                }
                ++i;
            }
            if (equal) hit = i - 1;
        } else {
            double best = INFINITY;                             /* nearest gated landmark, lowest index on ties */
            for (int i = 0; i < cnt; ++i)
                if (is_equal(rows + 6 * i, q, tolA, tolB, tol, &dmin) && dmin < best) { best = dmin; hit = i; }
            int w = 0, newhit = -1;                             /* every other landmark ages; expired ones go */
            for (int i = 0; i < cnt; ++i) {
                if (i == hit) newhit = w;
                else {
                    if (life[i] > 0) life[i] -= 1;
                    if (life[i] == 0) continue;
                }
                if (w != i) { memcpy(rows + 6 * w, rows + 6 * i, 6 * sizeof(double)); life[w] = life[i]; ids[w] = ids[i]; }
                ++w;
            }
            cnt = w;
            hit = newhit;
        }
        double pa = q[0], pb = q[1], hx = NAN, hy = NAN;
        if (hit >= 0) {
            life[hit] = life_init;                              /* reset_life() */
            match[qi] = hit; match_id[qi] = ids[hit];
            pa = rows[6 * hit]; pb = rows[6 * hit + 1]; hx = rows[6 * hit + 2]; hy = rows[6 * hit + 3];
        } else {
            match_id[qi] = qid;
            if (append && cnt < M) {                            /* ransac_functions.py:75-76 */
                memcpy(rows + 6 * cnt, q, sizeof q);
                life[cnt] = life_init; ids[cnt] = qid;
                ++cnt;
                match[qi] = -1;
            } else {
                match[qi] = append ? -3 : -1;
            }
        }
        if (proj_ab) { proj_ab[2 * qi] = pa; proj_ab[2 * qi + 1] = pb; }
        if (hit_pos) { hit_pos[2 * qi] = hx; hit_pos[2 * qi + 1] = hy; }
    }
    memset(rows + 6 * cnt, 0, sizeof(double) * 6 * (size_t)(M - cnt));
    memset(life + cnt, 0, sizeof(int32_t) * (size_t)(M - cnt));
    memset(ids + cnt, 0, sizeof(int32_t) * (size_t)(M - cnt));
    *count = cnt; *next_id = nid;
}

int oracle_associate(const double *query, int32_t qstride, int32_t layout, const int32_t *q_valid, int32_t T,
                     int32_t Q, int32_t M, double *table, int32_t *life, int32_t *ids, int32_t *count,
                     int32_t *next_id, double tolA, double tolB, double tol, int32_t life_init, int32_t mode,
                     int32_t append, int32_t *match, int32_t *match_id, double *proj_ab, double *hit_pos,
                     int32_t nthreads)
{
    if (T < 0 || Q < 0 || M < 1 || !query || !table || !life || !ids || !count || !next_id || !match || !match_id)
        return -1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
#else
    (void)nthreads;
#endif
    for (int t = 0; t < T; ++t)
        stream_associate(query + (size_t)t * Q * qstride, qstride, layout, q_valid ? q_valid + (size_t)t * Q : 0, Q, M,
                         table + (size_t)t * M * 6, life + (size_t)t * M, ids + (size_t)t * M, count + t, next_id + t,
                         tolA, tolB, tol, life_init, mode, append, match + (size_t)t * Q, match_id + (size_t)t * Q,
                         proj_ab ? proj_ab + (size_t)t * Q * 2 : 0, hit_pos ? hit_pos + (size_t)t * Q * 2 : 0);
    return 0;
}
