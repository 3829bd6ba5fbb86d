/* lslam.h - C ABI of the B200-native LiDAR_SLAM hot path (liblslam.so).
 *
 * The reference (Farofeiro231/LiDAR_SLAM) is pure Python and has no FFI of its own; its
 * boundary for this path is the Python module API of ransac_functions.py, landmarking.py,
 * UKFMethods.py and systemClass.py.  The drop-in modules in lidar_slam_b200/ keep that
 * API and bind the entry points below with ctypes (INTEGRATION.md shows the stub).  Each
 * entry point cites the reference interface it replaces (file:line in the reference).
 *
 * Conventions
 *   - plain pointers and sizes only; every data pointer is a DEVICE pointer unless its
 *     name ends in _h; the caller owns all buffers; nothing is allocated per call
 *   - `stream` is a cudaStream_t cast to void* (NULL = the legacy default stream); calls
 *     are asynchronous on that stream and never synchronise the device
 *   - return value: LSLAM_OK or a negative lslam_status; lslam_last_error() gives text
 *   - a handle is bound to one device, holds no per-call state and may be used from any
 *     one thread at a time (the reference calls the path from one worker thread,
 *     ransac_functions.py:104); create it AFTER fork (SLAM.py:20 forks the worker)
 *   - all arithmetic is IEEE fp64 like the reference's Python floats; indices are int32
 */
#ifndef LSLAM_H
#define LSLAM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum lslam_status {
    LSLAM_OK = 0,
    LSLAM_ERR_BAD_ARG = -1,        /* null pointer, size out of range, misaligned buffer */
    LSLAM_ERR_TOO_FEW_POINTS = -2, /* chunk slot with <= 2 points (functions.py:71 never queues one) */
    LSLAM_ERR_NO_MODEL = -3,       /* no trial found an inlier: upstream ransac returns (None, None)
                                      and the reference raises AttributeError at ransac_functions.py:25 */
    LSLAM_ERR_CUDA = -4,           /* CUDA runtime error, text in lslam_last_error() */
    LSLAM_ERR_NONFINITE = -5,      /* NaN/Inf (or |coordinate| >= 2^400) in a chunk */
    LSLAM_ERR_CAPACITY = -6        /* problem does not fit (shared memory, table capacity) */
} lslam_status;

/* ---- handle ---------------------------------------------------------------------- */
int lslam_create(int device, void **handle);
int lslam_destroy(void *handle);
const char *lslam_last_error(void *handle);
/* ABI version of the loaded library (LSLAM_ABI_VERSION at build time). */
#define LSLAM_ABI_VERSION 6
int lslam_abi_version(void);
/* Launch-shape tuning knobs; 0 restores the built-in heuristic.  Results never depend on them.
 *   "ransac.kernel"  0 default, 1 phase-structured kernel, 2 warp-specialised persistent kernel, 3 tile kernel
 *   "ransac.tpt"     trials per work item (1, 2, 4, 6, 8);  "ransac.slots_per_cta"  chunk slots per tile
 *   "ransac.scorers", "ransac.posts", "ransac.ring"   warp counts / ring depth of kernel 2
 *   "ransac.threads" CTA size of kernel 1 */
int lslam_set_option(void *handle, const char *key, int64_t value);
/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
int64_t lslam_launch_count(void *handle);

/* ---- input stage: polar -> cartesian ------------------------------------------------
 * Replaces the per-measure conversion of scanning(), functions.py:59-60:
 *   x = d * cos(-angle_deg * pi/180 + pi/2),  y = d * sin(-angle_deg * pi/180 + pi/2)
 * dist[npts] mm, ang_deg[npts] degrees -> xy[npts,2] (x,y interleaved). */
int lslam_polar_to_cart(void *handle, const double *dist, const double *ang_deg, double *xy,
                        int64_t npts, void *stream);

/* ---- RANSAC line / landmark extraction ---------------------------------------------
 * Replaces, per chunk, ransac_functions.py:20-31: skimage.measure.ransac(data,
 * LineModelND, min_samples=2, residual_threshold=thr, max_trials=trials) followed by the
 * slope/intercept/tip computation.  One revolution is cut into chunks the way scanning()
 * cuts it (functions.py:64-76): C = ceil(N/chunk) chunk slots per scan; slot c of scan b
 * holds points [c*chunk, min((c+1)*chunk, npts[b])) and is processed when it is a full
 * chunk or a remainder of more than 2 points.
 *
 *   xy        [B,N,2]  points, sensor frame, mm (x,y interleaved; 16-byte aligned)
 *   npts      [B] valid points per scan (<= N), or NULL for all N
 *   trials    hypotheses per chunk (MAX_TRIALS, ransac_functions.py:10)
 *   thr       inlier distance, strict '<'   (THRESHOLD, ransac_functions.py:9)
 *   seed, scan_base   Philox4x32-10 key and global index of scan 0 of this batch: trial t
 *             of chunk c of scan b samples the pair derived from counter
 *             (t>>1, c, scan_base+b) - see oracle/philox.py; results do not depend on
 *             how a job is split into batches or over GPUs
 *   pairs     NULL = that sampler, or [B,C,trials,2] explicit sample index pairs (8-byte aligned): trial t of
 *             chunk slot (b,c) uses points pairs[b,c,t,0..1] of the chunk - what upstream's
 *             random_state.choice(n, 2, replace=False) drew, so a seeded skimage.measure.ransac run
 *             (ransac_functions.py:23-24 with random_state=s) can be replayed pair for pair; indices are
 *             clamped into the chunk; a chunk upstream left early (sum of squares <= 0) is padded with
 *             repeats of any pair already tried (a repeat never displaces its first occurrence)
 * outputs (per chunk slot, [B,C,...])
 *   line      [B,C,8]  a, b, origin_x, origin_y, dir_x, dir_y, tip_x, tip_y
 *                      (dir normalised to dir_x > 0; a = dir_y/dir_x unguarded like :26)
 *   ninl      [B,C]    inlier count (>= 2), or 0 for an unused slot, or a negative
 *                      lslam_status (LSLAM_ERR_NO_MODEL, LSLAM_ERR_NONFINITE)
 *   win_trial [B,C]    index of the selected hypothesis (first best in trial order)
 *   win_pair  [B,C,2]  its two sample indices within the chunk
 *   ssum      [B,C]    its sum of squared residuals over all points of the chunk; may be NULL
 *                      (not needed by the reference's outputs: the kernel then decides a chunk whose
 *                      winner is unambiguous without the literal residual pass)
 *   inl_mask  [B,C,W]  W = ceil(chunk/32) words, bit k of word w = point 32*w+k is an inlier
 */
int lslam_ransac_lines(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N,
                       int32_t chunk, int32_t trials, double thr, uint64_t seed,
                       uint64_t scan_base, const int32_t *pairs, double *line, int32_t *ninl, int32_t *win_trial,
                       int32_t *win_pair, double *ssum, uint32_t *inl_mask, void *stream);

/* The points landmark_extraction returns (ransac_functions.py:28,46,49,53,55): for every chunk slot with a model,
 * xBase = the x of its inliers in data order, yBase = a' * xBase + b' (a product and a sum, each rounded, like NumPy)
 * with (a', b') = ab[slot*ab_stride + 0..1]: `proj_ab` of lslam_associate (ab_stride 2: the matched STORED landmark's
 * coefficients on a hit, :46, the new line's on a miss, :49,53) or the `line` records themselves (ab_stride 8).
 *   xy, npts, B, N, chunk, inl_mask, ninl   as given to / written by lslam_ransac_lines
 *   offsets [B*C] or NULL  first output row of each slot (e.g. an exclusive prefix sum of max(ninl, 0): compact
 *                          layout); NULL = slot * chunk (padded layout [B,C,chunk,2], rows >= ninl untouched)
 *   points  [rows,2]       (x, y) per inlier, 16-byte aligned */
int lslam_project_inliers(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N, int32_t chunk,
                          const uint32_t *inl_mask, const int32_t *ninl, const double *ab, int32_t ab_stride,
                          const int64_t *offsets, double *points, void *stream);

/* Diagnostics of the last lslam_ransac_lines calls on this handle (synchronises `stream`):
 *   counters_h[0] trials re-scored with the literal formula (threshold band / degenerate pair)
 *   counters_h[1] literal candidate evaluations in the tie-break stage
 *   counters_h[2] consistency faults (fast and literal inlier counts disagree; must be 0)
 * The counters are reset by this call. */
int lslam_ransac_counters(void *handle, int64_t counters_h[3], void *stream);

/* ---- landmark data association ---------------------------------------------------------
 * Replaces the association loop of landmark_extraction (ransac_functions.py:34-54) with
 * Landmark.is_equal / decrease_life / reset_life (landmarking.py:48-77), plus what its caller
 * check_ransac does with the result (ransac_functions.py:74-78): append the new landmark on a
 * miss, advance the extraction counter on every extraction.  T independent streams, each with
 * its own landmark table; the Q queries of a stream are applied IN ORDER (the list they are
 * matched against changes as they go).
 *
 *   query    [T,Q,qstride] doubles; layout 0: a, b, pos_x, pos_y, end_x, end_y (qstride >= 6)
 *            layout 1: a `line` record of lslam_ransac_lines (qstride >= 8; pos = origin,
 *            end = tip, exactly the arguments of Landmark(...) at ransac_functions.py:31)
 *   q_valid  [T,Q] a query is applied iff q_valid > 0 (ninl of lslam_ransac_lines fits), or NULL
 *   table    [T,M,6] in/out, rows a, b, pos_x, pos_y, end_x, end_y in list order
 *   life, ids [T,M] in/out; count [T] in/out live rows; next_id [T] in/out extraction counter
 *   tolA, tolB, tol, life_init   TOLERANCE_A, TOLERANCE_B, TOLERANCE, LIFE (landmarking.py:3-6)
 *   mode     0 first match in list order, reference-exact: a landmark whose life reaches 0 is
 *              removed and the element that slides into its slot is skipped (the reference's
 *              remove-while-indexing behaviour)
 *            1 first match, every element visited
 *            2 nearest gated landmark (smallest of the two gate distances); all others age
 *   append   non-zero: a miss appends the query as a new landmark (id = next_id at the time)
 * outputs
 *   match    [T,Q] row index of the hit in the table as it is after this query's removals;
 *            -1 miss (appended if `append`), -3 miss and table full, -2 query not applied
 *   match_id [T,Q] id of the matched landmark, or the id given to the new one, -1 if not applied
 *   proj_ab  [T,Q,2] or NULL: the (a, b) the reference projects the inliers with
 *            (ransac_functions.py:46 on a hit: the STORED landmark's; :49,53 on a miss: the new)
 *   hit_pos  [T,Q,2] or NULL: get_pos() of the matched STORED landmark (what hx measures against,
 *            UKFMethods.py:30); NaN on a miss
 * Any alignment and any M work.  The fast load path (16-byte asynchronous copies of the table and its queries in one
 * round trip) needs `query`, `table`, `life` and `ids` 16-byte aligned and M a multiple of 4 (any cudaMalloc'ed,
 * densely packed buffer with the reference's LANDMARK_NUMBER = 8 qualifies); results are identical either way.
 */
int lslam_associate(void *handle, const double *query, int32_t qstride, int32_t layout,
                    const int32_t *q_valid, int32_t T, int32_t Q, int32_t M, double *table,
                    int32_t *life, int32_t *ids, int32_t *count, int32_t *next_id, double tolA,
                    double tolB, double tol, int32_t life_init, int32_t mode, int32_t append,
                    int32_t *match, int32_t *match_id, double *proj_ab, double *hit_pos, void *stream);

/* ---- UKF sigma-point predict / update ---------------------------------------------------
 * Replaces filterpy's UnscentedKalmanFilter.predict(u=...) / update(z, landmarks=...) with
 * MerweScaledSigmaPoints as systemClass.System wires them (/root/reference/systemClass.py:20-29)
 * around the callbacks of /root/reference/UKFMethods.py:10-71 (fx = transition_function, hx =
 * transfer_function, circular means state_mean / z_mean, angle-wrapping residual_x / residual_h),
 * for F independent n = 3 filters (state x, y, theta).  Operation order: oracle/ukf.c.
 *
 *   x [F,3], P [F,3,3]   state and covariance, in/out
 *   u [F,2]              wheel speeds (vl, vr) of the predict, NULL = zero
 *   z [F,L,zstride]      measurement slots; zmode 0: (range, bearing) at [0],[1] of a slot
 *                        zmode 1: a `line` record of lslam_ransac_lines - the measurement is the
 *                        range/bearing of its inlier centroid, sqrt(ox^2+oy^2), atan2(oy, ox)
 *   lm_pos [F,L,2]       landmark positions the slots are measured against (Landmark.get_pos(),
 *                        UKFMethods.py:30), e.g. `hit_pos` of lslam_associate
 *   lm_valid [F,L]       slot l is used iff lm_valid >= valid_min (NULL = all); e.g. `match` of
 *                        lslam_associate with valid_min 0.  The used slots, in order, form
 *                        z = [d0, b0, d1, b1, ...] (UKFMethods.py:33) and R = diag([var_dist,
 *                        var_angle] * used) (systemClass.py:28); no used slot = no update
 *   params_h [17] HOST   dt, wheel_r (UKFMethods.R), base_l (UKFMethods.L), alpha, beta, kappa,
 *                        var_dist, var_angle, Q[3][3]   (systemClass.py:8-10,20,28-29)
 *   flags                LSLAM_UKF_PREDICT | LSLAM_UKF_UPDATE | LSLAM_UKF_REDRAW (re-draw the sigma
 *                        points from the prior at the end of predict, filterpy >= 1.4.5)
 *   status [F] or NULL   0, or LSLAM_UKF_ST_* bits: a Cholesky factorisation met a matrix that is
 *                        not positive definite (scipy raises LinAlgError there)
 *   sigmas_f [F,7,3] or NULL  the propagated sigma points: written by a predict, read by an
 *                        update-only call (NULL: re-drawn from x, P)
 *   L <= 32.
 */
#define LSLAM_UKF_PREDICT 1
#define LSLAM_UKF_UPDATE 2
#define LSLAM_UKF_REDRAW 4
#define LSLAM_UKF_ST_SIGMA 1  /* P not positive definite: step not applied                   */
#define LSLAM_UKF_ST_REDRAW 2 /* prior P not positive definite: prior kept, update skipped    */
#define LSLAM_UKF_ST_S 4      /* innovation covariance not positive definite: update skipped  */
int lslam_ukf_step(void *handle, double *x, double *P, const double *u, const double *z, int32_t zstride,
                   int32_t zmode, const double *lm_pos, const int32_t *lm_valid, int32_t valid_min,
                   int32_t F, int32_t L, const double *params_h, int32_t flags, int32_t *status,
                   double *sigmas_f, void *stream);
/* Fused UKF step + all-gather of its results over peer memory (NVLink / NVSwitch): same arguments and
 * arithmetic as lslam_ukf_step, and in the same kernel every filter's (x[3], P[9]) - and, when `ids` is
 * given, its id_slots per-slot landmark ids (lslam_associate's match_id) - is stored into record
 * first_record + f of EVERY rank's gather buffers: state [records,12] double, ids [records,id_slots]
 * int32 (rank r publishing F_r filters per step uses first_record = sum of F over the ranks before it,
 * plus the offset of this call's first filter when a step is issued in slices).
 * Replaces the result gather north_star names (one NCCL all-gather per step) by stores issued while the
 * remaining filters still compute.  Buffers: either the multicast address of the ranks' buffers
 * (mc_state / mc_ids, e.g. torch symmetric memory's multicast_ptr: one store per element, NVSwitch
 * replicates it) or host arrays peer_state_h / peer_ids_h of `world` device addresses: every rank's
 * buffer as mapped on THIS GPU, the local one included.  The caller synchronises the ranks after the
 * kernel (a barrier on the stream) before anyone reads the gathered records.  world <= 8. */
int lslam_ukf_step_gather(void *handle, double *x, double *P, const double *u, const double *z,
                          int32_t zstride, int32_t zmode, const double *lm_pos, const int32_t *lm_valid,
                          int32_t valid_min, int32_t F, int32_t L, const double *params_h, int32_t flags,
                          int32_t *status, const int32_t *ids, int32_t id_slots, int32_t world,
                          int64_t first_record, void *const *peer_state_h, void *mc_state,
                          void *const *peer_ids_h, void *mc_ids, void *stream);

/* filterpy-shaped halves of the same kernel: predict(u=u) and update(z, landmarks=...). */
int lslam_ukf_predict(void *handle, double *x, double *P, const double *u, int32_t F,
                      const double *params_h, int32_t redraw, int32_t *status, double *sigmas_f,
                      void *stream);
int lslam_ukf_update(void *handle, double *x, double *P, const double *z, int32_t zstride, int32_t zmode,
                     const double *lm_pos, const int32_t *lm_valid, int32_t valid_min, int32_t F,
                     int32_t L, const double *params_h, int32_t *status, const double *sigmas_f,
                     void *stream);

/* ---- augmented-state UKF (BASELINE.json configs[3]: pose + L landmarks in the state) ------------
 * The same filterpy engine and UKFMethods callbacks as above (/root/reference/systemClass.py:20-29,
 * UKFMethods.py:10-71) for F independent filters of dimension n = 3 + 2L: x = (x, y, theta, l0x, l0y,
 * l1x, ...), fx moves the pose and leaves the landmarks, hx is UKFMethods.transfer_function with the
 * landmark positions read from the state (oracle/ukf.py transfer_function_augmented), dim_z = 2L,
 * R = diag([var_dist, var_angle] * L), Q = params' 3x3 on the pose block and zero elsewhere,
 * 2n + 1 sigma points.  The gain is applied through the Cholesky factor of S and one forward solve
 * W = U^-T [Pzx | r]: x = xbar + W^T v, P = Pbar - W^T W (filterpy forms inv(S) and K; same x, P to rounding).
 *   x [F,n], P [F,n,n] in/out;  u [F,2] or NULL;  z [F,L,2] (range, bearing) per landmark;
 *   params_h, flags, status as for lslam_ukf_step, with the same resulting state: LSLAM_UKF_ST_SIGMA leaves (x, P) untouched;
 *   after a predict, LSLAM_UKF_ST_REDRAW / LSLAM_UKF_ST_S leave the PRIOR (filterpy's predict() has stored it before the
 *   re-draw, and only update() fails) - bit for bit what a predict-only call returns.
 *   workspace: device scratch; filters are processed in batches of as many as fit
 *   (lslam_ukf_aug_workspace_bytes(L, k) bytes hold k filters; at least one must fit).  1 <= L <= 1024.
 *   The workspace must be 16-byte aligned (its arrays are read through TMA tensor maps).
 */
int lslam_ukf_aug_workspace_bytes(int32_t L, int32_t filters, int64_t *bytes);
int lslam_ukf_aug_step(void *handle, double *x, double *P, const double *u, const double *z, int32_t F,
                       int32_t L, const double *params_h, int32_t flags, int32_t *status, void *workspace,
                       int64_t workspace_bytes, void *stream);

/* The same step with NO allocation anywhere: the scratch is a file-scope static device array of
 * LSLAM_STATIC_SCRATCH_BYTES inside the library (every GPU that loads liblslam.so holds it until the
 * process exits).  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Arguments and results are those of lslam_ukf_aug_step, bit for bit; the caller
 * still passes every input and output.  All F filters must fit the array in ONE batch:
 * lslam_ukf_aug_workspace_bytes(L, F) <= LSLAM_STATIC_SCRATCH_BYTES (e.g. 2 filters at L = 200,
 * 545 at L = 12), otherwise LSLAM_ERR_CAPACITY.  Calls no cudaMalloc / cudaFree / allocating
 * routine.  Every call on a device shares the one array, which keeps the last call's contents
 * (nothing in it has to start at zero: each array is written before it is read); two calls that
 * overlap on the GPU - different streams, or different handles / threads on one device - RACE on
 * it: serialise them on one stream. */
#define LSLAM_STATIC_SCRATCH_BYTES (32u << 20)
int lslam_ukf_aug_step_static(void *handle, double *x, double *P, const double *u, const double *z, int32_t F,
                              int32_t L, const double *params_h, int32_t flags, int32_t *status, void *stream);

/* ---- OLS-flavour RANSAC ------------------------------------------------------------
 * Replaces sklearn.linear_model.RANSACRegressor(residual_threshold=3).fit(x[i:k], y[i:k]) +
 * inlier_mask_ / estimator_.coef_ / predict of /root/reference/line_detection.py:51-73 for every
 * window of `window` (PNT_NBR = 50, :9) consecutive points of every scan.
 *   xy [B,N,2], npts [B] or NULL as for lslam_ransac_lines; C = ceil(N/window) window slots per scan
 *   (the last window of a scan may be short, :53-56).
 *   thr: inlier iff |y - y_hat| <= thr (non-strict); stop_probability: sklearn's (default 0.99);
 *   trials: max_trials (default 100).
 *   pairs: NULL = the library's Philox sampler (seed, scan_base as for lslam_ransac_lines), or
 *          [B,C,trials,2] explicit sample index pairs per trial (what sklearn's RandomState drew).
 * out: fit [B,C,8] = coef, intercept, R^2 score of the accepted trial on its inliers, min x, max x of
 *          the window (the span line_detection.py:66 draws), mean x, mean y of the inliers, 0
 *      ninl [B,C] inlier count, or 0 (empty slot), LSLAM_ERR_TOO_FEW_POINTS (1-point window: sklearn's
 *          "min_samples may not be larger than number of samples" ValueError), LSLAM_ERR_NO_MODEL
 *          (no trial accepted: sklearn's "could not find a valid consensus set" ValueError)
 *      ntrials [B,C] = n_trials_ (trials actually run), best_trial [B,C] (index of the accepted trial)
 *      inl_mask [B,C,ceil(window/32)] = inlier_mask_ as bit words */
int lslam_ransac_ols(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N,
                     int32_t window, int32_t trials, double thr, double stop_probability, uint64_t seed,
                     uint64_t scan_base, const int32_t *pairs, double *fit, int32_t *ninl,
                     int32_t *ntrials, int32_t *best_trial, uint32_t *inl_mask, void *stream);

/* ---- input stage from raw sensor bytes: RPLidar express packets -> cartesian -------
 * Replaces ExpressPacket.decode (/root/reference/lidar.py:66-91), Lidar._process_express_scan
 * (lidar.py:179-187) and the conversion inside scanning() (functions.py:59-60) for a batch of
 * revolutions.  packets [nscans][packets_per_scan][84] bytes (4-byte aligned): one revolution =
 * packets_per_scan consecutive packets; the start angle after a revolution's last packet is its
 * first packet's.  The first `beams` (<= 32*packets_per_scan) measures of each revolution are
 * written to xy [nscans,beams,2]; meas (nullable) [nscans,beams,2] receives (angle_deg, distance_mm),
 * bit-identical to the reference's Python floats; status (nullable) [nscans] receives
 * LSLAM_EXPRESS_BAD_SYNC / _BAD_CHECKSUM bits where the reference raises ValueError (lidar.py:70-77),
 * and the points of a rejected packet are NaN (lslam_ransac_lines then reports LSLAM_ERR_NONFINITE
 * for the chunks they fall in). */
#define LSLAM_EXPRESS_BAD_SYNC 1
#define LSLAM_EXPRESS_BAD_CHECKSUM 2
#define LSLAM_EXPRESS_PACKET_BYTES 84
int lslam_express_to_cart(void *handle, const uint8_t *packets, int64_t nscans, int32_t packets_per_scan,
                          int32_t beams, double *xy, double *meas, int32_t *status, void *stream);

/* Synthetic sensor: revolutions scan_base.. of the random-room scene of lslam_synth_scans, emitted
 * as express packets (uniform rotation, random start phase, +-0.5 degree correction codes, integer
 * millimetre ranges, new-scan flag on the first packet).  packets [nscans][packets_per_scan][84]. */
int lslam_synth_express(void *handle, uint8_t *packets, int64_t nscans, int32_t packets_per_scan,
                        uint64_t seed, uint64_t scan_base, double noise_mm, double dropout, void *stream);

/* ---- synthetic scans (headless replacement for the capture process) ----------------
 * Device twin of lidar_slam_b200/synth.py random_room_scans(): scan s (global index
 * scan_base+s) is a `beams`-beam revolution in a random rectangular room, range noise
 * N(0, noise_mm), `dropout` probability of a zero range, converted like functions.py:59-60.
 * xy [B,beams,2]. */
int lslam_synth_scans(void *handle, double *xy, int32_t B, int32_t beams, uint64_t seed,
                      uint64_t scan_base, double noise_mm, double dropout, void *stream);

/* ---- fp64 issue-rate probe (roofline denominator for the compute-bound kernels) ------
 * Runs a dependent-chain DFMA kernel sized to fill the device for about `ms_target`
 * milliseconds and returns the achieved rate in TFLOP/s (2 flop per DFMA) in *tflops_h.
 * Synchronises the device. */
int lslam_fp64_peak(void *handle, double ms_target, double *tflops_h);

#ifdef __cplusplus
}
#endif
#endif /* LSLAM_H */
