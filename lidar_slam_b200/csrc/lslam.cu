// lslam.cu - the C ABI of include/lslam.h: handle, argument checks, launch shapes.
// One translation unit; kernels live in the .cuh files next to it.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared
//        -Xcompiler -fPIC  (lidar_slam_b200/_build.py)
#include "lslam_common.cuh"
#include "ransac.cuh"
#include "ransac_ws.cuh"
#include "ransac_tile.cuh"
#include "aux_kernels.cuh"
#include "assoc.cuh"
#include "ukf.cuh"
#include "ukf_aug.cuh"
#include "express.cuh"
#include "ransac_ols.cuh"
#include <cfloat>
#include <vector>
#include <cmath>

#include "../../include/lslam.h"

using namespace lslam;


// ---- handle ------------------------------------------------------------------------------
extern "C" int lslam_abi_version(void) { return LSLAM_ABI_VERSION; }

extern "C" int lslam_create(int device, void **handle)
{
    if (handle == nullptr) return LSLAM_ERR_BAD_ARG;
    *handle = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return LSLAM_ERR_CUDA;
    Handle *h = new Handle();
    h->device = device;
    DeviceGuard g(device);
    cudaError_t e = cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device);
    if (e == cudaSuccess) e = cudaMalloc((void **)&h->counters_d, 8 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->counters_d, 0, 8 * sizeof(unsigned long long));
    if (e == cudaSuccess) {
        // sin/cos of whole degrees and of the 2048 sub-degree steps, rounded from long double
        std::vector<double> tab(kExpressTabDoubles);
        const long double pi = 3.14159265358979323846264338327950288L;
        for (int d = 0; d < 360; ++d) {
            const long double th = (long double)d * pi / 180.0L;
            tab[2 * d] = (double)sinl(th); tab[2 * d + 1] = (double)cosl(th);
        }
        for (int b = 0; b < 2048; ++b) {
            const long double th = (long double)b * pi / (180.0L * 2048.0L);
            tab[720 + 2 * b] = (double)sinl(th); tab[720 + 2 * b + 1] = (double)cosl(th);
        }
        // exact values where the true result is exactly representable (sin 180 = 0, cos 90 = 0, ...)
        for (int q = 0; q < 4; ++q) {
            const double sv[4] = {0.0, 1.0, 0.0, -1.0}, cv[4] = {1.0, 0.0, -1.0, 0.0};
            tab[2 * 90 * q] = sv[q]; tab[2 * 90 * q + 1] = cv[q];
        }
        e = cudaMalloc((void **)&h->express_tab_d, sizeof(double) * kExpressTabDoubles);
        if (e == cudaSuccess) e = cudaMemcpy(h->express_tab_d, tab.data(), sizeof(double) * kExpressTabDoubles, cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) { delete h; return LSLAM_ERR_CUDA; }
    *handle = h;
    return LSLAM_OK;
}

extern "C" int lslam_destroy(void *handle)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    DeviceGuard g(h->device);
    cudaFree(h->counters_d);
    cudaFree(h->express_tab_d);
    delete h;
    return LSLAM_OK;
}

extern "C" const char *lslam_last_error(void *handle)
{
    Handle *h = (Handle *)handle;
    return h ? h->err : "null handle";
}

extern "C" int64_t lslam_launch_count(void *handle)
{
    Handle *h = (Handle *)handle;
    return h ? h->launches : -1;
}

extern "C" int lslam_set_option(void *handle, const char *key, int64_t value)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr || key == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!strcmp(key, "ransac.tpt")) {
        if (value != 0 && value != 1 && value != 2 && value != 4 && value != 6 && value != 8) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.tpt must be 0, 1, 2, 4, 6 or 8");
        h->opt_tpt = (int)value;
    } else if (!strcmp(key, "ransac.slots_per_cta")) {
        if (value < 0 || value > 4096) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.slots_per_cta out of range");
        h->opt_slots = (int)value;
    } else if (!strcmp(key, "ransac.threads")) {
        if (value < 0 || value > 1024 || value % 32) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.threads must be a multiple of 32 <= 1024");
        h->opt_threads = (int)value;
    } else if (!strcmp(key, "ransac.kernel")) {
        if (value < 0 || value > 3) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.kernel: 0 default, 1 phase-structured, 2 warp-specialised, 3 tile");
        h->opt_kernel = (int)value;
    } else if (!strcmp(key, "ransac.scorers") || !strcmp(key, "ransac.posts")) {
        if (value < 0 || value > 29) return fail(h, LSLAM_ERR_BAD_ARG, "%s out of range (0..29)", key);
        (key[7] == 's' ? h->opt_scorers : h->opt_posts) = (int)value;
    } else if (!strcmp(key, "ransac.ring")) {
        if (value < 0 || value == 1 || value > 16) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.ring must be 0 or 2..16");
        h->opt_ring = (int)value;
    } else if (!strcmp(key, "ukf_aug.pzx")) {
        if (value < 0 || value > 1) return fail(h, LSLAM_ERR_BAD_ARG, "ukf_aug.pzx: 1 default (paired + / - points), 0 literal sum over all sigma points");
        h->opt_aug_pzx = (int)value;
    } else {
        return fail(h, LSLAM_ERR_BAD_ARG, "unknown option '%s'", key);
    }
    return LSLAM_OK;
}

// ---- RANSAC ------------------------------------------------------------------------------
template <int TPT, int MAXT, int MINB>
static int launch_ransac(Handle *h, RansacArgs a, int threads_target, cudaStream_t st)
{
    auto kern = ransac_lines_kernel<TPT, MAXT, MINB>;
    a.ipc = (a.trials + TPT - 1) / TPT;
    if (threads_target > MAXT) threads_target = MAXT;
    // chunk slots per CTA: enough work items to fill the CTA, bounded by shared memory
    // (measured on B200: about 3.2 items per thread and 256-thread CTAs, 4 CTAs per SM)
    int G = h->opt_slots ? h->opt_slots : (int)((3.2 * threads_target + a.ipc / 2) / a.ipc);
    if (G < 1) G = 1;
    if ((long long)G > a.nslots) G = (int)a.nslots;
    int threads = 0;
    size_t smem = 0;
    for (;; --G) {
        // threads: split the CTA's work items into equal rounds so the last round is not
        // mostly empty (items = G * ipc; every thread takes one item per round)
        const long long items = (long long)G * a.ipc;
        const long long rounds = (items + threads_target - 1) / threads_target;
        threads = (int)((items + rounds - 1) / rounds);
        threads = (threads + 31) & ~31;
        if (threads > MAXT) threads = MAXT & ~31;
        if (threads < 4 * G) threads = ((4 * G + 31) & ~31) > MAXT ? (MAXT & ~31) : ((4 * G + 31) & ~31);
        smem = ransac_smem_bytes(G, a.chunk, a.trials, a.W, threads / 32);
        if (smem <= (size_t)h->smem_optin || G == 1) break;
    }
    if (smem > (size_t)h->smem_optin)
        return fail(h, LSLAM_ERR_CAPACITY, "chunk=%d with trials=%d needs %zu B of shared memory (> %d)",
                    a.chunk, a.trials, smem, h->smem_optin);
    a.G = G;
    const long long grid = (a.nslots + G - 1) / G;
    if (grid > 0x7fffffffLL) return fail(h, LSLAM_ERR_CAPACITY, "batch too large for one launch");
    LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LSLAM_LAUNCH(kern, (unsigned)grid, (unsigned)threads, smem, st, a);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

// tile kernel (ransac_tile.cuh): one work item per thread, G = threads / ipc chunk slots per CTA.
// Returns LSLAM_ERR_CAPACITY (without touching the handle's error text) when the shape does not fit it.
template <int TPT, int MAXT, int MINB>
static int launch_ransac_tile(Handle *h, RansacArgs a, int threads, cudaStream_t st, bool probe_only = false)
{
    auto kern = ransac_tile_kernel<TPT, MAXT, MINB>;
    a.ipc = (a.trials + TPT - 1) / TPT;
    if (threads > MAXT) threads = MAXT;
    threads &= ~31;
    if (threads < 32 || a.ipc > threads) return LSLAM_ERR_CAPACITY;
    int G = threads / a.ipc;
    if (h->opt_slots && h->opt_slots < G) G = h->opt_slots;
    if ((long long)G > a.nslots) G = (int)a.nslots;
    size_t smem = 0;
    for (;; --G) {
        smem = tile_smem_bytes(G, a.chunk, a.ipc, TPT, a.W, threads / 32);
        if (smem <= (size_t)h->smem_optin || G == 1) break;
    }
    if (smem > (size_t)h->smem_optin) return LSLAM_ERR_CAPACITY;
    if (probe_only) return LSLAM_OK;
    // no more threads than the work items of G slots need (rounded up to warps, at least the lanes of the refit groups)
    int need = G * a.ipc > G * kTileLG ? G * a.ipc : G * kTileLG;
    need = (need + 31) & ~31;
    if (need < threads) threads = need;
    a.G = G;
    a.inv_thr = 1.0 / a.thr;
    LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long grid = (a.nslots + G - 1) / G;
    if (grid > 0x7fffffffLL) return fail(h, LSLAM_ERR_CAPACITY, "batch too large for one launch");
    LSLAM_LAUNCH(kern, (unsigned)grid, (unsigned)threads, smem, st, a);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

static int dispatch_ransac_tile(Handle *h, const RansacArgs &a, int tpt, int threads, cudaStream_t st, bool probe_only = false)
{
    switch (tpt) {
    case 2: return launch_ransac_tile<2, 256, 6>(h, a, threads, st, probe_only);
    case 6: return launch_ransac_tile<6, 256, 4>(h, a, threads, st, probe_only);
    case 8: return launch_ransac_tile<8, 256, 3>(h, a, threads, st, probe_only);
    default: return threads > 512 ? launch_ransac_tile<4, 1024, 1>(h, a, threads, st, probe_only)
                  : threads > 256 ? launch_ransac_tile<4, 512, 2>(h, a, threads, st, probe_only)
                                  : launch_ransac_tile<4, 256, 4>(h, a, threads, st, probe_only);
    }
}

// warp-specialised persistent kernel (ransac_ws.cuh): one CTA per SM
template <int TPT, int MAXT>
static int launch_ransac_ws(Handle *h, RansacArgs a, cudaStream_t st)
{
    auto kern = ransac_ws_kernel<TPT, MAXT>;
    // chunk slots per tile: about 600 work items, bounded by the ring in shared memory (C3 at 4 trials per item: 2 slots,
    // 0.607 ms against 0.62 at 4 and 0.79 at 1 - profiles/r2i_ws_sweep_c3.txt)
    int G = h->opt_slots ? h->opt_slots : (600 + a.ipc / 2) / a.ipc;
    if (G < 2 && !h->opt_slots) G = 2;
    if (G < 1) G = 1;
    if ((long long)G > a.nslots) G = (int)a.nslots;
    const int want_nb = h->opt_ring ? h->opt_ring : 3;
    int NB = 0;
    WsLayout L;
    for (;; --G) {
        for (NB = want_nb; NB >= 2; --NB) {
            L = ws_layout(G, a.chunk, a.trials, a.W, NB, a.ws_np);
            if (L.total <= (size_t)h->smem_optin) break;
        }
        if (NB >= 2 || G == 1) break;
    }
    if (NB < 2)
        return fail(h, LSLAM_ERR_CAPACITY, "chunk=%d with trials=%d does not fit a 2-deep ring in %d B of shared memory",
                    a.chunk, a.trials, h->smem_optin);
    a.G = G;
    a.ws_nb = NB;
    const long long ntiles = (a.nslots + G - 1) / G;
    const long long grid = ntiles < (long long)h->sms ? ntiles : (long long)h->sms;
    const int threads = (a.ws_ns + a.ws_np + 2) * 32;
    LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
    LSLAM_LAUNCH(kern, (unsigned)grid, (unsigned)threads, L.total, st, a);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

// warp-specialised persistent kernel (ransac_ws.cuh): one CTA per SM; register budget by CTA size
template <int TPT>
static int dispatch_ransac_ws(Handle *h, RansacArgs a, cudaStream_t st)
{
    a.ipc = (a.trials + TPT - 1) / TPT;
    // measured (profiles/r2i_ws_sweep_c3.txt, r01e_ransac_sweep_ws.json): big chunks want scorers, small chunks want post warps
    const bool big = (long long)a.chunk * a.trials >= 65536;
    a.ws_ns = h->opt_scorers ? h->opt_scorers : (big ? 14 : 8);
    a.ws_np = h->opt_posts ? h->opt_posts : (big ? 2 : 8);
    const int warps = a.ws_ns + a.ws_np + 2;
    if (warps > 32) return fail(h, LSLAM_ERR_BAD_ARG, "ransac.scorers + ransac.posts must not exceed 30 warps");
    if (warps <= 20) return launch_ransac_ws<TPT, 640>(h, a, st);
    return launch_ransac_ws<TPT, 1024>(h, a, st);
}

extern "C" int lslam_ransac_lines(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N,
                                  int32_t chunk, int32_t trials, double thr, uint64_t seed,
                                  uint64_t scan_base, const int32_t *pairs, double *line, int32_t *ninl, int32_t *win_trial,
                                  int32_t *win_pair, double *ssum, uint32_t *inl_mask, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!xy || !line || !ninl || !win_trial || !win_pair || !inl_mask)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_lines: null buffer");
    if (B < 0 || N < 1 || chunk < 2 || trials < 1 || trials > (1 << 20) || chunk > 8192)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_lines: need B >= 0, N >= 1, 2 <= chunk <= 8192, 1 <= trials <= 2^20");
    if (!(thr == thr)) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_lines: thr is NaN");
    if (((uintptr_t)xy & 15) != 0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_lines: xy must be 16-byte aligned");
    if (B == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    RansacArgs a;
    a.xy = reinterpret_cast<const double2 *>(xy);
    a.npts = npts;
    a.B = B; a.N = N; a.chunk = chunk;
    a.C = (N + chunk - 1) / chunk;
    a.W = (chunk + 31) / 32;
    a.trials = trials; a.thr = thr; a.seed = seed; a.scan_base = scan_base;
    a.line = line; a.ninl = ninl; a.win_trial = win_trial; a.win_pair = win_pair; a.ssum = ssum;
    a.mask = inl_mask;
    a.counters = h->counters_d;
    a.nslots = (long long)B * a.C;
    a.G = 0; a.ipc = 0;
    // default kernel by problem shape (profiles/r01e_ransac_sweep_ws.json): the warp-specialised kernel wins
    // when a chunk carries enough scoring work to amortise its per-chunk stages (C3: 410 points x 1024
    // trials, 22.5 vs 20.6 TFLOP/s); the phase-structured kernel wins on small chunks (C2: 14.3 vs 13.1)
    const bool big_chunks = (long long)chunk * trials >= 65536;
    cudaStream_t st = (cudaStream_t)stream;
    a.ws_ns = a.ws_np = a.ws_nb = 0;
    a.inv_thr = 0.0;
    a.pairs = pairs;
    if (pairs != nullptr && ((uintptr_t)pairs & 7) != 0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_lines: pairs must be 8-byte aligned");
    // tile kernel: 512-thread CTAs (20 chunk slots at 100 trials, two CTAs per SM) measured best on C2: 4.33 against 4.55 ms at 256
    const int threads = h->opt_threads ? h->opt_threads : 512, threads1 = h->opt_threads ? h->opt_threads : 256;
    // small chunks (the reference's 100 x 100): the tile kernel (3) when the shape fits it - one work item per thread,
    // every trial's inlier words in shared memory - else the phase-structured kernel (1)
    int kernel = h->opt_kernel;
    if (kernel == 0)
        kernel = (big_chunks && pairs == nullptr) ? 2 : (dispatch_ransac_tile(h, a, h->opt_tpt ? h->opt_tpt : 4, threads, st, true) == LSLAM_OK ? 3 : 1);
    if (kernel == 3) {
        const int rc = dispatch_ransac_tile(h, a, h->opt_tpt ? h->opt_tpt : 4, threads, st);
        if (rc != LSLAM_ERR_CAPACITY || h->opt_kernel == 0) return rc;
        return fail(h, LSLAM_ERR_CAPACITY, "ransac.kernel = 3 needs ceil(trials / tpt) <= threads and the trials' inlier words in shared memory "
                                           "(chunk=%d, trials=%d)", chunk, trials);
    }
    if (kernel == 2 && pairs != nullptr) kernel = 1;              // the warp-specialised kernel has no explicit-pairs path
    // warp-specialised kernel: 4 trials per work item (measured on C3: 0.607 ms against 0.615 at 2 and 0.69 at 8); phased kernel: 2
    const int tpt = h->opt_tpt ? h->opt_tpt : (kernel == 2 ? 4 : 2);
    if (kernel == 2) {
        switch (tpt) {
        case 2: return dispatch_ransac_ws<2>(h, a, st);
        case 4: return dispatch_ransac_ws<4>(h, a, st);
        default: return dispatch_ransac_ws<8>(h, a, st);
        }
    }
    switch (tpt) {
    case 1: return launch_ransac<1, 512, 2>(h, a, threads1, st);
    case 2: return launch_ransac<2, 512, 2>(h, a, threads1, st);
    default: return launch_ransac<4, 512, 1>(h, a, threads1, st);
    }
}

extern "C" int lslam_project_inliers(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N, int32_t chunk,
                                     const uint32_t *inl_mask, const int32_t *ninl, const double *ab, int32_t ab_stride,
                                     const int64_t *offsets, double *points, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!xy || !inl_mask || !ninl || !ab || !points) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_project_inliers: null buffer");
    if (B < 0 || N < 1 || chunk < 2 || chunk > 8192 || ab_stride < 2)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_project_inliers: need B >= 0, N >= 1, 2 <= chunk <= 8192, ab_stride >= 2");
    if ((((uintptr_t)xy | (uintptr_t)points) & 15) != 0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_project_inliers: xy and points must be 16-byte aligned");
    if (B == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    const int C = (N + chunk - 1) / chunk, W = (chunk + 31) / 32;
    const long long nslots = (long long)B * C;
    const int threads = 256;
    long long grid = (nslots * 32 + threads - 1) / threads;
    const long long cap = (long long)h->sms * 16;                 // grid-stride beyond 16 CTAs per SM
    if (grid > cap) grid = cap;
    LSLAM_LAUNCH(project_inliers_kernel, (unsigned)grid, threads, 0, (cudaStream_t)stream, reinterpret_cast<const double2 *>(xy), npts,
                 nslots, (int)N, (int)chunk, C, W, inl_mask, ninl, ab, (int)ab_stride, (const long long *)offsets,
                 reinterpret_cast<double2 *>(points));
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

extern "C" int lslam_ransac_counters(void *handle, int64_t counters_h[3], void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr || counters_h == nullptr) return LSLAM_ERR_BAD_ARG;
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long tmp[3];
    LSLAM_CUDA_TRY(h, cudaMemcpyAsync(tmp, h->counters_d, sizeof tmp, cudaMemcpyDeviceToHost, st));
    LSLAM_CUDA_TRY(h, cudaMemsetAsync(h->counters_d, 0, sizeof tmp, st));
    LSLAM_CUDA_TRY(h, cudaStreamSynchronize(st));
    for (int k = 0; k < 3; ++k) counters_h[k] = (int64_t)tmp[k];
    return LSLAM_OK;
}

// ---- association --------------------------------------------------------------------------
extern "C" int lslam_associate(void *handle, const double *query, int32_t qstride, int32_t layout,
                               const int32_t *q_valid, int32_t T, int32_t Q, int32_t M, double *table,
                               int32_t *life, int32_t *ids, int32_t *count, int32_t *next_id, double tolA,
                               double tolB, double tol, int32_t life_init, int32_t mode, int32_t append,
                               int32_t *match, int32_t *match_id, double *proj_ab, double *hit_pos,
                               void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!query || !table || !life || !ids || !count || !next_id || !match || !match_id)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_associate: null buffer");
    if (T < 0 || Q < 0 || M < 1 || mode < 0 || mode > 2 || layout < 0 || layout > 1 ||
        qstride < (layout ? 8 : 6) || life_init < 1)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_associate: bad sizes / mode / layout");
    if (T == 0 || Q == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    // lane group per table: the smallest of 8 / 16 / 32 lanes that covers the table in one tile
    const int GS = M <= 8 ? 8 : (M <= 16 ? 16 : 32);
    const int NG = 32 / GS;
    // the table's queries ride along with the table in shared memory while at least four warps still fit
    int stage_q = 1;
    size_t per_table = assoc_smem_per_table(M, Q, qstride, 1);
    if ((size_t)h->smem_optin / (per_table * NG) < 4) { stage_q = 0; per_table = assoc_smem_per_table(M, Q, qstride, 0); }
    int nwarps = (int)((size_t)h->smem_optin / (per_table * NG));
    if (nwarps < 1) return fail(h, LSLAM_ERR_CAPACITY, "lslam_associate: table capacity M=%d does not fit shared memory", M);
    if (nwarps > 8) nwarps = 8;
    AssocArgs a;
    a.query = query; a.q_valid = q_valid; a.qstride = qstride; a.layout = layout;
    a.T = T; a.Q = Q; a.M = M;
    a.table = table; a.life = life; a.ids = ids; a.count = count; a.next_id = next_id;
    a.tolA = tolA; a.tolB = tolB; a.tol = tol; a.tol_sq = assoc_gate_sq(tol); a.life_init = life_init; a.mode = mode; a.append = append;
    a.match = match; a.match_id = match_id; a.proj_ab = proj_ab; a.hit_pos = hit_pos;
    a.stage_q = stage_q;
    // 16-byte asynchronous copies need every per-table block to start and end on 16 bytes
    a.async16 = (M % 4 == 0) && (((size_t)Q * qstride) % 2 == 0 || !stage_q) &&
                ((((uintptr_t)table | (uintptr_t)life | (uintptr_t)ids | (stage_q ? (uintptr_t)query : 0)) & 15) == 0);
    const size_t smem = per_table * NG * nwarps;
    const long long per_cta = (long long)nwarps * NG;
    const unsigned grid = (unsigned)((T + per_cta - 1) / per_cta);
    auto launch = [&](auto kern) -> int {
        LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        LSLAM_LAUNCH(kern, grid, (unsigned)(nwarps * 32), smem, (cudaStream_t)stream, a);
        h->launches++;
        LSLAM_CUDA_TRY(h, cudaGetLastError());
        return LSLAM_OK;
    };
    if (GS == 8) return launch(associate_kernel<8>);
    if (GS == 16) return launch(associate_kernel<16>);
    return launch(associate_kernel<32>);
}

// ---- UKF ------------------------------------------------------------------------------------
template <int G>
static int launch_ukf(Handle *h, const UkfArgs &a, cudaStream_t st)
{
    auto kern = ukf_kernel<G>;
    const int nwarps = G >= 16 ? 2 : 4;
    const size_t smem = ukf_smem_doubles_per_warp(G) * sizeof(double) * nwarps;
    if (smem > (size_t)h->smem_optin) return fail(h, LSLAM_ERR_CAPACITY, "lslam_ukf: shared memory");
    LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long per_cta = (long long)nwarps * 32;        // a warp owns 32 consecutive filters (ukf_kernel: fb)
    const long long grid = (a.F + per_cta - 1) / per_cta;
    if (grid > 0x7fffffffLL) return fail(h, LSLAM_ERR_CAPACITY, "too many filters for one launch");
    LSLAM_LAUNCH(kern, (unsigned)grid, (unsigned)(nwarps * 32), smem, st, a);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

struct UkfGather {
    int world = 0, id_slots = 0;
    long long first_record = 0;
    const int32_t *ids = nullptr;
    void *const *peer_state_h = nullptr; void *mc_state = nullptr;
    void *const *peer_ids_h = nullptr; void *mc_ids = nullptr;
};

static int ukf_step_impl(void *handle, double *x, double *P, const double *u, const double *z,
                         int32_t zstride, int32_t zmode, const double *lm_pos, const int32_t *lm_valid,
                         int32_t valid_min, int32_t F, int32_t L, const double *params_h, int32_t flags,
                         int32_t *status, double *sigmas_f, const UkfGather &gd, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!x || !P || !params_h) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step: null buffer");
    if (F < 0 || L < 0 || L > 32 || (flags & ~7) || !(flags & 3))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step: need F >= 0, 0 <= L <= 32, flags of PREDICT|UPDATE|REDRAW");
    const bool upd = (flags & LSLAM_UKF_UPDATE) && L > 0;
    if (upd && (!z || !lm_pos || zmode < 0 || zmode > 1 || zstride < (zmode ? 4 : 2)))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step: update needs z, lm_pos and a zstride that fits zmode");
    if (F == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    UkfArgs a;
    a.x = x; a.P = P; a.u = u; a.z = z; a.lm_pos = lm_pos; a.lm_valid = lm_valid; a.status = status;
    a.sigmas_f = sigmas_f;
    a.F = F; a.L = upd ? L : 0; a.zstride = zstride; a.zmode = zmode; a.valid_min = valid_min;
    a.flags = upd ? flags : (flags & ~LSLAM_UKF_UPDATE);
    a.gather_world = gd.world; a.gather_first = gd.first_record; a.gather_L = gd.id_slots; a.gather_ids = gd.ids;
    a.gather_state_mc = (double *)gd.mc_state; a.gather_ids_mc = (int32_t *)gd.mc_ids;
    for (int p = 0; p < kUkfMaxPeers; ++p) {
        a.gather_state[p] = (p < gd.world && gd.peer_state_h) ? (double *)gd.peer_state_h[p] : nullptr;
        a.gather_idbuf[p] = (p < gd.world && gd.peer_ids_h) ? (int32_t *)gd.peer_ids_h[p] : nullptr;
    }
    if (gd.world > 0 && !(a.flags & 3))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: nothing to run (update without slots)");
    if (!(a.flags & 3)) {                          // update-only call without slots: nothing to do
        if (status) LSLAM_CUDA_TRY(h, cudaMemsetAsync(status, 0, sizeof(int32_t) * (size_t)F, (cudaStream_t)stream));
        return LSLAM_OK;
    }
    const double dt = params_h[0], wheel_r = params_h[1], base_l = params_h[2];
    const double alpha = params_h[3], beta = params_h[4], kappa = params_h[5];
    // weights exactly as MerweScaledSigmaPoints._compute_weights (and oracle/ukf.c) round them
    const int n = 3;
    const double lam = alpha * alpha * (n + kappa) - n;
    const double c = n + lam;
    a.dt = dt; a.half_r = wheel_r / 2.0; a.r_over_l = wheel_r / base_l; a.cscale = c;
    a.wi = 0.5 / c; a.wm0 = lam / c; a.wc0 = lam / c + (1 - alpha * alpha + beta);
    a.var_dist = params_h[6]; a.var_angle = params_h[7];
    for (int i = 0; i < 9; ++i) a.Q[i] = params_h[8 + i];
    if (!(c == c) || c == 0.0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step: alpha/kappa give n + lambda = 0");
    cudaStream_t st = (cudaStream_t)stream;
    const int Lk = a.L;
    if (Lk <= 4) return launch_ukf<4>(h, a, st);
    if (Lk <= 8) return launch_ukf<8>(h, a, st);
    if (Lk <= 16) return launch_ukf<16>(h, a, st);
    return launch_ukf<32>(h, a, st);
}

extern "C" int lslam_ukf_step(void *handle, double *x, double *P, const double *u, const double *z,
                              int32_t zstride, int32_t zmode, const double *lm_pos, const int32_t *lm_valid,
                              int32_t valid_min, int32_t F, int32_t L, const double *params_h, int32_t flags,
                              int32_t *status, double *sigmas_f, void *stream)
{
    return ukf_step_impl(handle, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params_h, flags,
                         status, sigmas_f, UkfGather(), stream);
}

extern "C" int lslam_ukf_step_gather(void *handle, double *x, double *P, const double *u, const double *z,
                                     int32_t zstride, int32_t zmode, const double *lm_pos, const int32_t *lm_valid,
                                     int32_t valid_min, int32_t F, int32_t L, const double *params_h, int32_t flags,
                                     int32_t *status, const int32_t *ids, int32_t id_slots, int32_t world,
                                     int64_t first_record, void *const *peer_state_h, void *mc_state,
                                     void *const *peer_ids_h, void *mc_ids, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (world < 1 || world > kUkfMaxPeers || first_record < 0 || id_slots < 0)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: need 1 <= world <= 8, first_record >= 0");
    if (!mc_state && !peer_state_h) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: no gather buffers");
    if (ids && id_slots > 0 && !mc_ids && !peer_ids_h)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: ids given without id gather buffers");
    for (int p = 0; p < world; ++p) {
        if (!mc_state && !peer_state_h[p]) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: null peer buffer");
        if (ids && id_slots > 0 && !mc_ids && !peer_ids_h[p])
            return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_step_gather: null peer id buffer");
    }
    UkfGather gd;
    gd.world = world; gd.first_record = first_record; gd.id_slots = (ids && id_slots > 0) ? id_slots : 0; gd.ids = gd.id_slots ? ids : nullptr;
    gd.peer_state_h = peer_state_h; gd.mc_state = mc_state; gd.peer_ids_h = peer_ids_h; gd.mc_ids = mc_ids;
    return ukf_step_impl(handle, x, P, u, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params_h, flags,
                         status, nullptr, gd, stream);
}

extern "C" int lslam_ukf_predict(void *handle, double *x, double *P, const double *u, int32_t F,
                                 const double *params_h, int32_t redraw, int32_t *status, double *sigmas_f,
                                 void *stream)
{
    return lslam_ukf_step(handle, x, P, u, nullptr, 2, 0, nullptr, nullptr, 0, F, 0, params_h,
                          LSLAM_UKF_PREDICT | (redraw ? LSLAM_UKF_REDRAW : 0), status, sigmas_f, stream);
}

extern "C" int lslam_ukf_update(void *handle, double *x, double *P, const double *z, int32_t zstride,
                                int32_t zmode, const double *lm_pos, const int32_t *lm_valid, int32_t valid_min,
                                int32_t F, int32_t L, const double *params_h, int32_t *status,
                                const double *sigmas_f, void *stream)
{
    return lslam_ukf_step(handle, x, P, nullptr, z, zstride, zmode, lm_pos, lm_valid, valid_min, F, L, params_h,
                          LSLAM_UKF_UPDATE, status, const_cast<double *>(sigmas_f), stream);
}

// ---- augmented-state UKF ------------------------------------------------------------------------
namespace {
struct AugLayout {
    long long per_filter;           // doubles
    long long U, X, Y, H, Kt, Pbar, Dinv, xbar, xnew, zhat;   // sizes in doubles per filter (all even: 16-byte aligned arrays)
};
AugLayout aug_layout(int L)
{
    // arrays with n columns are stored with the even row pitch ldn = n + 1 (TMA needs 16-byte pitches)
    const long long n = 3 + 2LL * L, ldn = n + 1, m = 2 * n + 1, dz = 2LL * L;
    AugLayout y;
    y.U = n * ldn; y.X = m * ldn; y.Y = m * ldn; y.H = m * dz; y.Kt = dz * ldn;
    y.Pbar = n * ldn; y.xbar = ldn; y.xnew = ldn; y.zhat = dz;
    y.Dinv = ((n + 31) / 32) * 1024;                              // inverted 32 x 32 diagonal blocks of the factor in progress
    y.per_filter = y.U + y.X + y.Y + y.H + y.Kt + y.Pbar + y.Dinv + y.xbar + y.xnew + y.zhat;
    return y;
}
constexpr int kAugMaxBatch = 32768;

// tensor map of a k-major operand: element (col, row, filter) at base[filter * sF + row * ld + col]; the
// box is box_cols x kAugBK x 1.  Columns >= cols, rows >= rows read as zero.
#ifndef LSLAM_EMU
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}
#endif
int tmap_encode_3d(Handle *h, Tmap *map, const double *base, long long cols, long long rows, long long nf, long long ld,
                   long long sF, int box_cols)
{
#ifndef LSLAM_EMU
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return fail(h, LSLAM_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    if (((uintptr_t)base & 15) || (ld & 1) || (sF & 1))
        return fail(h, LSLAM_ERR_BAD_ARG, "TMA operand needs a 16-byte aligned base and even pitches (base %p, ld %lld, stride %lld)",
                    (const void *)base, ld, sF);
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)(rows > 0 ? rows : 1), (cuuint64_t)nf};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 8, (cuuint64_t)sF * 8};
    const cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)kAugBK, 1}, es[3] = {1, 1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (void *)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(h, LSLAM_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
#else
    (void)h;
    map->base = base; map->dim[0] = cols; map->dim[1] = rows; map->dim[2] = nf; map->stride[0] = ld; map->stride[1] = sF;
    map->box[0] = box_cols; map->box[1] = kAugBK; map->box[2] = 1;
#endif
    return LSLAM_OK;
}

struct AugRun {
    Handle *h; cudaStream_t st; int F;
    double *Dinv = nullptr; long long sDinv = 0;                  // inverted diagonal blocks (aug_chol_diag_kernel), [F][blocks][32][32]
    template <int TM, int TN, int WM, int WN, int S = 4, bool SOLVE = false, int MINB = 1>
    int gemm_tma(const AugGemm &g, bool &attr_set)
    {
        auto kern = aug_gemm_tma_kernel<TM, TN, WM, WN, S, SOLVE, MINB>;
        constexpr size_t smem = aug_tma_smem<TM, TN, WN, S, SOLVE>();
        if (!attr_set) {
            LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set = true;
        }
        Tmap ma, mb;
        int rc = tmap_encode_3d(h, &ma, g.A, g.M, g.K, F, g.lda, g.sA, TM + 4);
        if (rc == LSLAM_OK) rc = tmap_encode_3d(h, &mb, g.B, g.N, g.K, F, g.ldb, g.sB, TN + 4);
        if (rc != LSLAM_OK) return rc;
        LSLAM_LAUNCH(kern, dim3((g.N + TN - 1) / TN, (g.M + TM - 1) / TM, F), 32 * (WM * WN + 1), smem, st, g, ma, mb);
        return LSLAM_OK;
    }
    // epilogue extras of the next gemm() call (reset by it): noise 1 = + Q (nq[9]) on the pose block, 2 = + R (nq[0], nq[1])
    // on the diagonal; C2 = a second, scaled copy of the result
    int noise = 0; double nq[9] = {0}; double *C2 = nullptr; long long sC2 = 0; int ldc2 = 0; double scale2 = 0.0; int ktri = 0, mirror = 0;
    int gemm(const double *A, long long sA, int lda, const double *B, long long sB, int ldb, double *C, long long sC, int ldc,
             int M, int N, int K, double w0, double alpha, double beta, const int32_t *skip, int upper)
    {
        if (M <= 0 || N <= 0) return LSLAM_OK;
        AugGemm g;
        g.A = A; g.B = B; g.C = C; g.sA = sA; g.sB = sB; g.sC = sC; g.M = M; g.N = N; g.K = K;
        g.lda = lda; g.ldb = ldb; g.ldc = ldc; g.alpha = alpha; g.beta = beta; g.w0 = w0; g.skip = skip; g.upper = upper;
        g.noise = noise; for (int i = 0; i < 9; ++i) g.nq[i] = nq[i];
        g.C2 = C2; g.sC2 = sC2; g.ldc2 = ldc2; g.scale2 = scale2; g.Dinv = nullptr; g.sDinv = 0;
        g.ktri = ktri; g.mirror = mirror;
        noise = 0; C2 = nullptr; ktri = 0; mirror = 0;
        // CTA tile of the big products: 32 x 64, eight consumer warps of 32 x 8, FOUR CTAs per SM (53 KB, 55 registers) = 32 DMMA
        // warps per SM: 29.6 ms per 1,024 filters.  Measured and dropped: 64 x 64 tiles with 4 or 8 warps
        // (30.8 - 33.9 ms), 128 x 128 (one CTA per SM: slower, L2 was never the limit), 64 x 128, 5 CTAs per SM with 3 stages.
        {
            const int rc = M <= 32 ? gemm_tma<32, 64, 1, 4>(g, h->aug_attr_set[0])
                                   : gemm_tma<32, 64, 1, 8, 4, false, 4>(g, h->aug_attr_set[8]);
            if (rc != LSLAM_OK) return rc;
        }
        h->launches++;
        return LSLAM_OK;
    }
    // block-row step of a blocked forward solve on the tensor pipe (SOLVE kernel):
    //   C[0:nb, 0:N] = Dinv_blk^T (C - A^T B),  A: K x nb, B: K x N (k-major), Dinv_blk = block `blk` of this->Dinv
    int gemm_solve(const double *A, long long sA, int lda, const double *B, long long sB, int ldb, double *C, long long sC, int ldc,
                   int nb, int N, int K, int blk, const int32_t *skip)
    {
        if (nb <= 0 || N <= 0) return LSLAM_OK;
        AugGemm g;
        g.A = A; g.B = B; g.C = C; g.sA = sA; g.sB = sB; g.sC = sC; g.M = nb; g.N = N; g.K = K;
        g.lda = lda; g.ldb = ldb; g.ldc = ldc; g.alpha = -1.0; g.beta = 1.0; g.w0 = 0.0; g.skip = skip; g.upper = 0;
        g.noise = 0; for (int i = 0; i < 9; ++i) g.nq[i] = 0.0;
        g.C2 = nullptr; g.sC2 = 0; g.ldc2 = 0; g.scale2 = 0.0;
        g.Dinv = Dinv + (size_t)blk * 1024; g.sDinv = sDinv; g.ktri = 0; g.mirror = 0;
        // EIGHT consumer warps per 32 x 64 tile (warp tile 32 x 8), the V tiles parked in the operand ring: 4 CTAs per SM.
        // Measured and dropped: four warps of 32 x 16 (+ 0.8 ms per 1,024 filters), 32 x 128 tiles (+ 6 %).
        const int rc = gemm_tma<32, 64, 1, 8, 3, true, 4>(g, h->aug_attr_set[7]);
        if (rc == LSLAM_OK) h->launches++;
        return rc;
    }
    // in-place upper Cholesky of F matrices (dim x dim, row pitch ld, stride sA), LEFT-looking: block row jb
    // takes the contributions of all the rows above it in products of depth K = jb (each element of the
    // factor is written once; the right-looking form did dim/32 rank-32 updates of the whole trailing matrix):
    //   diagonal block  D = A_jj - U[0:jb, jb:jb+nb]^T U[0:jb, jb:jb+nb] = Ud^T Ud, Dinv = Ud^-1   (aug_chol_diag_kernel)
    //   block row       U[jb:jb+nb, jb+nb:] = Dinv^T (A[jb:jb+nb, jb+nb:] - U[0:jb, jb:jb+nb]^T U[0:jb, jb+nb:])   (SOLVE GEMM)
    // The inverted diagonal blocks stay in this->Dinv for a following solve with the same factor.
    int chol(double *A, long long sA, int ld, int dim, int32_t *status, int fail_bit)
    {
        for (int jb = 0; jb < dim; jb += 32) {
            const int nb = dim - jb < 32 ? dim - jb : 32;
            Tmap mu;                                              // the block's 32 columns of the rows above it
            int rc0 = tmap_encode_3d(h, &mu, A + jb, nb, jb, F, ld, sA, 36);
            if (rc0 != LSLAM_OK) return rc0;
            LSLAM_LAUNCH(aug_chol_diag_kernel, dim3(F), 160, kAugDiagSmem, st, A, sA, ld, jb, nb, Dinv + (size_t)(jb / 32) * 1024, sDinv,
                         status, fail_bit, mu);
            h->launches++;
            const int rc = gemm_solve(A + jb, sA, ld, A + jb + nb, sA, ld, A + (size_t)jb * ld + jb + nb, sA, ld, nb, dim - jb - nb, jb,
                                      jb / 32, status);
            if (rc != LSLAM_OK) return rc;
        }
        return LSLAM_OK;
    }
    // W = U^-T W in place (U: dim x dim factor just produced by chol(), W: dim x ncols), block rows top down
    int forward_solve(const double *U, long long sU, int ldu, int dim, double *W, long long sW, int ldw, int ncols, const int32_t *status)
    {
        for (int jb = 0; jb < dim; jb += 32) {
            const int nb = dim - jb < 32 ? dim - jb : 32;
            const int rc = gemm_solve(U + jb, sU, ldu, W, sW, ldw, W + (size_t)jb * ldw, sW, ldw, nb, ncols, jb, jb / 32, status);
            if (rc != LSLAM_OK) return rc;
        }
        return LSLAM_OK;
    }
};
#define AUG_TRY(expr) do { const int rc_ = (expr); if (rc_ != LSLAM_OK) return rc_; } while (0)
} // namespace

extern "C" int lslam_ukf_aug_workspace_bytes(int32_t L, int32_t filters, int64_t *bytes)
{
    if (L < 1 || L > 1024 || filters < 1 || !bytes) return LSLAM_ERR_BAD_ARG;
    const AugLayout y = aug_layout(L);
    const long long n = 3 + 2LL * L;
    *bytes = (int64_t)sizeof(double) * (y.per_filter * filters + 2 * (2 * n + 2));
    return LSLAM_OK;
}

extern "C" int lslam_ukf_aug_step(void *handle, double *x, double *P, const double *u, const double *z, int32_t F,
                                  int32_t L, const double *params_h, int32_t flags, int32_t *status, void *workspace,
                                  int64_t workspace_bytes, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (!x || !P || !params_h || !status || !workspace)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step: null buffer (x, P, params_h, status, workspace)");
    if (F < 0 || L < 1 || L > 1024 || (flags & ~7) || !(flags & 3))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step: need F >= 0, 1 <= L <= 1024, flags of PREDICT|UPDATE|REDRAW");
    const bool pred = flags & LSLAM_UKF_PREDICT, upd = flags & LSLAM_UKF_UPDATE;
    if (upd && !z) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step: update needs z");
    if ((uintptr_t)workspace & 15) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step: workspace must be 16-byte aligned");
    if (F == 0) return LSLAM_OK;
    const int n = 3 + 2 * L, ldn = n + 1, m = 2 * n + 1, dz = 2 * L;
    const AugLayout lay = aug_layout(L);
    const long long avail = workspace_bytes / (long long)sizeof(double) - 2LL * (m + 1);
    long long cap = avail > 0 ? avail / lay.per_filter : 0;
    if (cap < 1) return fail(h, LSLAM_ERR_CAPACITY, "lslam_ukf_aug_step: workspace holds no filter (lslam_ukf_aug_workspace_bytes)");
    if (cap > kAugMaxBatch) cap = kAugMaxBatch;
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const double alpha = params_h[3], beta = params_h[4], kappa = params_h[5];
    const double lam = alpha * alpha * (n + kappa) - n;
    const double c = n + lam;
    if (!(c == c) || c == 0.0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step: alpha/kappa give n + lambda = 0");
    double *ws = (double *)workspace;
    double *wm = ws, *wc = ws + m + 1;
    ws += 2 * (m + 1);
    const double wi = 0.5 / c, wc0 = lam / c + (1 - alpha * alpha + beta);
    LSLAM_LAUNCH(aug_weights_kernel, dim3((m + 255) / 256), 256, 0, st, wm, wc, m, wi, lam / c, wc0);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaMemsetAsync(status, 0, sizeof(int32_t) * (size_t)F, st));
    const unsigned ew = (unsigned)(2 * h->sms);                 // elementwise grids: x-blocks per filter
    for (long long f0 = 0; f0 < F; f0 += cap) {
        const int Fb = (int)(F - f0 < cap ? F - f0 : cap);
        AugArgs a;
        a.F = Fb; a.L = L; a.n = n; a.m = m; a.dz = dz; a.ldn = ldn;
        a.x = x + (size_t)f0 * n; a.P = P + (size_t)f0 * n * n;
        a.u = u ? u + (size_t)f0 * 2 : nullptr; a.z = z ? z + (size_t)f0 * dz : nullptr;
        a.status = status + f0;
        double *p = ws;
        a.U = p; p += lay.U * Fb; a.X = p; p += lay.X * Fb; a.Y = p; p += lay.Y * Fb; a.H = p; p += lay.H * Fb;
        a.Kt = p; p += lay.Kt * Fb;
        a.Pbar = p; p += lay.Pbar * Fb;
        double *dinv = p; p += lay.Dinv * Fb;
        a.xbar = p; p += lay.xbar * Fb;
        p += lay.xnew * Fb;                                       // (spare vector; the update writes x in place)
        a.zhat = p; a.wm = wm; a.wc = wc;
        a.dt = params_h[0]; a.half_r = params_h[1] / 2.0; a.r_over_l = params_h[1] / params_h[2]; a.cscale = c;
        a.var_dist = params_h[6]; a.var_angle = params_h[7];
        for (int i = 0; i < 9; ++i) a.Q[i] = params_h[8 + i];
        AugRun run{h, st, Fb};
        run.Dinv = dinv; run.sDinv = lay.Dinv;
        const long long nn = (long long)n * n, nl = (long long)n * ldn, ml = (long long)m * ldn;
        const dim3 gew(Fb < 64 ? ew : 8, Fb), gcol((n + 31) / 32, Fb);
        const bool redraw = pred && upd && (flags & LSLAM_UKF_REDRAW);
        // paired cross-covariance (ukf_aug.pzx = 1, default): when the residuals Y are those of sigma points about their own
        // centre (re-draw, or no predict), Pzx = wi sum_r (e_{1+r} - e_{1+n+r}) y_{1+r}^T - half the depth, and only the
        // + half of Y is needed (see the update below)
        const bool paired = h->opt_aug_pzx != 0 && (redraw || !pred);
        // sigma points of (x, P): U^T U = (n + lambda) P.  The points themselves are not stored: every consumer
        // recomputes them from (x, U) (aug_sigma_value)
        LSLAM_LAUNCH(aug_scale_upper_kernel, gew, 256, 0, st, (const double *)a.P, nn, n, a.U, nl, ldn, n, c, (const int32_t *)a.status);
        h->launches++;
        AUG_TRY(run.chol(a.U, nl, ldn, n, a.status, LSLAM_UKF_ST_SIGMA));
        const double *xs = a.x;                                   // centre of the sigma points hx sees, and its stride
        int xs_stride = n;
        bool hx_from_U = true;
        if (pred) {
            // fx, circular mean, residuals in one pass over the factor; X kept only if hx needs the fx'd points
            LSLAM_LAUNCH(aug_sigma_resid_kernel, gcol, 256, kAugMeanSmem, st, a, (const double *)a.x, n, 1, 0, (upd && !redraw) ? a.X : nullptr, m);
            h->launches++;
            // Pbar = Y^T diag(Wc) Y + Q; with a re-draw the same epilogue also writes (n + lambda) Pbar, the matrix
            // the next Cholesky factors, into U
            run.noise = 1; for (int i = 0; i < 9; ++i) run.nq[i] = a.Q[i];
            if (redraw) { run.C2 = a.U; run.sC2 = nl; run.ldc2 = ldn; run.scale2 = c; }
            AUG_TRY(run.gemm(a.Y, ml, ldn, a.Y, ml, ldn, a.Pbar, nl, ldn, n, n, m, wc0 - wi, wi, 0.0, a.status, 1));
            if (redraw) {
                AUG_TRY(run.chol(a.U, nl, ldn, n, a.status, LSLAM_UKF_ST_REDRAW));
                LSLAM_LAUNCH(aug_sigma_resid_kernel, gcol, 256, kAugMeanSmem, st, a, (const double *)a.xbar, ldn, 0, 1, (double *)nullptr, paired ? n + 1 : m);
                h->launches++;
                xs = a.xbar; xs_stride = ldn;
            } else {
                hx_from_U = false;
            }
        } else {
            LSLAM_CUDA_TRY(h, cudaMemcpy2DAsync(a.xbar, sizeof(double) * ldn, a.x, sizeof(double) * n, sizeof(double) * n, (size_t)Fb,
                                                cudaMemcpyDeviceToDevice, st));
            LSLAM_CUDA_TRY(h, cudaMemcpy2DAsync(a.Pbar, sizeof(double) * ldn, a.P, sizeof(double) * n, sizeof(double) * n, (size_t)Fb * n,
                                                cudaMemcpyDeviceToDevice, st));
            LSLAM_LAUNCH(aug_sigma_resid_kernel, gcol, 256, kAugMeanSmem, st, a, (const double *)a.x, n, 0, 1, (double *)nullptr, upd && paired ? n + 1 : m);
            h->launches++;
        }
        const double *xres = a.xbar;
        if (upd) {
            const int n1 = ldn;                                   // W: dz x (n + 1), right-hand sides [Pzx | r]
            const long long zn = (long long)dz * n1, mz = (long long)m * dz;
            LSLAM_LAUNCH(aug_hx_kernel, gew, 256, 0, st, a, xs, xs_stride, hx_from_U ? 1 : 0);
            LSLAM_LAUNCH(aug_meas_resid_kernel, dim3((L + 31) / 32, Fb), 256, kAugMeasSmem, st, a,
                         paired ? a.X : (double *)nullptr);     // D = E+ - E- parks in X (unused when the points are not stored)
            LSLAM_LAUNCH(aug_innovation_kernel, dim3(Fb), 256, 0, st, a);
            h->launches += 3;
            // S = E^T diag(Wc) E + R straight into U (pitch dz; a.U keeps its n * ldn stride per filter), where its
            // Cholesky is taken in place - S itself is not needed again (W-form update)
            run.noise = 2; run.nq[0] = a.var_dist; run.nq[1] = a.var_angle;
            AUG_TRY(run.gemm(a.H, mz, dz, a.H, mz, dz, a.U, nl, dz, dz, dz, m, wc0 - wi, wi, 0.0, a.status, 1));   // S (upper)
            // Pzx = E^T diag(Wc) Y.  When Y holds residuals of points about their own centre (re-draw, or no predict) its rows
            // are 0, +T, -T with T upper triangular and exact zeros below.
            if (paired) {
                // The - point of factor row r has the residual -y_{1+r} up to one rounding of x (y = fl(fl(x + u) - x) against
                // fl(fl(x - u) - x): both are u to within ulp(x)), so  Pzx = wi sum_r (e_{1+r} - e_{1+n+r}) y_{1+r}^T = wi D^T T:
                // depth n instead of 2n + 1, the central point drops out (y_0 = 0), and a column tile walks rows [0, j0 + TN) of T.
                // Differs from the literal sum by O(ulp(x) |e|) per term - the size of the rounding noise the literal sum carries.
                run.ktri = -1;
                AUG_TRY(run.gemm(a.X, (long long)n * dz, dz, a.Y + ldn, ml, ldn, a.Kt, zn, n1, dz, n, n, 0.0, wi, 0.0, a.status, 0));
            } else {
                if (hx_from_U) run.ktri = n;                      // literal sum, k-tiles that are all zero skipped
                AUG_TRY(run.gemm(a.H, mz, dz, a.Y, ml, ldn, a.Kt, zn, n1, dz, n, m, wc0 - wi, wi, 0.0, a.status, 0));
            }
            AUG_TRY(run.chol(a.U, nl, dz, dz, a.status, LSLAM_UKF_ST_S));
            AUG_TRY(run.forward_solve(a.U, nl, dz, dz, a.Kt, zn, n1, n1, a.status));   // U^T W = [Pzx | r]
            // x = xbar + W^T v and P = Pbar - W^T W go straight into the caller's arrays (the epilogue stores both
            // triangles); a filter that failed a factorisation is skipped by both kernels and keeps its state
            LSLAM_LAUNCH(aug_update_x_kernel, dim3((n + 31) / 32, Fb), 256, kAugUpdateXSmem, st, a, (double *)nullptr);
            h->launches++;
            run.mirror = 1; run.C2 = a.P; run.sC2 = nn; run.ldc2 = n; run.scale2 = 1.0;
            AUG_TRY(run.gemm(a.Kt, zn, n1, a.Kt, zn, n1, a.Pbar, nl, ldn, n, n, dz, 0.0, -1.0, 1.0, a.status, 1));
            if (pred) {                                           // failed re-draw / S after a predict: the prior stands
                LSLAM_LAUNCH(aug_commit_prior_kernel, gew, 256, 0, st, a);
                h->launches++;
            }
        } else {
            LSLAM_LAUNCH(aug_commit_kernel, gew, 256, 0, st, a, xres, (const double *)a.Pbar);
            h->launches++;
        }
        LSLAM_CUDA_TRY(h, cudaGetLastError());
    }
    return LSLAM_OK;
}

// ---- augmented-state UKF without any allocation: scratch in a static device array ---------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the same
// step as lslam_ukf_aug_step with its workspace in the file-scope array below (LSLAM_STATIC_SCRATCH_BYTES, held by every
// GPU that loads the library until the process exits).  In host code the array's NAME is a host-side stand-in, not a device
// address: the address the kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[LSLAM_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticScratchMaxDevices = 64;
void *g_static_scratch_addr[kStaticScratchMaxDevices] = {};     // device address of g_static_scratch on device d, or null

// device address of g_static_scratch on the handle's device (which must be current); no allocation
int static_scratch_address(Handle *h, void **out)
{
    if (h->device < 0 || h->device >= kStaticScratchMaxDevices)
        return fail(h, LSLAM_ERR_CAPACITY, "static scratch: device ordinal %d beyond %d", h->device, kStaticScratchMaxDevices);
    void *p = __atomic_load_n(&g_static_scratch_addr[h->device], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(h, LSLAM_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) on device %d: %s", h->device,
                        e != cudaSuccess ? cudaGetErrorString(e) : "null address");
        __atomic_store_n(&g_static_scratch_addr[h->device], p, __ATOMIC_RELEASE);   // every thread computes the same value
    }
    *out = p;
    return LSLAM_OK;
}
} // namespace

extern "C" int lslam_ukf_aug_step_static(void *handle, double *x, double *P, const double *u, const double *z, int32_t F,
                                         int32_t L, const double *params_h, int32_t flags, int32_t *status, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (F < 0 || L < 1 || L > 1024)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ukf_aug_step_static: need F >= 0, 1 <= L <= 1024");
    if (F == 0) return LSLAM_OK;
    int64_t need = 0;
    if (lslam_ukf_aug_workspace_bytes(L, F, &need) != LSLAM_OK || need > (int64_t)LSLAM_STATIC_SCRATCH_BYTES)
        return fail(h, LSLAM_ERR_CAPACITY,
                    "lslam_ukf_aug_step_static: %d filters of L = %d need %lld bytes of scratch, the static array holds %lld "
                    "(LSLAM_STATIC_SCRATCH_BYTES); use lslam_ukf_aug_step with a caller-owned workspace",
                    (int)F, (int)L, (long long)need, (long long)LSLAM_STATIC_SCRATCH_BYTES);
    void *scratch = nullptr;
    {
        DeviceGuard g(h->device);
        const int rc = static_scratch_address(h, &scratch);
        if (rc != LSLAM_OK) return rc;
    }
    // nothing in the workspace has to start at zero: every array is written before it is read (the NaN-filled-workspace
    // tests of lslam_ukf_aug_step), so the previous call's contents are simply overwritten
    return lslam_ukf_aug_step(handle, x, P, u, z, F, L, params_h, flags, status, scratch, (int64_t)LSLAM_STATIC_SCRATCH_BYTES, stream);
}

// ---- input stage, synthetic scans, fp64 probe --------------------------------------------
extern "C" int lslam_polar_to_cart(void *handle, const double *dist, const double *ang_deg, double *xy,
                                   int64_t npts, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (npts < 0 || (npts > 0 && (!dist || !ang_deg || !xy)))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_polar_to_cart: bad arguments");
    if (((uintptr_t)xy & 15) != 0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_polar_to_cart: xy must be 16-byte aligned");
    if (npts == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    const int threads = 256;
    const long long grid = (npts + threads - 1) / threads;
    if (grid > 0x7fffffffLL) return fail(h, LSLAM_ERR_CAPACITY, "too many points for one launch");
    auto kern = polar_to_cart_kernel;
    LSLAM_LAUNCH(kern, (unsigned)grid, threads, 0, (cudaStream_t)stream, dist, ang_deg,
                 reinterpret_cast<double2 *>(xy), (long long)npts);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

extern "C" int lslam_synth_scans(void *handle, double *xy, int32_t B, int32_t beams, uint64_t seed,
                                 uint64_t scan_base, double noise_mm, double dropout, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (B < 0 || beams < 1 || !xy || ((uintptr_t)xy & 15) != 0)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_synth_scans: bad arguments");
    if (B == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    const long long total = (long long)B * beams;
    const int threads = 256;
    const long long grid = (total + threads - 1) / threads;
    if (grid > 0x7fffffffLL) return fail(h, LSLAM_ERR_CAPACITY, "too many points for one launch");
    auto kern = synth_scans_kernel;
    LSLAM_LAUNCH(kern, (unsigned)grid, threads, 0, (cudaStream_t)stream, reinterpret_cast<double2 *>(xy),
                 total, (int)beams, seed, scan_base, noise_mm, dropout);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

// ---- R7 / N4: OLS-flavour RANSAC (line_detection.py:57-58) -------------------------------------
extern "C" int lslam_ransac_ols(void *handle, const double *xy, const int32_t *npts, int32_t B, int32_t N,
                                int32_t window, int32_t trials, double thr, double stop_probability, uint64_t seed,
                                uint64_t scan_base, const int32_t *pairs, double *fit, int32_t *ninl,
                                int32_t *ntrials, int32_t *best_trial, uint32_t *inl_mask, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (B < 0 || N < 1 || window < 2 || trials < 0 || !(thr >= 0.0) || !(stop_probability >= 0.0 && stop_probability <= 1.0))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_ols: bad sizes or parameters");
    if (B > 0 && (!xy || !fit || !ninl || !ntrials || !best_trial || !inl_mask))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_ols: null buffer");
    if (((uintptr_t)xy & 15) != 0) return fail(h, LSLAM_ERR_BAD_ARG, "lslam_ransac_ols: xy must be 16-byte aligned");
    if (window > 1024) return fail(h, LSLAM_ERR_CAPACITY, "lslam_ransac_ols: window > 1024 points");
    if (B == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    OlsArgs a;
    a.xy = reinterpret_cast<const double2 *>(xy); a.npts = npts; a.pairs = pairs;
    a.B = B; a.N = N; a.window = window; a.C = (N + window - 1) / window; a.W = (window + 31) / 32; a.trials = trials;
    a.thr = thr;
    a.nom = std::fmax(DBL_EPSILON, 1.0 - stop_probability);
    a.seed = seed; a.scan_base = scan_base;
    a.fit = fit; a.ninl = ninl; a.ntrials = ntrials; a.best_trial = best_trial; a.inl_mask = inl_mask;
    const long long total = (long long)B * a.C;
    int nwarps = 8;
    size_t smem = (size_t)nwarps * 2 * window * sizeof(double);
    while (nwarps > 1 && smem > (size_t)h->smem_optin) { nwarps >>= 1; smem = (size_t)nwarps * 2 * window * sizeof(double); }
    auto kern = ransac_ols_kernel;
#ifndef LSLAM_EMU
    if (smem > 48 * 1024) LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
#endif
    long long blocks = (total + nwarps - 1) / nwarps;
    const long long cap = (long long)h->sms * 32;
    if (blocks > cap) blocks = cap;
    LSLAM_LAUNCH(kern, (unsigned)blocks, nwarps * 32, smem, (cudaStream_t)stream, a);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

// ---- N2 input stage: express packets -> points; synthetic packets --------------------------
static int express_shape(Handle *h, int64_t nscans, int P, unsigned *grid, int *threads)
{
    *threads = 32 * (P < kExpressMaxWarps ? P : kExpressMaxWarps);
    const long long cap = (long long)h->sms * 5;          // persistent: 5 CTAs per SM by shared memory
    *grid = (unsigned)(nscans < cap ? nscans : cap);
    return 0;
}

extern "C" int lslam_express_to_cart(void *handle, const uint8_t *packets, int64_t nscans, int32_t packets_per_scan,
                                     int32_t beams, double *xy, double *meas, int32_t *status, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (nscans < 0 || packets_per_scan < 1 || beams < 1 || beams > 32 * (int64_t)packets_per_scan ||
        (nscans > 0 && (!packets || !xy)))
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_express_to_cart: bad arguments");
    if (((uintptr_t)xy & 15) != 0 || ((uintptr_t)meas & 15) != 0 || ((uintptr_t)packets & 3) != 0)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_express_to_cart: xy/meas must be 16-byte, packets 4-byte aligned");
    if (nscans == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    const int threads = 32 * kExpressMaxWarps;
    if (status) LSLAM_CUDA_TRY(h, cudaMemsetAsync(status, 0, sizeof(int32_t) * (size_t)nscans, (cudaStream_t)stream));
    auto kern = express_to_cart_kernel;
#ifndef LSLAM_EMU
    LSLAM_CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kExpressDecodeSmem));
#endif
    // persistent warps (a pair of packets per iteration): exactly the CTAs that are resident at a time - a grid larger than
    // that runs its tail as a second, half-empty wave
    int per_sm = 0;
    LSLAM_CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, (size_t)kExpressDecodeSmem));
    if (per_sm < 1) per_sm = 1;
    const long long pairs = ((long long)nscans * packets_per_scan + 1) / 2, want = (pairs + kExpressMaxWarps - 1) / kExpressMaxWarps;
    const long long cap = (long long)per_sm * h->sms;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    LSLAM_LAUNCH(kern, grid, threads, kExpressDecodeSmem, (cudaStream_t)stream, reinterpret_cast<const uint32_t *>(packets),
                 (long long)nscans, (int)packets_per_scan, (int)beams, reinterpret_cast<double2 *>(xy),
                 reinterpret_cast<double2 *>(meas), status, reinterpret_cast<const double2 *>(h->express_tab_d));
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

extern "C" int lslam_synth_express(void *handle, uint8_t *packets, int64_t nscans, int32_t packets_per_scan,
                                   uint64_t seed, uint64_t scan_base, double noise_mm, double dropout, void *stream)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr) return LSLAM_ERR_BAD_ARG;
    if (nscans < 0 || packets_per_scan < 1 || (nscans > 0 && !packets) || ((uintptr_t)packets & 3) != 0)
        return fail(h, LSLAM_ERR_BAD_ARG, "lslam_synth_express: bad arguments");
    if (nscans == 0) return LSLAM_OK;
    DeviceGuard g(h->device);
    unsigned grid; int threads;
    express_shape(h, nscans, packets_per_scan, &grid, &threads);
    auto kern = synth_express_kernel;
    LSLAM_LAUNCH(kern, grid, threads, kExpressSmem, (cudaStream_t)stream, reinterpret_cast<uint32_t *>(packets),
                 (long long)nscans, (int)packets_per_scan, seed, scan_base, noise_mm, dropout);
    h->launches++;
    LSLAM_CUDA_TRY(h, cudaGetLastError());
    return LSLAM_OK;
}

extern "C" int lslam_fp64_peak(void *handle, double ms_target, double *tflops_h)
{
    Handle *h = (Handle *)handle;
    if (h == nullptr || tflops_h == nullptr) return LSLAM_ERR_BAD_ARG;
    DeviceGuard g(h->device);
    const int threads = 256, blocks = h->sms * 8;
    double *sink = nullptr;
    LSLAM_CUDA_TRY(h, cudaMalloc((void **)&sink, sizeof(double) * (size_t)blocks * threads));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    auto kern = dfma_probe_kernel;
    int iters = 1 << 12;
    float ms = 0.f;
    int rc = LSLAM_OK;
    for (int round = 0; round < 8; ++round) {
        LSLAM_LAUNCH(kern, blocks, threads, 0, (cudaStream_t)0, sink, iters);   // warm-up at this size
        cudaEventRecord(e0, 0);
        LSLAM_LAUNCH(kern, blocks, threads, 0, (cudaStream_t)0, sink, iters);
        cudaEventRecord(e1, 0);
        h->launches += 2;
        if (cudaEventSynchronize(e1) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
            rc = fail(h, LSLAM_ERR_CUDA, "fp64 probe failed");
            break;
        }
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms >= ms_target || iters >= (1 << 24)) break;
        iters *= 2;
    }
    if (rc == LSLAM_OK) {
        const double flops = 2.0 * kDfmaChains * kDfmaUnroll * (double)iters * (double)blocks * threads;
        *tflops_h = ms > 0.f ? flops / (ms * 1e-3) * 1e-12 : 0.0;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    return rc;
}
