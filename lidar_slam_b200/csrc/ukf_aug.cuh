// ukf_aug.cuh - augmented-state UKF: pose + L landmarks in the state,
// n = 3 + 2L (403 at L = 200), 2n + 1 sigma points, dim_z = 2L.
//
// The reference's filter is a localiser with n = 3 (ukf.cuh).  BASELINE.json configs[3] asks for the same
// filterpy engine (UnscentedKalmanFilter + MerweScaledSigmaPoints, systemClass.py:20-21) with the landmark
// positions INSIDE the state: fx leaves them fixed, hx reads them from the state vector
// (oracle/ukf.py transfer_function_augmented), x_mean / residual_x treat index 2 as an angle and the rest
// linearly.  At this size one filter step is dense linear algebra, ~1.3 GFLOP of fp64, and the matrices
// (P: 1.3 MB) do not fit on chip, so the step is a sequence of batched kernels over global-memory
// matrices, all filters of a batch side by side:
//   Cholesky        U^T U = (n + lambda) P        blocked left-looking, per 32-row block: aug_chol_diag_kernel (bracket of the
//                                                 diagonal block on TMA + DMMA, factorisation, inverse) + a SOLVE GEMM (block row)
//   sigma points    x, x + U[k], x - U[k] - never stored: every consumer recomputes them from (x, U)
//   fx, mean, residuals   aug_sigma_resid_kernel: fx on the pose part, xbar (circular mean for the heading), Y = residual_x(., xbar)
//   Pbar            = Y^T diag(Wc) Y + Q          GEMM (upper triangle); the epilogue adds Q and also stores (n + lambda) Pbar,
//                                                 the matrix the re-draw factors
//   redraw          Cholesky of (n + lambda) Pbar; Y = residuals of the new points about xbar (+ half only)   (filterpy >= 1.4.5)
//   hx              H[k] = (range, bearing) of every landmark of sigma point k (aug_hx_kernel, from the factor)
//   zhat, E, D      aug_meas_resid_kernel: zhat = z_mean(H), E = residual_h(H, zhat) in place, D = E+ - E-
//   S               = E^T diag(Wc) E + R (upper, straight into the factor buffer);  Pzx = wi D^T Y+ (paired form; ukf_aug.pzx = 0:
//                   the literal E^T diag(Wc) Y with the all-zero k-tiles skipped)
//   gain            through the Cholesky S = U^T U and ONE blocked forward solve W = U^-T [Pzx | r] (SOLVE GEMMs with the
//                   inverted diagonal blocks the factorisation left behind):
//                   K S K^T = Pzx^T S^-1 Pzx = W^T W and K r = W^T (U^-T r), so neither K nor inv(S) is formed
//                   (filterpy forms inv(S) and K explicitly; same x, P to rounding, 129 MFMA less per step)
//   update          x = xbar + W^T v;  P = Pbar - W^T W, both triangles stored from the GEMM epilogue into the caller's P
// One GEMM form serves all of them: C[i,j] = beta C[i,j] + alpha sum_k w_k A[k,i] B[k,j] with A and B
// stored k-major ("TN"), which is how the sigma-point arrays and the factor rows lie in memory: aug_gemm_tma_kernel
// (TMA producer warp + DMMA consumer warps).  The earlier formulations (a DFMA tile kernel, DMMA fed by cp.async, substitution-
// based panel / solve kernels: 2 x slower) were removed from the library in round 2; DESIGN.md 4.6 keeps their numbers.
#pragma once
#include "lslam_core.cuh"
#include "async_copy.cuh"
#include "ransac.cuh"
#include "ukf.cuh"

namespace lslam {

// ---- batched TN GEMM -------------------------------------------------------------------------
// C (M x N, ldc) = beta * C + alpha * sum_{k < K} A[k, i] * B[k, j] + w0 * A[0, i] * B[0, j]
// A: K x M (lda), B: K x N (ldb), both k-major.  The w0 term carries the central sigma point's weight:
// sum_k Wc[k] a_k b_k = wi * sum_k a_k b_k + (Wc[0] - wi) a_0 b_0 (Merwe weights are equal for k >= 1).
// grid = (ceil(N/64), ceil(M/TM), batch); 256 threads as 16 x 16, (TM/16) x 4 outputs per thread, rows
// {2ty, 2ty+1 (+32)} and columns {2tx, 2tx+1, 32+2tx, 33+2tx} so that each operand pair is one 16-byte
// shared-memory load.  Tiles of 16 k-rows go global -> shared with 8-byte cp.async (LDGSTS; the row pitch
// n = 403 is odd, so rows are only 8-byte aligned) through a 3-stage ring: the loads of tiles t+1, t+2
// are in flight while tile t is multiplied.
struct AugGemm {
    const double *A, *B;
    double *C;
    long long sA, sB, sC;           // element strides between consecutive filters
    int M, N, K, lda, ldb, ldc;
    double alpha, beta, w0;
    const int32_t *skip;            // per-filter status: non-zero -> leave C alone (may be null)
    int upper;                      // 1: C is square and only its upper triangle (j >= i) is computed / stored
    // epilogue extras (fused former elementwise passes):
    int noise;                      // 1: + nq[3 i + j] on the 3 x 3 pose block (Q);  2: + nq[i & 1] on the diagonal (R)
    double nq[9];
    double *C2;                     // second output C2[i, j] = scale2 * C[i, j] (the matrix the next Cholesky factors), or null
    long long sC2;
    int ldc2;
    double scale2;
    const double *Dinv;             // SOLVE kernels: the inverted 32 x 32 diagonal block, row-major [k][r] (aug_chol_diag_kernel)
    long long sDinv;
    int mirror;                     // 1: upper-triangle result v goes to C2[i, j] and C2[j, i] (C itself is not written)
    int ktri;                       // < 0: B is upper triangular (K x K), a column tile walks rows [0, j0 + TN) only;
                                    // > 0: B holds residuals of sigma points about their own centre, B = [0; +T; -T] with T
                                    // (ktri x ktri) upper triangular: B[k, j] is an exact 0 for 1 + j < k <= ktri and for
                                    // ktri + 1 + j < k, so a column tile only walks the k-tiles that can be non-zero
};

// one output element: C = beta C + alpha acc (+ the central-weight term) (+ noise), optional scaled copy
__device__ __forceinline__ void aug_gemm_store(const AugGemm &g, double *C, double *C2, const double *B, int i, int j, double a0,
                                               double accv)
{
    double *c = C + (size_t)i * g.ldc + j;
    double v = g.alpha * accv;
    if (g.w0 != 0.0) v = fma(a0, B[j], v);
    v = (g.beta == 0.0 ? 0.0 : g.beta * *c) + v;
    if (g.noise == 1) { if (i < 3 && j < 3) v = __dadd_rn(v, g.nq[i * 3 + j]); }
    else if (g.noise == 2) { if (i == j) v = __dadd_rn(v, g.nq[i & 1]); }
    if (g.mirror) {                                                   // final covariance: both triangles straight into the caller's P
        C2[(size_t)i * g.ldc2 + j] = v;
        C2[(size_t)j * g.ldc2 + i] = v;
        return;
    }
    *c = v;
    if (C2 != nullptr) C2[(size_t)i * g.ldc2 + j] = __dmul_rn(g.scale2, v);
}

constexpr int kAugBK = 16;
constexpr size_t kAugMeanSmem = (8 * 32 * 2 + 32) * sizeof(double);
constexpr size_t kAugUpdateXSmem = 8 * 33 * sizeof(double);

// ---- the same GEMM on the fp64 tensor-core instruction (DMMA) ---------------------------------------------
// mma.sync.m8n8k4.f64: one warp instruction = an 8 x 8 x 4 product, 256 FMA, one operand double per lane.
// tools/probe/dmma_probe.cu measures it at the DFMA rate on B200 (37.1 TFLOP/s against 36.9; the two share
// the fp64 pipe, mixed streams add up) - so it buys no flops, it buys everything AROUND the flops: the
// DFMA kernel above needs one 16-byte shared-memory operand load per 4 FMA-instructions and runs into the
// shared-memory pipe (ncu: l1tex 93 %, fp64 pipe 48 %); here a warp tile of WM x WN takes
// (WM + WN) / 8 8-byte loads per (WM / 8)(WN / 8) DMMAs, 8 x fewer shared-memory wavefronts per flop, and
// one issue slot per 256 FMA.
// Fragment layout (PTX ISA, mma.m8n8k4 .f64): A (8 x 4, row): lane holds A[lane >> 2][lane & 3];
// B (4 x 8, col): lane holds B[lane & 3][lane >> 2]; C/D: lane holds C[lane >> 2][2 (lane & 3) + {0, 1}].
// With the tiles k-major in shared memory (as[k][i], bs[k][j]) both operand loads are
// tile[(k0 + (lane & 3)) * LD + c0 + (lane >> 2)]; LD = T + 4 = 4 (mod 16) makes the 16 addresses of each
// half-warp fall in 16 different 8-byte banks.
#ifndef LSLAM_EMU
__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
#else
inline void dmma_m8n8k4(double &c0, double &c1, double a, double b)
{
    double av[32], bv[32];
    emu::gather_all(a, av);
    emu::gather_all(b, bv);
    const int r = emu::ts.lane >> 2, c = (emu::ts.lane & 3) * 2;
    for (int k = 0; k < 4; ++k) {
        c0 = fma(av[r * 4 + k], bv[c * 4 + k], c0);
        c1 = fma(av[r * 4 + k], bv[(c + 1) * 4 + k], c1);
    }
}
#endif

// ---- the DMMA GEMM fed by the TMA engine (default) ---------------------------------------------------------
// ncu on the cp.async version above (profiles/r01m_aug_gemm_ncu_summary.txt): DMMA is 7.6 % of the issued
// warp-instructions, the 64-bit address arithmetic and predicates of the 8-byte LDGSTS copies are 70 %, and
// the tensor pipe is 58 % busy.  Here one elected lane of a PRODUCER warp issues two tiled TMA copies per
// k-tile (cp.async.bulk.tensor.3d, SASS UTMALDG): boxes of 16 k-rows x (T + 4) columns of the k-major
// operands, through tensor maps the host encodes per call (dims = (columns, K, filters), so rows >= K and
// columns >= M / N arrive as zeros from the copy engine - no predicates, no address arithmetic in the
// kernel).  The 4 extra box columns make the shared-memory row pitch T + 4 = 4 (mod 16) doubles, the
// bank-conflict-free pitch of the fragment loads (see above); they are over-fetch, never used.  The
// consumer warps only wait (mbarrier full[s]), load fragments, issue DMMAs and release the stage
// (mbarrier empty[s], one arrival per live warp): no block-wide barrier inside the k-loop.
// Needs 16-byte aligned operands with even pitches - lslam_ukf_aug_step pads the workspace pitch of the
// n-column arrays to n + 1 (n = 3 + 2L is odd).
//
// SOLVE variant (TM = 32, one row of warps): the block-row step of a blocked triangular solve.  After the
// product the 32 x TN tile V = beta C + alpha acc is multiplied from the left by the transposed inverse of
// the 32 x 32 diagonal block, out = Dinv^T V (Dinv = D^-1 upper triangular, from aug_chol_diag_kernel): the
// forward substitution with D^T as one more K = 32 product on the tensor pipe - each warp parks its V tile
// in shared memory as the k-major B operand, Dinv is the k-major A operand.  Used for the block rows of the
// Cholesky factor and of W = U^-T [Pzx | r].
template <int TM, int TN, int WARPS_N, int S, bool SOLVE> constexpr size_t aug_tma_smem()
{
    return (size_t)S * kAugBK * (TM + 4 + TN + 4) * sizeof(double) + (2 * S + 2) * sizeof(Mbar) + 128 +
           (SOLVE ? (size_t)(32 * 36) * sizeof(double) : 0);          // Dinv; the V tiles reuse the operand ring
}

template <int TM, int TN, int WARPS_M, int WARPS_N, int S, bool SOLVE, int MINB = 1>
__global__ void __launch_bounds__(32 * (WARPS_M * WARPS_N + 1), MINB)
aug_gemm_tma_kernel(const AugGemm g, const LSLAM_GRID_CONSTANT Tmap mapA, const LSLAM_GRID_CONSTANT Tmap mapB)
{
    LSLAM_SMEM_DECL;
    constexpr int NW = WARPS_M * WARPS_N, WM = TM / WARPS_M, WN = TN / WARPS_N, FM = WM / 8, FN = WN / 8;
    constexpr int LDA = TM + 4, LDB = TN + 4, STAGE = kAugBK * (LDA + LDB);
    static_assert(!SOLVE || (TM == 32 && WARPS_M == 1), "SOLVE: one 32-row block");
    static_assert(TM % 16 == 0 && TN % 16 == 0 && WM % 8 == 0 && WN % 8 == 0, "tile shape");
    static_assert((kAugBK * LDA * sizeof(double)) % 128 == 0 && (kAugBK * LDB * sizeof(double)) % 128 == 0, "TMA box alignment");
    const int f = blockIdx.z;
    if (g.skip != nullptr && g.skip[f] != 0) return;
    const int i0 = blockIdx.y * TM, j0 = blockIdx.x * TN;
    if (g.upper && i0 > j0 + TN - 1) return;
    // 128-byte aligned tile ring, then the barriers
    // (the pad is computed on the shared-window address so that the pointer stays a SHARED pointer for the
    // compiler: fragment loads must be LDS, not generic LD)
#ifndef LSLAM_EMU
    double *sm = reinterpret_cast<double *>(smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u));
#else
    double *sm = reinterpret_cast<double *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
#endif
    Mbar *full = reinterpret_cast<Mbar *>(sm + (size_t)S * STAGE), *empty = full + S;
    Mbar *dbar = empty + S, *cbar = empty + S + 1;                    // SOLVE: Dinv has landed; every consumer warp is out of the k-loop
    double *Ds = reinterpret_cast<double *>(empty + S + 2);           // SOLVE: Dinv [32][36]
    static_assert(!SOLVE || (size_t)S * STAGE >= (size_t)NW * 32 * (WN + 4), "SOLVE: the V tiles must fit in the operand ring");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // fragment liveness per consumer warp (see aug_gemm_dmma_kernel); a warp without a live fragment leaves
    auto live_mask = [&](int w) {
        const int m0 = (w / WARPS_N) * WM, n0 = (w % WARPS_N) * WN;
        unsigned live = 0;
        for (int a = 0; a < FM; ++a)
            for (int b = 0; b < FN; ++b) {
                const int fi = i0 + m0 + 8 * a, fj = j0 + n0 + 8 * b;
                if (fi < g.M && fj < g.N && !(g.upper && fi > fj + 7)) live |= 1u << (a * FN + b);
            }
        return live;
    };
    if (threadIdx.x == 0) {
        int nlive = 0;
        for (int w = 0; w < NW; ++w) nlive += live_mask(w) != 0;
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nlive > 0 ? nlive : 1); }
        mbar_init(dbar, 1);
        mbar_init(cbar, nlive > 0 ? nlive : 1);
        mbar_fence_init();
    }
    __syncthreads();
    // k-tiles this CTA walks: all of them, or (ktri) the two runs that can hold non-zeros of B's columns
    // [j0, j0 + TN): rows [0, j0 + TN + 1) and [ktri + 1, ktri + 1 + j0 + TN) - skipped terms are products
    // with exact zeros, the sums do not change by a bit
    const int ntiles_all = (g.K + kAugBK - 1) / kAugBK;
    int run1 = ntiles_all, skip12 = 0;                                // tiles [0, run1) then [run1 + skip12, ...)
    int ntiles = ntiles_all;
    if (g.ktri < 0) {                                                 // B = T itself (K x K upper triangular): rows [0, j0 + TN)
        const int e1 = j0 + TN < g.K ? j0 + TN : g.K;
        ntiles = run1 = (e1 + kAugBK - 1) / kAugBK;
    } else if (g.ktri > 0) {
        const int e1 = j0 + TN + 1 < g.K ? j0 + TN + 1 : g.K;          // end of the first run (rows)
        const int b2 = g.ktri + 1, e2 = b2 + j0 + TN < g.K ? b2 + j0 + TN : g.K;
        const int t1 = (e1 + kAugBK - 1) / kAugBK, tb2 = b2 / kAugBK, te2 = (e2 + kAugBK - 1) / kAugBK;
        if (tb2 > t1) { run1 = t1; skip12 = tb2 - t1; ntiles = t1 + (te2 - tb2); }
        else { ntiles = te2 > t1 ? te2 : t1; run1 = ntiles; }
    }
    if (warp == NW) {                                                 // ---- producer
        if (lane == 0) {
            bool any = false;
            for (int w = 0; w < NW; ++w) any = any || live_mask(w) != 0;
            if (any && SOLVE) {                                       // the inverted diagonal block: 32 rows of 256 B into pitch-36 rows
                const double *Dinv = g.Dinv + (size_t)f * g.sDinv;
                mbar_expect_tx(dbar, 32 * 32 * (unsigned)sizeof(double));
                for (int r = 0; r < 32; ++r) bulk_g2s(Ds + r * 36, Dinv + r * 32, 32 * (unsigned)sizeof(double), dbar);
            }
            if (any)
                for (int t = 0; t < ntiles; ++t) {
                    const int s = t % S;
                    if (t >= S) mbar_wait_relaxed(&empty[s], ((t / S) - 1) & 1);
                    double *as = sm + (size_t)s * STAGE, *bs = as + kAugBK * LDA;
                    mbar_expect_tx(&full[s], (unsigned)(STAGE * sizeof(double)));
                    const int kt = t < run1 ? t : t + skip12;
                    tma_load_3d(as, &mapA, i0, kt * kAugBK, f, &full[s]);
                    tma_load_3d(bs, &mapB, j0, kt * kAugBK, f, &full[s]);
                }
        }
        return;
    }
    const int wm0 = (warp / WARPS_N) * WM, wn0 = (warp % WARPS_N) * WN;
    const unsigned live = live_mask(warp);
    if (live == 0) return;
    const bool fullw = live == (FM * FN == 32 ? 0xffffffffu : (1u << (FM * FN)) - 1u);
    if (SOLVE) {                                                      // the epilogue reads this warp's C tile: start it towards L2 now
        const double *crow = g.C + (size_t)f * g.sC + (size_t)lane * g.ldc + j0 + wn0;
        if (lane < g.M && j0 + wn0 < g.N) { prefetch_l2(crow); prefetch_l2(crow + WN - 1 < crow + (g.N - j0 - wn0) ? crow + WN - 1 : crow); }
    }
    double acc[FM][FN][2];
#pragma unroll
    for (int a = 0; a < FM; ++a)
#pragma unroll
        for (int b = 0; b < FN; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
    const int frag = (lane & 3), fcol = (lane >> 2);
    for (int t = 0; t < ntiles; ++t) {
        const int s = t % S;
        mbar_wait(&full[s], (t / S) & 1);
        const double *as = sm + (size_t)s * STAGE + frag * LDA + wm0 + fcol;
        const double *bs = sm + (size_t)s * STAGE + kAugBK * LDA + frag * LDB + wn0 + fcol;
        // fragments double-buffered in registers: the loads of k-step q + 1 are in flight under the DMMAs of q
        double av[2][FM], bv[2][FN];
#pragma unroll
        for (int a = 0; a < FM; ++a) av[0][a] = as[8 * a];
#pragma unroll
        for (int b = 0; b < FN; ++b) bv[0][b] = bs[8 * b];
#pragma unroll
        for (int q = 0; q < kAugBK / 4; ++q) {
            if (q + 1 < kAugBK / 4) {
#pragma unroll
                for (int a = 0; a < FM; ++a) av[(q + 1) & 1][a] = as[4 * (q + 1) * LDA + 8 * a];
#pragma unroll
                for (int b = 0; b < FN; ++b) bv[(q + 1) & 1][b] = bs[4 * (q + 1) * LDB + 8 * b];
            }
            if (fullw) {
#pragma unroll
                for (int a = 0; a < FM; ++a)
#pragma unroll
                    for (int b = 0; b < FN; ++b) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], av[q & 1][a], bv[q & 1][b]);
            } else {
#pragma unroll
                for (int a = 0; a < FM; ++a)
#pragma unroll
                    for (int b = 0; b < FN; ++b)
                        if ((live >> (a * FN + b)) & 1u) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], av[q & 1][a], bv[q & 1][b]);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    const double *A = g.A + (size_t)f * g.sA, *B = g.B + (size_t)f * g.sB;
    double *C = g.C + (size_t)f * g.sC, *C2 = g.C2 ? g.C2 + (size_t)f * g.sC2 : nullptr;
    if (SOLVE) {
        // the V tiles go where the operand ring was (4 CTAs per SM instead of 3): wait until every live consumer warp
        // has left the k-loop - the producer issued its last copy long before, nothing writes the ring any more
        constexpr int LDV = WN + 4;
        __syncwarp();
        if (lane == 0) mbar_arrive(cbar);
        mbar_wait(cbar, 0);
        double *Vs = sm + (size_t)warp * 32 * LDV;
#pragma unroll
        for (int a = 0; a < FM; ++a) {
            const int i = 8 * a + fcol;
#pragma unroll
            for (int b = 0; b < FN; ++b)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int jl = 8 * b + 2 * frag + q, j = j0 + wn0 + jl;
                    double v = 0.0;
                    if (i < g.M && j < g.N) v = fma(g.alpha, acc[a][b][q], g.beta == 0.0 ? 0.0 : g.beta * C[(size_t)i * g.ldc + j]);
                    Vs[i * LDV + jl] = v;
                    acc[a][b][0 + q] = 0.0;
                }
        }
        __syncwarp();
        mbar_wait(dbar, 0);
#pragma unroll
        for (int ks = 0; ks < 32; ks += 4) {
            double av[FM], bv[FN];
#pragma unroll
            for (int a = 0; a < FM; ++a) av[a] = Ds[(ks + frag) * 36 + 8 * a + fcol];
#pragma unroll
            for (int b = 0; b < FN; ++b) bv[b] = Vs[(ks + frag) * LDV + 8 * b + fcol];
#pragma unroll
            for (int a = 0; a < FM; ++a)
#pragma unroll
                for (int b = 0; b < FN; ++b) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], av[a], bv[b]);
        }
#pragma unroll
        for (int a = 0; a < FM; ++a) {
            const int i = 8 * a + fcol;
            if (i >= g.M) continue;
#pragma unroll
            for (int b = 0; b < FN; ++b)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int j = j0 + wn0 + 8 * b + 2 * frag + q;
                    if (j < g.N) C[(size_t)i * g.ldc + j] = acc[a][b][q];
                }
        }
    } else {
#pragma unroll
        for (int a = 0; a < FM; ++a) {
            const int i = i0 + wm0 + 8 * a + fcol;
            if (i >= g.M) continue;
            const double a0 = g.w0 != 0.0 ? g.w0 * A[i] : 0.0;
#pragma unroll
            for (int b = 0; b < FN; ++b)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int j = j0 + wn0 + 8 * b + 2 * frag + q;
                    if (j >= g.N || (g.upper && j < i)) continue;
                    aug_gemm_store(g, C, C2, B, i, j, a0, acc[a][b][q]);
                }
        }
    }
}

// ---- diagonal block of the left-looking blocked Cholesky -------------------------------------------------------
// One CTA per matrix, block step jb (nb <= 32 rows), 4 DMMA warps + 1 TMA producer warp:
//   D = A[jb:jb+nb, jb:jb+nb] - U[0:jb, jb:jb+nb]^T U[0:jb, jb:jb+nb]     the GEMM pipeline of aug_gemm_tma_kernel with ONE
//                                                                          operand (A = B = the 32 columns of the block):
//                                                                          2 x 2 warps of 16 x 16, fragments below the
//                                                                          diagonal skipped
//   D = Ud^T Ud                                                            right-looking in shared memory, all warps
//   Dinv = Ud^-1                                                           blocked: the four 8 x 8 diagonal blocks by back
//                                                                          substitution (a warp each), then the block
//                                                                          super-diagonals X_ij = -X_ii sum_k U_ik X_kj
// Ud goes back into A, Dinv (row-major, zero-padded to 32 x 32) to the workspace: the block row
// U[jb:jb+nb, jb+nb:] = Ud^-T (A - U^T U) = Dinv^T (...) is then the SOLVE epilogue of the GEMM that forms
// the bracket - the triangular solve of every block step runs on the tensor pipe instead of as a serial
// substitution per column.  A non-positive pivot sets bit `fail_bit` of status[f] (scipy raises
// LinAlgError) and the matrix is left alone from then on.
// (History, B200, 256 matrices of 403: DFMA accumulation from a shared-memory tile + one-warp inverse 68 us
// per launch - profiles/r01r; a fully unrolled in-register factorisation was no faster: ~80 KB of
// straight-line code executed once per CTA runs at the speed of instruction-cache misses.)
constexpr int kAugDiagStages = 2;      // 29 KB of shared memory: 7 CTAs per SM, so 1,024 matrices are one wave (4 stages: 5 per SM, two waves, 2 x the time)
constexpr size_t kAugDiagSmem = 128 + (size_t)kAugDiagStages * kAugBK * 36 * sizeof(double) + 2 * kAugDiagStages * sizeof(Mbar) +
                                (2 * 32 * 33 + 4 * 8 * 9 + 64) * sizeof(double) + 16;
__global__ void __launch_bounds__(160) aug_chol_diag_kernel(double *Abase, long long sA, int lda, int jb, int nb, double *Dinv_base,
                                                            long long sD, int32_t *status, int fail_bit,
                                                            const LSLAM_GRID_CONSTANT Tmap mapU)
{
    LSLAM_SMEM_DECL;
    constexpr int S = kAugDiagStages, LDT = 36, STAGE = kAugBK * LDT, NW = 4;
    const int f = blockIdx.x;
    if (status[f] != 0) return;
#ifndef LSLAM_EMU
    double *sm = reinterpret_cast<double *>(smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u));
#else
    double *sm = reinterpret_cast<double *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
#endif
    Mbar *full = reinterpret_cast<Mbar *>(sm + (size_t)S * STAGE), *empty = full + S;
    double (*Ub)[33] = reinterpret_cast<double (*)[33]>(empty + S);
    double (*Xs)[33] = Ub + 32;
    double (*Ts)[8][9] = reinterpret_cast<double (*)[8][9]>(Xs + 32);
    double *rinv = reinterpret_cast<double *>(Ts + 4), *dsave = rinv + 32;
    int &bad = *reinterpret_cast<int *>(dsave + 32);
    double *A = Abase + (size_t)f * sA;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nwarp = nthr >> 5;
    // ---- phase 1: D = A_jj - U^T U into Ub
    auto live_mask = [&](int w) {
        const int m0 = (w >> 1) * 16, n0 = (w & 1) * 16;
        unsigned live = 0;
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) {
                const int fi = m0 + 8 * a, fj = n0 + 8 * b;
                if (fi < nb && fj < nb && fi <= fj + 7) live |= 1u << (a * 2 + b);
            }
        return live;
    };
    for (int e = tid; e < 32 * 33; e += nthr) { (&Ub[0][0])[e] = 0.0; (&Xs[0][0])[e] = 0.0; }
    if (tid < 32) rinv[tid] = 0.0;
    if (tid == 0) {
        int nlive = 0;
        for (int w = 0; w < NW; ++w) nlive += live_mask(w) != 0;
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nlive); }
        mbar_fence_init();
        bad = 0;
    }
    __syncthreads();
    const int ntiles = (jb + kAugBK - 1) / kAugBK;
    if (warp == NW) {
        if (lane == 0)
            for (int t = 0; t < ntiles; ++t) {
                const int s = t % S;
                if (t >= S) mbar_wait_relaxed(&empty[s], ((t / S) - 1) & 1);
                mbar_expect_tx(&full[s], (unsigned)(STAGE * sizeof(double)));
                tma_load_3d(sm + (size_t)s * STAGE, &mapU, 0, t * kAugBK, f, &full[s]);
            }
    } else if (warp < NW) {
        const unsigned live = live_mask(warp);
        if (live != 0) {
            const int wm0 = (warp >> 1) * 16, wn0 = (warp & 1) * 16, frag = lane & 3, fcol = lane >> 2;
            double acc[2][2][2] = {};
            for (int t = 0; t < ntiles; ++t) {
                const int s = t % S;
                mbar_wait(&full[s], (t / S) & 1);
                const double *ts = sm + (size_t)s * STAGE + frag * LDT + fcol;
#pragma unroll
                for (int q = 0; q < kAugBK / 4; ++q) {
                    double av[2], bv[2];
#pragma unroll
                    for (int a = 0; a < 2; ++a) av[a] = ts[4 * q * LDT + wm0 + 8 * a];
#pragma unroll
                    for (int b = 0; b < 2; ++b) bv[b] = ts[4 * q * LDT + wn0 + 8 * b];
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int b = 0; b < 2; ++b)
                            if ((live >> (a * 2 + b)) & 1u) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], av[a], bv[b]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[s]);
            }
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int i = wm0 + 8 * a + fcol, j = wn0 + 8 * b + 2 * frag + q;
                        if (((live >> (a * 2 + b)) & 1u) && i < nb && j < nb && j >= i)
                            Ub[i][j] = __dsub_rn(A[(size_t)(jb + i) * lda + jb + j], acc[a][b][q]);
                    }
        }
    }
    __syncthreads();
    // ---- phase 2: right-looking factorisation in shared memory, all warps: row r is scaled by warp 0, the trailing
    // rows are updated by the warps round-robin, lanes over columns.  The diagonal keeps the pivots until the end
    // (dsave holds the roots), so reading the next pivot needs no extra barrier.
    int ok = 1;
    for (int r = 0; r < nb; ++r) {
        if (warp == 0) {                                              // root and reciprocal once, not by every warp: with several
            const double piv = Ub[r][r];                              // CTAs per SM the redundant fp64 sqrt / divide chains of all
            if (!(piv > 0.0)) { if (lane == 0) bad = 1; }             // warps were a third of the kernel
            else {
                const double d = sqrt(piv), ri = 1.0 / d;
                if (lane > r) Ub[r][lane] = __dmul_rn(Ub[r][lane], ri);
                if (lane == 0) { rinv[r] = ri; dsave[r] = d; }
            }
        }
        __syncthreads();
        if (bad) { ok = 0; break; }                                   // uniform: read after the barrier
        for (int ii = r + 1 + warp; ii < nb; ii += nwarp)
            if (lane >= ii) Ub[ii][lane] = fma(-Ub[r][ii], Ub[r][lane], Ub[ii][lane]);
        __syncthreads();
    }
    if (!ok) {
        if (tid == 0) status[f] |= fail_bit;
        return;
    }
    if (tid < nb) Ub[tid][tid] = dsave[tid];
    // ---- phase 3: Dinv = Ud^-1, 8 x 8 blocks.  Padded rows / columns (>= nb) have rinv = 0 and zero entries: they stay 0.
    if (warp < 4 && lane < 8) {                                       // diagonal block `warp`: column lane by back substitution
        const int b0 = 8 * warp, c = lane;
        double x[8];
#pragma unroll
        for (int rr = 7; rr >= 0; --rr) {
            double sum = 0.0;
#pragma unroll
            for (int kk = rr + 1; kk < 8; ++kk) sum = fma(Ub[b0 + rr][b0 + kk], kk <= c ? x[kk] : 0.0, sum);
            x[rr] = rr > c ? 0.0 : (rr == c ? rinv[b0 + rr] : -sum * rinv[b0 + rr]);
        }
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) Xs[b0 + rr][b0 + c] = x[rr];
    }
    __syncthreads();
    for (int dd = 1; dd < 4; ++dd) {                                  // block super-diagonal dd: X_ij = -X_ii (sum_{i < k <= j} U_ik X_kj), j = i + dd
        if (warp < 4 - dd) {
            const int bi = warp, bj = warp + dd, c = lane & 7, r0 = lane >> 3;   // this lane: rows r0 and r0 + 4, column c
            double t0 = 0.0, t1 = 0.0;
            for (int k = 8 * (bi + 1); k < 8 * (bj + 1); ++k) {
                const double xk = Xs[k][8 * bj + c];
                t0 = fma(Ub[8 * bi + r0][k], xk, t0);
                t1 = fma(Ub[8 * bi + r0 + 4][k], xk, t1);
            }
            Ts[warp][r0][c] = t0; Ts[warp][r0 + 4][c] = t1;
            __syncwarp();
            double y0 = 0.0, y1 = 0.0;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const double tk = Ts[warp][kk][c];
                y0 = fma(Xs[8 * bi + r0][8 * bi + kk], tk, y0);
                y1 = fma(Xs[8 * bi + r0 + 4][8 * bi + kk], tk, y1);
            }
            Xs[8 * bi + r0][8 * bj + c] = -y0; Xs[8 * bi + r0 + 4][8 * bj + c] = -y1;
        }
        __syncthreads();
    }
    double *Dinv = Dinv_base + (size_t)f * sD;
    for (int e = tid; e < 32 * 32; e += nthr) Dinv[e] = Xs[e >> 5][e & 31];
    for (int e = tid; e < nb * nb; e += nthr) {
        const int r = e / nb, c = e % nb;
        if (c >= r) A[(size_t)(jb + r) * lda + jb + c] = Ub[r][c];
    }
}


// ---- elementwise stages -----------------------------------------------------------------------
struct AugArgs {
    int F, L, n, m, dz;             // filters in this batch, landmarks, 3 + 2L, 2n + 1, 2L
    int ldn;                        // row pitch of the workspace arrays with n columns (U, X, Y, Pbar): n + 1, even
    double *x, *P;                  // [F, n], [F, n, n] in / out
    const double *u, *z;            // [F, 2], [F, L, 2]
    int32_t *status;                // [F]
    double *U, *X, *Y, *H, *Kt, *Pbar, *xbar, *zhat, *wm, *wc;   // workspace (Kt: W = U^-T [Pzx | r], dz x (n + 1))
    double dt, half_r, r_over_l, cscale, var_dist, var_angle;
    double Q[9];
};

// Merwe weights (MerweScaledSigmaPoints._compute_weights): wi everywhere, wm0 / wc0 on the central point
__global__ void aug_weights_kernel(double *wm, double *wc, int m, double wi, double wm0, double wc0)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < m) { wm[k] = k ? wi : wm0; wc[k] = k ? wi : wc0; }
}

// dst = scale * src on the upper triangle (lower zeroed), n x n with row pitches lds / ldd
__global__ void aug_scale_upper_kernel(const double *src, long long s_src, int lds, double *dst, long long s_dst, int ldd, int n,
                                       double scale, const int32_t *status)
{
    const int f = blockIdx.y;
    if (status[f] != 0) return;
    src += (size_t)f * s_src; dst += (size_t)f * s_dst;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)n * n; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / n), c = (int)(e % n);
        dst[(size_t)r * ldd + c] = c >= r ? __dmul_rn(scale, src[(size_t)r * lds + c]) : 0.0;
    }
}

// ---- fused sigma points -> fx -> mean -> residuals (state side) ---------------------------------------------
// The sigma points of (x0, U) are never stored: element (k, c) is x0[c], x0[c] + U[k-1, c] or x0[c] - U[k-n-1, c]
// (MerweScaledSigmaPoints.sigma_points; U upper triangular - entries left of the diagonal are taken as 0
// whatever the memory holds), so every pass recomputes it from the factor (n x n) instead of reading a
// stored (2n+1) x n array.
__device__ __forceinline__ double aug_sigma_value(const double *x, const double *U, int ld, int n, int k, int c)
{
    double v = x[c];
    if (k >= 1) {
        const int r = k <= n ? k - 1 : k - n - 1;
        const double uv = c >= r ? U[(size_t)r * ld + c] : 0.0;
        v = k <= n ? __dadd_rn(v, uv) : __dsub_rn(v, uv);
    }
    return v;
}

// pose component c (< 3) of fx(sigma point k): the differential-drive transition (UKFMethods.py:17-24),
// x' = x + dt * B(theta) u,  B = [[R/2 cos, R/2 cos], [R/2 sin, R/2 sin], [-R/L, R/L]]  (oracle/ukf.py order)
__device__ __forceinline__ double aug_fx_pose(const AugArgs &a, const double *x, const double *U, int k, int c, double ul, double ur)
{
    const double p0 = aug_sigma_value(x, U, a.ldn, a.n, k, 0), p1 = aug_sigma_value(x, U, a.ldn, a.n, k, 1),
                 p2 = aug_sigma_value(x, U, a.ldn, a.n, k, 2);
    double sn, cs;
    sincos(p2, &sn, &cs);
    const double b0 = __dmul_rn(a.half_r, cs), b1 = __dmul_rn(a.half_r, sn);
    const double d0 = __dadd_rn(__dmul_rn(b0, ul), __dmul_rn(b0, ur));
    const double d1 = __dadd_rn(__dmul_rn(b1, ul), __dmul_rn(b1, ur));
    const double d2 = __dadd_rn(__dmul_rn(-a.r_over_l, ul), __dmul_rn(a.r_over_l, ur));
    return c == 0 ? __dadd_rn(p0, __dmul_rn(a.dt, d0)) : c == 1 ? __dadd_rn(p1, __dmul_rn(a.dt, d1)) : __dadd_rn(p2, __dmul_rn(a.dt, d2));
}

// One CTA per (32-column group, filter), 8 warps over the sigma points, lanes across 32 consecutive columns
// (coalesced rows of U and Y).  Pass 1 (skipped when the mean is given - residuals of re-drawn points about
// the prior): weighted column sums, circular for the heading (column 2), partial sums of the 8 warps added
// in warp order; pass 2: Y[k] = residual_x(sigma_k, mean), and the points themselves into Xout if the
// caller still needs them (update without re-draw).
__global__ void __launch_bounds__(256) aug_sigma_resid_kernel(const AugArgs a, const double *x0, int sx, int do_fx, int mean_given,
                                                              double *Xout, int rows)
{
    LSLAM_SMEM_DECL;
    double *part = reinterpret_cast<double *>(smem_raw);             // [8][32][2], then mu at [512 + lane]
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n = a.n, ld = a.ldn;
    const int c = blockIdx.x * 32 + lane;
    const bool live = c < n, ang = c == 2, posefx = do_fx && c < 3;
    const double *x = x0 + (size_t)f * sx, *U = a.U + (size_t)f * n * ld;
    double *mean = a.xbar + (size_t)f * ld, *Y = a.Y + (size_t)f * a.m * ld;
    const double ul = a.u ? a.u[2 * f] : 0.0, ur = a.u ? a.u[2 * f + 1] : 0.0;
    if (!mean_given) {
        double s0 = 0.0, s1 = 0.0;
        if (live) {
            for (int k = warp; k < a.m; k += 8) {
                const double v = posefx ? aug_fx_pose(a, x, U, k, c, ul, ur) : aug_sigma_value(x, U, ld, n, k, c), w = a.wm[k];
                if (ang) { double sn, cs; sincos(v, &sn, &cs); s0 = fma(sn, w, s0); s1 = fma(cs, w, s1); }
                else s0 = fma(v, w, s0);
            }
        }
        part[(warp * 32 + lane) * 2] = s0; part[(warp * 32 + lane) * 2 + 1] = s1;
        __syncthreads();
        if (warp == 0) {
            double t0 = 0.0, t1 = 0.0;
            for (int w = 0; w < 8; ++w) { t0 = __dadd_rn(t0, part[(w * 32 + lane) * 2]); t1 = __dadd_rn(t1, part[(w * 32 + lane) * 2 + 1]); }
            const double mu = ang ? atan2(t0, t1) : t0;
            part[512 + lane] = mu;
            if (live) mean[c] = mu;
        }
        __syncthreads();
    } else {
        if (warp == 0) part[512 + lane] = live ? mean[c] : 0.0;
        __syncthreads();
    }
    if (!live) return;
    const double mu = part[512 + lane];
    double *X = Xout ? Xout + (size_t)f * a.m * ld : nullptr;
    for (int k = warp; k < rows; k += 8) {                            // rows = 2n + 1, or n + 1 when only the + points are needed
        const double v = posefx ? aug_fx_pose(a, x, U, k, c, ul, ur) : aug_sigma_value(x, U, ld, n, k, c);
        double d = __dsub_rn(v, mu);
        if (ang) d = ukf_normalize_angle(d);
        Y[(size_t)k * ld + c] = d;
        if (X) X[(size_t)k * ld + c] = v;
    }
}

// Measurement side: zhat = z_mean(H, Wm) (UKFMethods.py:47-57: weighted mean of the ranges, circular mean of the
// bearings), then E = residual_h(H, zhat) in place, and optionally D[r] = E[1 + r] - E[1 + n + r] (the operand
// of the paired cross-covariance product, lslam_ukf_aug_step).  One CTA per (32 landmarks, filter); lane l owns
// the (range, bearing) PAIR of one landmark - one 16-byte load per sigma point, and every lane of a warp runs
// the bearing's sincos (with a column per lane half of the lanes idle through it: the pass was
// instruction-bound on exactly that).  8 warps over the sigma points, partial sums added in warp order - the
// same order of additions per column as a column-per-lane layout.
constexpr size_t kAugMeasSmem = (8 * 32 * 3 + 64) * sizeof(double);
__global__ void __launch_bounds__(256) aug_meas_resid_kernel(const AugArgs a, double *Dbase)
{
    LSLAM_SMEM_DECL;
    double *part = reinterpret_cast<double *>(smem_raw);             // [8][32][3], then (mu_r, mu_b) at [768 + 2 lane]
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, dz = a.dz, n = a.n;
    const int l = blockIdx.x * 32 + lane;
    const bool live = l < a.L;
    double *H = a.H + (size_t)f * a.m * dz + 2 * l;                   // this landmark's column pair, row pitch dz
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    if (live) {
        for (int k = warp; k < a.m; k += 8) {
            const double2 v = *reinterpret_cast<const double2 *>(H + (size_t)k * dz);
            const double w = a.wm[k];
            double sn, cs;
            sincos(v.y, &sn, &cs);
            s0 = fma(v.x, w, s0); s1 = fma(sn, w, s1); s2 = fma(cs, w, s2);
        }
    }
    part[(warp * 32 + lane) * 3] = s0; part[(warp * 32 + lane) * 3 + 1] = s1; part[(warp * 32 + lane) * 3 + 2] = s2;
    __syncthreads();
    if (warp == 0) {
        double t0 = 0.0, t1 = 0.0, t2 = 0.0;
        for (int w = 0; w < 8; ++w) {
            t0 = __dadd_rn(t0, part[(w * 32 + lane) * 3]);
            t1 = __dadd_rn(t1, part[(w * 32 + lane) * 3 + 1]);
            t2 = __dadd_rn(t2, part[(w * 32 + lane) * 3 + 2]);
        }
        const double mb = atan2(t1, t2);
        part[768 + 2 * lane] = t0; part[768 + 2 * lane + 1] = mb;
        if (live) { a.zhat[(size_t)f * dz + 2 * l] = t0; a.zhat[(size_t)f * dz + 2 * l + 1] = mb; }
    }
    __syncthreads();
    if (!live) return;
    const double mr = part[768 + 2 * lane], mb = part[768 + 2 * lane + 1];
    auto resid = [&](int k) {
        double2 v = *reinterpret_cast<const double2 *>(H + (size_t)k * dz);
        v.x = __dsub_rn(v.x, mr);
        v.y = ukf_normalize_angle(__dsub_rn(v.y, mb));
        *reinterpret_cast<double2 *>(H + (size_t)k * dz) = v;
        return v;
    };
    if (Dbase != nullptr) {
        double *D = Dbase + (size_t)f * n * dz + 2 * l;
        if (warp == 0) resid(0);
        for (int r = warp; r < n; r += 8) {
            const double2 e1 = resid(1 + r), e2 = resid(1 + n + r);
            double2 d;
            d.x = __dsub_rn(e1.x, e2.x); d.y = __dsub_rn(e1.y, e2.y);
            *reinterpret_cast<double2 *>(D + (size_t)r * dz) = d;
        }
        return;
    }
    for (int k = warp; k < a.m; k += 8) resid(k);
}

// hx of every sigma point: (range, bearing) of every landmark read from the state (oracle/ukf.py
// transfer_function_augmented, UKFMethods.py:26-34).  from_U: the sigma points are those of (x0, U) and are
// recomputed from the factor (one thread does the + and the - point of a factor row: the row is read once);
// otherwise they are read from the stored array X.
__device__ __forceinline__ void aug_hx_one(double s0, double s1, double s2, double lx, double ly, double *h)
{
    const double dx = __dsub_rn(lx, s0), dy = __dsub_rn(ly, s1);
    h[0] = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    h[1] = ukf_normalize_angle(__dsub_rn(atan2(dy, dx), s2));
}
__global__ void aug_hx_kernel(const AugArgs a, const double *x0, int sx, int from_U)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int n = a.n, ld = a.ldn;
    double *H = a.H + (size_t)f * a.m * a.dz;
    if (!from_U) {
        for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)a.m * a.L; e += (long long)gridDim.x * blockDim.x) {
            const int k = (int)(e / a.L), l = (int)(e % a.L);
            const double *s = a.X + ((size_t)f * a.m + k) * ld;
            aug_hx_one(s[0], s[1], s[2], s[3 + 2 * l], s[4 + 2 * l], H + (size_t)k * a.dz + 2 * l);
        }
        return;
    }
    const double *x = x0 + (size_t)f * sx, *U = a.U + (size_t)f * n * ld;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < (long long)(n + 1) * a.L; e += (long long)gridDim.x * blockDim.x) {
        const int kk = (int)(e / a.L), l = (int)(e % a.L), cx = 3 + 2 * l;
        if (kk == 0) {
            aug_hx_one(x[0], x[1], x[2], x[cx], x[cx + 1], H + 2 * l);
            continue;
        }
        const int r = kk - 1;
        const double *Ur = U + (size_t)r * ld;
        const double u0 = 0 >= r ? Ur[0] : 0.0, u1 = 1 >= r ? Ur[1] : 0.0, u2 = 2 >= r ? Ur[2] : 0.0;
        const double ux = cx >= r ? Ur[cx] : 0.0, uy = cx + 1 >= r ? Ur[cx + 1] : 0.0;
        aug_hx_one(__dadd_rn(x[0], u0), __dadd_rn(x[1], u1), __dadd_rn(x[2], u2), __dadd_rn(x[cx], ux), __dadd_rn(x[cx + 1], uy),
                   H + (size_t)kk * a.dz + 2 * l);
        aug_hx_one(__dsub_rn(x[0], u0), __dsub_rn(x[1], u1), __dsub_rn(x[2], u2), __dsub_rn(x[cx], ux), __dsub_rn(x[cx + 1], uy),
                   H + (size_t)(n + kk) * a.dz + 2 * l);
    }
}

// innovation r = residual_h(z, zhat) into column n of W (W: dz x (n + 1), the right-hand sides of the
// forward solve: Pzx in columns 0..n-1, r in column n)
__global__ void __launch_bounds__(256) aug_innovation_kernel(const AugArgs a)
{
    const int f = blockIdx.x;
    if (a.status[f] != 0) return;
    const double *zh = a.zhat + (size_t)f * a.dz, *z = a.z + (size_t)f * a.dz;
    double *W = a.Kt + (size_t)f * a.dz * (a.n + 1);
    for (int j = threadIdx.x; j < a.dz; j += blockDim.x) {
        double d = __dsub_rn(z[j], zh[j]);
        if (j & 1) d = ukf_normalize_angle(d);
        W[(size_t)j * (a.n + 1) + a.n] = d;
    }
}

// x = xbar + K r with K r = Pxz S^-1 r = W^T v, W = U^-T Pzx (columns 0..n-1), v = U^-T r (column n);
// grid (ceil(n / 32), F): warp w sums rows w, w+8, ... of its 32 columns, partial sums added in warp order
__global__ void __launch_bounds__(256) aug_update_x_kernel(const AugArgs a, double *xnew_base)
{
    LSLAM_SMEM_DECL;
    double (*part)[33] = reinterpret_cast<double (*)[33]>(smem_raw);
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n1 = a.n + 1;
    const int i = blockIdx.x * 32 + lane;
    const double *W = a.Kt + (size_t)f * a.dz * n1;
    double s = 0.0;
    if (i < a.n)
        for (int j = warp; j < a.dz; j += 8) s = fma(W[(size_t)j * n1 + i], W[(size_t)j * n1 + a.n], s);
    part[warp][lane] = s;
    __syncthreads();
    if (warp == 0 && i < a.n) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t = __dadd_rn(t, part[w][lane]);
        const double xv = __dadd_rn(a.xbar[(size_t)f * n1 + i], t);
        if (xnew_base) xnew_base[(size_t)f * n1 + i] = xv;
        else a.x[(size_t)f * a.n + i] = xv;                           // committed directly (status is final by now)
    }
}

// commit: filters whose status is still 0 take (xnew, Pnew); the others keep their state
__global__ void aug_commit_kernel(const AugArgs a, const double *xnew, const double *Pnew)
{
    const int f = blockIdx.y;
    if (a.status[f] != 0) return;
    const long long nn = (long long)a.n * a.n, nl = (long long)a.n * a.ldn;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nn; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / a.n), c = (int)(e % a.n);             // Pnew (pitch ldn) holds the upper triangle
        a.P[(size_t)f * nn + e] = Pnew[(size_t)f * nl + (c >= r ? (long long)r * a.ldn + c : (long long)c * a.ldn + r)];
        if (e < a.n) a.x[(size_t)f * a.n + e] = xnew[(size_t)f * a.ldn + e];
    }
}

// A filter whose re-draw (prior not positive definite) or S factorisation failed after a predict keeps the PRIOR, as
// filterpy does (predict() has stored x, P before the sigma points are re-drawn) and as the n = 3 kernel does: the
// later kernels skipped it, so xbar and the upper triangle of Pbar still hold what the predict left.
__global__ void aug_commit_prior_kernel(const AugArgs a)
{
    const int f = blockIdx.y;
    const int st = a.status[f];
    if ((st & (kUkfStRedraw | kUkfStS)) == 0 || (st & kUkfStSigma) != 0) return;
    const long long nn = (long long)a.n * a.n, nl = (long long)a.n * a.ldn;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nn; e += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(e / a.n), c = (int)(e % a.n);
        a.P[(size_t)f * nn + e] = a.Pbar[(size_t)f * nl + (c >= r ? (long long)r * a.ldn + c : (long long)c * a.ldn + r)];
        if (e < a.n) a.x[(size_t)f * a.n + e] = a.xbar[(size_t)f * a.ldn + e];
    }
}

} // namespace lslam
