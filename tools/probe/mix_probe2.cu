// Micro-benchmark 2: which integer ops overlap with the fp64 pipe?  Variants of the per
// point-hypothesis step: V0 = 4 DP + SHF + FMNMX3/2 (current kernel); V1 = 3 DP + IMAD + IMAD.HI
// + VIMNMX3/2 (integer threshold test on the high word of s); V2 = V1 without the min (count
// only); V3 = 3 DP only + 1 IMAD.HI count (lower bound)
#include <cstdio>
#include <cuda_runtime.h>

template <int V, int CH>
__global__ void __launch_bounds__(256) mix(double *sink, int iters, double thr, int c1)
{
    __shared__ double2 pts[256];
    pts[threadIdx.x] = make_double2(1.0 + threadIdx.x * 1e-3, 2.0 - threadIdx.x * 1e-3);
    __syncthreads();
    double nx[CH], ny[CH], ng[CH], ss[CH];
    unsigned bits[CH], cnt[CH], mn[CH];
    float mab[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) { nx[c] = 0.6 + 1e-3 * c; ny[c] = 0.8 - 1e-3 * c; ng[c] = -1.0 - c + threadIdx.x * 1e-6; ss[c] = 0; bits[c] = 0; mab[c] = 1e30f; cnt[c] = 0; mn[c] = 0xffffffffu; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            const double2 p = pts[(it + k) & 255];
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const double s = fma(p.y, ny[c], fma(p.x, nx[c], ng[c]));
                ss[c] = fma(s, s, ss[c]);
                if (V == 0) {
                    const int hi = __double2hiint(fabs(s) - thr);
                    bits[c] = __funnelshift_l((unsigned)hi, bits[c], 1);
                    mab[c] = fminf(mab[c], fabsf(__int_as_float(hi)));
                } else {
                    const unsigned hi = (unsigned)__double2hiint(s);
                    unsigned kk2;
                    asm("mad.lo.u32 %0, %1, 2, %2;" : "=r"(kk2) : "r"(hi), "r"((unsigned)c1));
                    asm("mad.hi.u32 %0, %1, 2, %0;" : "+r"(cnt[c]) : "r"(kk2));
                    if (V == 1) mn[c] = min(mn[c], kk2);
                    if (V == 3) { }
                }
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += ss[c] + bits[c] + mab[c] + cnt[c] + mn[c];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int V, int CH>
void run(const char *name, int blocks_per_sm)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 2000;
    double *sink; cudaMalloc(&sink, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 20.0, -0x40340000 * 2);
    cudaEventRecord(e0);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 20.0, -0x40340000 * 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double phe = (double)blocks * threads * iters * 32.0 * CH;
    printf("%-34s blocks/SM %d  %.3f ms  %.2f G PHE/s  %.2f cycles/PHE-warp/SMSP @1.95GHz\n", name, blocks_per_sm, ms,
           phe / ms * 1e-6, (ms * 1e-3 * 1.95e9 * sms * 4) / (phe / 32.0));
    cudaFree(sink);
}

int main()
{
    for (int bps : {4, 8}) {
        run<0, 2>("V0 4DP+SHF+FMNMX3/2 CH2", bps);
        run<1, 2>("V1 3DP+IMAD+IMAD.HI+UMIN3/2 CH2", bps);
        run<1, 4>("V1 3DP+IMAD+IMAD.HI+UMIN3/2 CH4", bps);
        run<2, 2>("V2 3DP+IMAD+IMAD.HI CH2", bps);
        run<2, 4>("V2 3DP+IMAD+IMAD.HI CH4", bps);
    }
    return 0;
}
