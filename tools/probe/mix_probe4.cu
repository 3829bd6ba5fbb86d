// Micro-benchmark 4 (round 2): candidate per point-hypothesis instruction mixes for the RANSAC hot loop.
//  V0  round-1 loop: 3 DFMA (s: 2, ss: 1) + DADD (|s| - thr) + SHF + FMNMX3/2
//  V1  2 DFMA (s) + FADD on the HIGH WORDS (|hi(s)| as float - hi(thr) as float) + SHF + FMNMX3/2   (sum of squares from moments)
//  V2  2 DFMA (s) + FFMA.SAT (indicator in [0,1], multiples of 1/8 inside the band) + FADD (count) + FFMA (sum ind^2)
//  V3  V1 with the distance in fp32: 2 FFMA + FADD + SHF + FMNMX3/2
//  V4  V2 with the distance in fp32: 2 FFMA + FFMA.SAT + FADD + FFMA
// Reports cycles per PHE-warp per SM sub-partition (lower is better; a DFMA alone is 2).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cudart shared -o mix_probe4 mix_probe4.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int V, int CH>
__global__ void __launch_bounds__(256) mix(float *sink, int iters, double thr)
{
    __shared__ double2 pts[256];
    __shared__ float2 ptsf[256];
    pts[threadIdx.x] = make_double2(1.0 + threadIdx.x * 1e-3, 2.0 - threadIdx.x * 1e-3);
    ptsf[threadIdx.x] = make_float2(1.0f + threadIdx.x * 1e-3f, 2.0f - threadIdx.x * 1e-3f);
    __syncthreads();
    float acc = 0.f;
    const float thf = __int_as_float(__double2hiint(thr));
    if (V <= 2) {
        double nx[CH], ny[CH], ng[CH], ss[CH];
        float mab[CH], c1[CH], c2[CH];
        unsigned bits[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            nx[c] = 0.6 + 1e-3 * c + threadIdx.x * 1e-7; ny[c] = 0.8 - 1e-3 * c - threadIdx.x * 1e-7; ng[c] = -1.0 - c + threadIdx.x * 1e-6;
            ss[c] = 0; bits[c] = 0; mab[c] = 1e30f; c1[c] = 0.f; c2[c] = 0.f;
        }
        const float K = 1e6f, thK = thf * K + 0.5f;
        for (int it = 0; it < iters; ++it) {
#pragma unroll 8
            for (int k = 0; k < 32; ++k) {
                const double2 p = pts[(it + k) & 255];
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const double s = fma(p.y, ny[c], fma(p.x, nx[c], ng[c]));
                    if (V == 0) {
                        ss[c] = fma(s, s, ss[c]);
                        const int hi = __double2hiint(fabs(s) - thr);
                        bits[c] = __funnelshift_l((unsigned)hi, bits[c], 1);
                        mab[c] = fminf(mab[c], fabsf(__int_as_float(hi)));
                    } else if (V == 1) {
                        const float f = fabsf(__int_as_float(__double2hiint(s))) - thf;
                        bits[c] = __funnelshift_l(__float_as_uint(f), bits[c], 1);
                        mab[c] = fminf(mab[c], fabsf(f));
                    } else {
                        const float ind = __saturatef(fmaf(-fabsf(__int_as_float(__double2hiint(s))), K, thK));
                        c1[c] += ind;
                        c2[c] = fmaf(ind, ind, c2[c]);
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CH; ++c) acc += (float)ss[c] + bits[c] + mab[c] + c1[c] + c2[c];
    } else {
        float nx[CH], ny[CH], ng[CH], mab[CH], c1[CH], c2[CH];
        unsigned bits[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            nx[c] = 0.6f + 1e-3f * c + threadIdx.x * 1e-7f; ny[c] = 0.8f - 1e-3f * c - threadIdx.x * 1e-7f; ng[c] = -1.0f - c + threadIdx.x * 1e-6f;
            bits[c] = 0; mab[c] = 1e30f; c1[c] = 0.f; c2[c] = 0.f;
        }
        const float th = (float)thr, K = 1e6f, thK = th * K + 0.5f;
        for (int it = 0; it < iters; ++it) {
#pragma unroll 8
            for (int k = 0; k < 32; ++k) {
                const float2 p = ptsf[(it + k) & 255];
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const float s = fmaf(p.y, ny[c], fmaf(p.x, nx[c], ng[c]));
                    if (V == 3) {
                        const float f = fabsf(s) - th;
                        bits[c] = __funnelshift_l(__float_as_uint(f), bits[c], 1);
                        mab[c] = fminf(mab[c], fabsf(f));
                    } else {
                        const float ind = __saturatef(fmaf(-fabsf(s), K, thK));
                        c1[c] += ind;
                        c2[c] = fmaf(ind, ind, c2[c]);
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CH; ++c) acc += bits[c] + mab[c] + c1[c] + c2[c];
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int V, int CH>
void run(const char *name, int blocks_per_sm)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 2000;
    float *sink; cudaMalloc(&sink, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 20.0);
    cudaEventRecord(e0);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 20.0);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double phe = (double)blocks * threads * iters * 32.0 * CH;
    const double cyc = (ms * 1e-3 * khz * 1e3 * sms * 4) / (phe / 32.0);
    printf("%-44s CH%d blocks/SM %d  %8.3f ms  %7.2f G PHE/s  %5.2f cycles/PHE-warp/SMSP  (8 algorithmic flop/PHE = %5.1f %% of the DFMA peak)\n",
           name, CH, blocks_per_sm, ms, phe / ms * 1e-6, cyc, 100.0 * 8.0 / cyc);
    cudaFree(sink);
}

int main()
{
    for (int bps : {4, 8}) {
        run<0, 2>("V0 3DFMA+DADD+SHF+FMNMX3/2 (round 1)", bps);
        run<0, 4>("V0 3DFMA+DADD+SHF+FMNMX3/2 (round 1)", bps);
        run<1, 2>("V1 2DFMA+FADD(hi)+SHF+FMNMX3/2", bps);
        run<1, 4>("V1 2DFMA+FADD(hi)+SHF+FMNMX3/2", bps);
        run<1, 6>("V1 2DFMA+FADD(hi)+SHF+FMNMX3/2", bps);
        run<2, 4>("V2 2DFMA+FFMA.SAT+FADD+FFMA", bps);
        run<2, 6>("V2 2DFMA+FFMA.SAT+FADD+FFMA", bps);
        run<3, 4>("V3 fp32 2FFMA+FADD+SHF+FMNMX3/2", bps);
        run<3, 8>("V3 fp32 2FFMA+FADD+SHF+FMNMX3/2", bps);
        run<4, 4>("V4 fp32 2FFMA+FFMA.SAT+FADD+FFMA", bps);
        run<4, 8>("V4 fp32 2FFMA+FFMA.SAT+FADD+FFMA", bps);
    }
    return 0;
}
