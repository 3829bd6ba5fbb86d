// Micro-benchmark 3: fp32 screening variants of the per point-hypothesis step.
//  F0: 4 FFMA (s: 2, ss: 1, u = s*s - thr^2: 1) + SHF + FMNMX3/2 + LDS.64
//  F1: same with packed fma.rn.f32x2 over two trials (2 FFMA2 per PHE) + SHF + FMNMX3/2
//  F2: F1 but counting with one SHF per PAIR of trials is impossible -> count via PRMT+... (skip)
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long pack(float a, float b) {
    return ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(a);
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int V, int CH>
__global__ void __launch_bounds__(256) mix(float *sink, int iters, float thr2)
{
    __shared__ float2 pts[256];
    pts[threadIdx.x] = make_float2(1.0f + threadIdx.x * 1e-3f, 2.0f - threadIdx.x * 1e-3f);
    __syncthreads();
    float acc = 0.f;
    if (V == 0) {
        float nx[CH], ny[CH], ng[CH], ss[CH], mab[CH];
        unsigned bits[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) { nx[c] = 0.6f + 1e-3f * c + threadIdx.x * 1e-7f; ny[c] = 0.8f - 1e-3f * c - threadIdx.x * 1e-7f; ng[c] = -1.0f - c + threadIdx.x * 1e-6f; ss[c] = 0; bits[c] = 0; mab[c] = 1e30f; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll 8
            for (int k = 0; k < 32; ++k) {
                const float2 p = pts[(it + k) & 255];
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const float s = fmaf(p.y, ny[c], fmaf(p.x, nx[c], ng[c]));
                    ss[c] = fmaf(s, s, ss[c]);
                    const float u = fmaf(s, s, -thr2);
                    bits[c] = __funnelshift_l(__float_as_uint(u), bits[c], 1);
                    mab[c] = fminf(mab[c], fabsf(u));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CH; ++c) acc += ss[c] + bits[c] + mab[c];
    } else {
        // CH trials = CH/2 packed pairs
        unsigned long long nx[CH / 2], ny[CH / 2], ng[CH / 2], ss[CH / 2];
        float mab[CH / 2];
        unsigned bits[CH];
        const unsigned long long mthr = pack(-thr2, -thr2);
#pragma unroll
        for (int c = 0; c < CH / 2; ++c) { nx[c] = pack(0.6f + 1e-3f * c + threadIdx.x * 1e-7f, 0.61f + threadIdx.x * 1e-7f); ny[c] = pack(0.8f - 1e-3f * c - threadIdx.x * 1e-7f, 0.79f - threadIdx.x * 1e-7f); ng[c] = pack(-1.0f - c + threadIdx.x * 1e-6f, -2.f); ss[c] = 0; mab[c] = 1e30f; bits[2 * c] = bits[2 * c + 1] = 0; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll 8
            for (int k = 0; k < 32; ++k) {
                const float2 p = pts[(it + k) & 255];
                const unsigned long long px = pack(p.x, p.x), py = pack(p.y, p.y);
#pragma unroll
                for (int c = 0; c < CH / 2; ++c) {
                    const unsigned long long s = fma2(py, ny[c], fma2(px, nx[c], ng[c]));
                    ss[c] = fma2(s, s, ss[c]);
                    const unsigned long long u = fma2(s, s, mthr);
                    const unsigned u0 = (unsigned)u, u1 = (unsigned)(u >> 32);
                    bits[2 * c] = __funnelshift_l(u0, bits[2 * c], 1);
                    bits[2 * c + 1] = __funnelshift_l(u1, bits[2 * c + 1], 1);
                    mab[c] = fminf(mab[c], fminf(fabsf(__uint_as_float(u0)), fabsf(__uint_as_float(u1))));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CH / 2; ++c) acc += __uint_as_float((unsigned)ss[c]) + __uint_as_float((unsigned)(ss[c] >> 32)) + bits[2 * c] + bits[2 * c + 1] + mab[c];
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int V, int CH>
void run(const char *name, int blocks_per_sm)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 2000;
    float *sink; cudaMalloc(&sink, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 400.f);
    cudaEventRecord(e0);
    mix<V, CH><<<blocks, threads>>>(sink, iters, 400.f);
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
        run<0, 2>("F0 4FFMA+SHF+FMNMX3/2 CH2", bps);
        run<0, 4>("F0 4FFMA+SHF+FMNMX3/2 CH4", bps);
        run<0, 8>("F0 4FFMA+SHF+FMNMX3/2 CH8", bps);
        run<1, 2>("F1 2FFMA2+SHF+FMNMX3/2 CH2", bps);
        run<1, 4>("F1 2FFMA2+SHF+FMNMX3/2 CH4", bps);
        run<1, 8>("F1 2FFMA2+SHF+FMNMX3/2 CH8", bps);
    }
    return 0;
}
