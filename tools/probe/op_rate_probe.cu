// Issue rates of the RANSAC loop's tail operations on B200: cycles per warp instruction per SM sub-partition, 8 independent
// chains per thread, 8 blocks of 256 threads per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cudart shared -o op_rate_probe op_rate_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(256) rate(float *sink, int iters, float thf, unsigned seed)
{
    float v[8], m[8];
    unsigned b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = 1.0f + i + threadIdx.x * 1e-3f; m[i] = 1e30f; b[i] = seed + i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (OP == 0) v[i] = fabsf(v[i]) - thf;                                                   // FADD with |.|
                if (OP == 1) b[i] = __funnelshift_l(__float_as_uint(v[i]), b[i], 1);                      // SHF.L.W (funnel)
                if (OP == 2) { m[i] = fminf(m[i], fminf(fabsf(v[i]), fabsf(v[(i + 1) & 7]))); }             // FMNMX3 (v changes below)
                if (OP == 3) { m[i] = fminf(m[i], fabsf(v[i])); }                                             // FMNMX
                if (OP == 4) v[i] = fmaf(v[i], v[(i + 1) & 7], m[i]);                                     // FFMA 3-reg
                if (OP == 5) b[i] = (b[i] << 1) | (__float_as_uint(v[i]) >> 31);                          // shift + LOP3 variant
                if (OP == 6) b[i] = b[i] + b[i] + (__float_as_uint(v[i]) >> 31);                          // LEA-style
                if (OP == 7) { const float f = fabsf(v[i]) - thf; b[i] = __funnelshift_l(__float_as_uint(f), b[i], 1); v[i] = f + m[i]; }   // FADD + SHF + FADD
                if (OP == 8) { const float f = fabsf(v[i]) - thf; b[i] = __funnelshift_l(__float_as_uint(f), b[i], 1); m[i] = fminf(m[i], fabsf(f)); v[i] += 1.0f; }
                if (OP == 9) b[i] = __umulhi(__float_as_uint(v[i]), 2u) + b[i];                                // IMAD.HI: count += sign bit
                if (OP == 10) b[i] = b[i] + b[(i + 1) & 7] + __float_as_uint(v[i]);                             // IADD3
                if (OP == 11) b[i] = (b[i] & __float_as_uint(v[i])) ^ b[(i + 3) & 7];                           // LOP3
                if (OP == 12) b[i] = __byte_perm(b[i], __float_as_uint(v[i]), 0x7210);                          // PRMT
                if (OP == 13) v[i] = __saturatef(v[i] * -1e30f) + m[i];                                         // FMUL.SAT + FADD
                if (OP == 14) b[i] = b[i] * 3u + __float_as_uint(v[i]);                                         // IMAD
                if (OP == 15) { const float f = fabsf(v[i]) - thf; b[i] = __umulhi(__float_as_uint(f), 2u) + b[i]; m[i] = fminf(m[i], fabsf(f)); v[i] += 1.0f; }   // FADD + IMAD.HI + FMNMX + FADD
                if (OP == 16) { const float f = fabsf(v[i]) - thf; b[i] = __umulhi(__float_as_uint(f), 2u) + b[i]; v[i] = f + m[i]; }   // FADD + IMAD.HI + FADD
            }
            if (OP == 2 || OP == 3 || OP == 9 || OP == 10 || OP == 11 || OP == 12 || OP == 14) v[r] += 1.0f;
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += v[i] + m[i] + b[i];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP> void run(const char *name, double ops_per_inner)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int blocks = sms * 8, threads = 256, iters = 4000;
    float *sink; cudaMalloc(&sink, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    rate<OP><<<blocks, threads>>>(sink, iters, 20.0f, 12345u);
    cudaEventRecord(e0);
    rate<OP><<<blocks, threads>>>(sink, iters, 20.0f, 12345u);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double n = (double)blocks * (threads / 32) * iters * 32.0;           // inner statements per warp
    printf("%-34s %8.3f ms  %6.2f cycles per inner statement per SMSP (%g instructions each)\n", name, ms,
           ms * 1e-3 * khz * 1e3 * sms * 4 / n, ops_per_inner);
    cudaFree(sink);
}

int main()
{
    run<0>("FADD |v| - thr", 1); run<1>("SHF.L.W funnel", 1); run<2>("FMNMX3 |a| |b| m", 1); run<3>("FMNMX |a| m", 1);
    run<4>("FFMA 3-reg", 1); run<5>("SHL + LOP3/LEA", 2); run<6>("IADD3-style", 2); run<7>("FADD + SHF + FADD", 3);
    run<8>("FADD + SHF + FMNMX + FADD", 4);
    run<9>("IMAD.HI (count += sign)", 1); run<10>("IADD3", 1); run<11>("LOP3", 1); run<12>("PRMT", 1); run<13>("FMUL.SAT + FADD", 2);
    run<14>("IMAD", 1); run<15>("FADD + IMAD.HI + FMNMX + FADD", 4); run<16>("FADD + IMAD.HI + FADD", 3);
    return 0;
}
