// Micro-benchmark 5 (round 2): the RANSAC hot loop's distances on the tensor cores.
//  T0  reference: fp32 2 FFMA + FADD + SHF + FMNMX3/2 per point-hypothesis, 4 trials per thread (the shipped loop)
//  T1  mma.sync.m16n8k8 tf32 (16 trials x 8 points per instruction; K = 6 of 8 used by a 3-term split of the two products,
//      cp in the accumulator) + FADD + SHF + FMNMX3/2 per accumulator element; one 16-trial row tile per warp
//  T2  the same with two row tiles per warp (32 trials share every B fragment)
// Reports cycles per PHE-warp (32 point-hypotheses) per SM sub-partition, lower is better.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -cudart shared -o mix_probe5 mix_probe5.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_tf32(float &c0, float &c1, float &c2, float &c3, unsigned a0, unsigned a1, unsigned a2, unsigned a3,
                                         unsigned b0, unsigned b1)
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c0), "+f"(c1), "+f"(c2), "+f"(c3) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int V>
__global__ void __launch_bounds__(256) mix(float *sink, int iters, float thf)
{
    __shared__ float2 ptsf[256];
    __shared__ unsigned bfrag[64 * 32 * 2];                 // 64 column tiles x 32 lanes x 2 words, as the B fragments want them
    ptsf[threadIdx.x] = make_float2(1.0f + threadIdx.x * 1e-3f, 2.0f - threadIdx.x * 1e-3f);
    for (int i = threadIdx.x; i < 64 * 32 * 2; i += blockDim.x) bfrag[i] = __float_as_uint(1.0f + (i % 97) * 1e-3f) & 0xffffe000u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    float acc = 0.f;
    if (V == 0 || V >= 3) {
        constexpr int CH = 4;
        float nx[CH], ny[CH], ng[CH], mab[CH];
        unsigned bits[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) { nx[c] = 0.6f + 1e-3f * c + threadIdx.x * 1e-7f; ny[c] = 0.8f - 1e-3f * c; ng[c] = -1.0f - c; bits[c] = 0; mab[c] = 1e30f; }
        const float4 *p4 = reinterpret_cast<const float4 *>(ptsf);
        for (int it = 0; it < iters; ++it) {
#pragma unroll 4
            for (int k = 0; k < 32; k += 2) {
                const float4 pp = p4[((it + k) & 255) >> 1];
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const float fa = fabsf(fmaf(pp.y, ny[c], fmaf(pp.x, nx[c], ng[c]))) - thf;
                    const float fb = fabsf(fmaf(pp.w, ny[c], fmaf(pp.z, nx[c], ng[c]))) - thf;
                    if (V != 3) {
                        bits[c] = __funnelshift_l(__float_as_uint(fa), bits[c], 1);
                        bits[c] = __funnelshift_l(__float_as_uint(fb), bits[c], 1);
                    } else {
                        bits[c] += __float_as_uint(fa) >> 31;                 // count only (LEA.HI)
                        bits[c] += __float_as_uint(fb) >> 31;
                    }
                    if (V == 0 || V == 3) mab[c] = fminf(mab[c], fminf(fabsf(fa), fabsf(fb)));   // FMNMX3
                    if (V == 5) { mab[c] = fminf(mab[c], fabsf(fa)); mab[c] = fminf(mab[c], fabsf(fb)); }   // 2 x FMNMX
                    if (V == 6) mab[c] = fminf(mab[c], fabsf(fa * fb));                                    // FMUL + FMNMX
                    if (V == 7) mab[c] = fmaf(fa, fb, mab[c]);                                             // stand-in: one FFMA
                }
            }
        }
#pragma unroll
        for (int c = 0; c < CH; ++c) acc += bits[c] + mab[c];
    } else {
        constexpr int RT = (V == 1 || V == 2) ? V : 1;                    // row tiles (16 trials each) per warp
        unsigned a0[RT], a1[RT], a2[RT], a3[RT], bits0[RT], bits1[RT];
        float cp0[RT], cp1[RT], mab0[RT], mab1[RT];
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            a0[r] = __float_as_uint(0.6f + 1e-3f * (lane + r)) & 0xffffe000u; a1[r] = __float_as_uint(0.8f - 1e-3f * lane) & 0xffffe000u;
            a2[r] = __float_as_uint(1e-4f * (lane + 1)) & 0xffffe000u; a3[r] = __float_as_uint(2e-4f * (lane + 2 + r)) & 0xffffe000u;
            cp0[r] = -1.0f - lane; cp1[r] = -2.0f - lane - r; bits0[r] = bits1[r] = 0; mab0[r] = mab1[r] = 1e30f;
        }
        const uint2 *bf = reinterpret_cast<const uint2 *>(bfrag);
        for (int it = 0; it < iters; ++it) {
#pragma unroll 4
            for (int j = 0; j < 16; ++j) {                   // 16 column tiles of 8 points = 128 points per outer iteration
                const uint2 b = bf[((it + j) & 63) * 32 + lane];
#pragma unroll
                for (int r = 0; r < RT; ++r) {
                    float c0 = cp0[r], c1 = cp0[r], c2 = cp1[r], c3 = cp1[r];
                    mma_tf32(c0, c1, c2, c3, a0[r], a1[r], a2[r], a3[r], b.x, b.y);
                    const float f0 = fabsf(c0) - thf, f1 = fabsf(c1) - thf, f2 = fabsf(c2) - thf, f3 = fabsf(c3) - thf;
                    bits0[r] = __funnelshift_l(__float_as_uint(f0), bits0[r], 1);
                    bits0[r] = __funnelshift_l(__float_as_uint(f1), bits0[r], 1);
                    bits1[r] = __funnelshift_l(__float_as_uint(f2), bits1[r], 1);
                    bits1[r] = __funnelshift_l(__float_as_uint(f3), bits1[r], 1);
                    mab0[r] = fminf(mab0[r], fminf(fabsf(f0), fabsf(f1)));
                    mab1[r] = fminf(mab1[r], fminf(fabsf(f2), fabsf(f3)));
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RT; ++r) acc += bits0[r] + bits1[r] + mab0[r] + mab1[r];
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int V>
void run(const char *name, int blocks_per_sm)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 2000;
    float *sink; cudaMalloc(&sink, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<V><<<blocks, threads>>>(sink, iters, 20.0f);
    cudaEventRecord(e0);
    mix<V><<<blocks, threads>>>(sink, iters, 20.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    // PHE per thread and outer iteration: V0 32 points x 4 trials = 128; V>=1: 16 mma x RT row tiles x 128 PHE per warp / 32 lanes = 64 RT
    const double per_thread = (V == 0 || V >= 3) ? 128.0 : 64.0 * V;
    const double phe = (double)blocks * threads * iters * per_thread;
    const double cyc = (ms * 1e-3 * khz * 1e3 * sms * 4) / (phe / 32.0);
    printf("%-58s blocks/SM %d  %8.3f ms  %7.2f G PHE/s  %5.2f cycles/PHE-warp/SMSP\n", name, blocks_per_sm, ms, phe / ms * 1e-6, cyc);
    cudaFree(sink);
}


template <int K>
__global__ void __launch_bounds__(256) mmarate(float *sink, int iters)
{
    float c[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x; c[i][1] = 1.f; c[i][2] = 2.f; c[i][3] = 3.f; }
    unsigned a0 = 0x3f800000u + threadIdx.x, a1 = 0x3f000000u, a2 = 0x3e800000u, a3 = 0x3e000000u, b0 = 0x3f800000u, b1 = 0x3f000000u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (K == 0) asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            if (K == 1) asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};" : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(b0));
            if (K == 2) asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            if (K == 3) asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            if (K == 4) asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};" : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(b0));
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int K> void runrate(const char *name)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int blocks = sms * 8, threads = 256, iters = 4000;
    float *sink; cudaMalloc(&sink, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mmarate<K><<<blocks, threads>>>(sink, iters);
    cudaEventRecord(e0);
    mmarate<K><<<blocks, threads>>>(sink, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double n = (double)blocks * (threads / 32) * iters * 8.0;            // warp-level mma instructions
    printf("%-40s %8.3f ms  %6.2f cycles per mma per SMSP\n", name, ms, ms * 1e-3 * khz * 1e3 * sms * 4 / n);
    cudaFree(sink);
}

int main()
{
    runrate<0>("mma m16n8k8 tf32"); runrate<1>("mma m16n8k4 tf32"); runrate<2>("mma m16n8k16 bf16"); runrate<3>("mma m16n8k16 f16"); runrate<4>("mma m16n8k8 bf16");
    for (int bps : {4, 8}) {
        run<0>("T0 fp32 2FFMA+FADD+SHF+FMNMX3/2, 4 trials/thread", bps);
        run<1>("T1 mma.m16n8k8.tf32 + FADD+SHF+FMNMX3/2, 1 row tile", bps);
        run<2>("T2 mma.m16n8k8.tf32 + FADD+SHF+FMNMX3/2, 2 row tiles", bps);
        run<3>("T3 fp32 2FFMA+FADD+LEA.HI(count)+FMNMX3/2, 4 trials/thread", bps);
        run<4>("T4 fp32 2FFMA+FADD+SHF, no min tracking", bps);
        run<5>("T5 fp32 2FFMA+FADD+SHF+FMNMX (2-input, 1 per PHE)", bps);
        run<6>("T6 fp32 2FFMA+FADD+SHF+(FMUL+FMNMX)/2", bps);
        run<7>("T7 fp32 2FFMA+FADD+SHF+FFMA/2 (no min)", bps);
    }
    return 0;
}
