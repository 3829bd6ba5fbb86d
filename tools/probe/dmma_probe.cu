// dmma_probe.cu - fp64 tensor-core (mma.sync f64) rate on this GPU against the DFMA rate.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_probe dmma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int SHAPE>
__global__ void __launch_bounds__(256) dmma_kernel(double *out, int iters)
{
    double a[4], b[2], c[8][4];
    for (int i = 0; i < 4; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
    for (int i = 0; i < 2; ++i) b[i] = 1.0 - threadIdx.x * 1e-9 + i;
    for (int t = 0; t < 8; ++t)
        for (int i = 0; i < 4; ++i) c[t][i] = 0.0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            if (SHAPE == 0)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[t][0]), "+d"(c[t][1]) : "d"(a[0]), "d"(b[0]));
            else if (SHAPE == 1)
                asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                             : "+d"(c[t][0]), "+d"(c[t][1]), "+d"(c[t][2]), "+d"(c[t][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
            else if (SHAPE == 2)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+d"(c[t][0]), "+d"(c[t][1]), "+d"(c[t][2]), "+d"(c[t][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
        }
    }
    double s = 0;
    for (int t = 0; t < 8; ++t)
        for (int i = 0; i < 4; ++i) s += c[t][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dfma_kernel(double *out, int iters)
{
    double c[16], a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int i = 0; i < 16; ++i) c[i] = i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = fma(a, b, c[i]);
    double s = 0;
    for (int i = 0; i < 16; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 8 DMMA m8n8k4 + NF DFMA per iteration: do the two share a pipe (times add) or overlap (max)?
template <int NF>
__global__ void __launch_bounds__(256) mixed_kernel(double *out, int iters)
{
    double a0 = 1.0 + threadIdx.x * 1e-9, b0 = 1.0 - threadIdx.x * 1e-9, c[8][2], f[NF > 0 ? NF : 1];
    for (int t = 0; t < 8; ++t) c[t][0] = c[t][1] = 0.0;
    for (int i = 0; i < NF; ++i) f[i] = i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[t][0]), "+d"(c[t][1]) : "d"(a0), "d"(b0));
#pragma unroll
            for (int i = t * NF / 8; i < (t + 1) * NF / 8; ++i) f[i] = fma(a0, b0, f[i]);
        }
    }
    double s = 0;
    for (int t = 0; t < 8; ++t) s += c[t][0] + c[t][1];
    for (int i = 0; i < NF; ++i) s += f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class K> static double time_ms(K launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 4, iters = 20000;
    double *out; cudaMalloc(&out, sizeof(double) * grid * 256);
    const double warps = (double)grid * 8;
    double ms = time_ms([&] { dfma_kernel<<<grid, 256>>>(out, iters); });
    printf("dfma            %.2f ms  %.2f TFLOP/s\n", ms, 2.0 * 16 * 32 * iters * warps / ms / 1e9);
    ms = time_ms([&] { dmma_kernel<0><<<grid, 256>>>(out, iters); });
    printf("mma m8n8k4 f64  %.2f ms  %.2f TFLOP/s\n", ms, 2.0 * 8 * 8 * 4 * 8 * iters * warps / ms / 1e9);
    ms = time_ms([&] { dmma_kernel<1><<<grid, 256>>>(out, iters); });
    printf("mma m16n8k4 f64 %.2f ms  %.2f TFLOP/s\n", ms, 2.0 * 16 * 8 * 4 * 8 * iters * warps / ms / 1e9);
    ms = time_ms([&] { dmma_kernel<2><<<grid, 256>>>(out, iters); });
    printf("mma m16n8k8 f64 %.2f ms  %.2f TFLOP/s\n", ms, 2.0 * 16 * 8 * 8 * 8 * iters * warps / ms / 1e9);
    ms = time_ms([&] { mixed_kernel<0><<<grid, 256>>>(out, iters); });
    printf("mixed 8 dmma + 0 dfma   %.2f ms\n", ms);
    ms = time_ms([&] { mixed_kernel<8><<<grid, 256>>>(out, iters); });
    printf("mixed 8 dmma + 8 dfma   %.2f ms\n", ms);
    ms = time_ms([&] { mixed_kernel<16><<<grid, 256>>>(out, iters); });
    printf("mixed 8 dmma + 16 dfma  %.2f ms\n", ms);
    ms = time_ms([&] { mixed_kernel<32><<<grid, 256>>>(out, iters); });
    printf("mixed 8 dmma + 32 dfma  %.2f ms\n", ms);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
