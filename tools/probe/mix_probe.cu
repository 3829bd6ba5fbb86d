// Micro-benchmark: how much fp64-pipe throughput survives when ALU / LDS instructions share the
// issue port?  Each thread runs CH independent chains; per step and chain: NDP fp64 ops and
// NALU integer-pipe ops (funnel shift / fminf), optionally one LDS.128 per step.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mix_probe mix_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int NALU, bool LDS, int CH>
__global__ void __launch_bounds__(256) mix(double *sink, int iters, double thr)
{
    __shared__ double2 pts[256];
    pts[threadIdx.x] = make_double2(1.0 + threadIdx.x * 1e-3, 2.0 - threadIdx.x * 1e-3);
    __syncthreads();
    double nx[CH], ny[CH], ng[CH], ss[CH];
    unsigned bits[CH];
    float mab[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) { nx[c] = 0.6 + 1e-3 * c; ny[c] = 0.8 - 1e-3 * c; ng[c] = -1.0 - c + threadIdx.x * 1e-6; ss[c] = 0; bits[c] = 0; mab[c] = 1e30f; }
    double2 p = pts[0];
    for (int it = 0; it < iters; ++it) {
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            if (LDS) p = pts[(it + k) & 255];
            else { p.x += 1e-9; }
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const double s = fma(p.y, ny[c], fma(p.x, nx[c], ng[c]));
                ss[c] = fma(s, s, ss[c]);
                const int hi = __double2hiint(fabs(s) - thr);
                if (NALU >= 1) bits[c] = __funnelshift_l((unsigned)hi, bits[c], 1);
                if (NALU >= 2) mab[c] = fminf(mab[c], fabsf(__int_as_float(hi)));
                if (NALU >= 3) bits[c] ^= (unsigned)hi >> 3;
                if (NALU >= 4) bits[c] += (unsigned)hi & 0xff;
                if (NALU == 0) ss[c] += __hiloint2double(hi, 0) * 1e-300;   // keep the DADD alive (extra DP op!)
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += ss[c] + bits[c] + mab[c];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NALU, bool LDS, int CH>
void run(const char *name, int blocks_per_sm)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 2000;
    double *sink; cudaMalloc(&sink, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<NALU, LDS, CH><<<blocks, threads>>>(sink, iters, 20.0);
    cudaEventRecord(e0);
    mix<NALU, LDS, CH><<<blocks, threads>>>(sink, iters, 20.0);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double phe = (double)blocks * threads * iters * 32.0 * CH;
    const int ndp = NALU == 0 ? 6 : 4;
    printf("%-28s blocks/SM %d  %.3f ms  %.2f G PHE/s  fp64 pipe %.1f%% of 2-cycle peak (ndp=%d)\n", name, blocks_per_sm, ms,
           phe / ms * 1e-6, 100.0 * phe * ndp / 32.0 * 2.0 / (ms * 1e-3 * 1.95e9 * sms * 4), ndp);
    cudaFree(sink);
}

int main()
{
    for (int bps : {2, 4, 8}) {
        run<1, false, 2>("4DP+1ALU noLDS CH2", bps);
        run<2, false, 2>("4DP+1.5ALU noLDS CH2", bps);
        run<2, true, 2>("4DP+1.5ALU+LDS CH2", bps);
        run<2, true, 4>("4DP+1.5ALU+LDS CH4", bps);
        run<3, true, 2>("4DP+2.5ALU+LDS CH2", bps);
        run<4, true, 2>("4DP+3.5ALU+LDS CH2", bps);
    }
    return 0;
}
