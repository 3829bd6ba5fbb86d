// lslam_common.cuh - handle, error plumbing and the launch macro shared by the C-ABI TU.
#pragma once
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#ifndef LSLAM_EMU
#include <cuda_runtime.h>
#define LSLAM_LAUNCH(kern, grid, block, smem, stream, ...) kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#else
// tests/emu/cuda_emu.h is force-included by the emulation build (tests/emu/build_emu.py)
#define LSLAM_LAUNCH(kern, grid, block, smem, stream, ...) emu::launch(kern, dim3(grid), dim3(block), (smem), __VA_ARGS__)
#endif

namespace lslam {

struct Handle {
    int device = 0;
    int sms = 0;
    int smem_optin = 0;
    int smem_sm = 0;
    int64_t launches = 0;
    int opt_tpt = 0, opt_slots = 0, opt_threads = 0;
    int opt_kernel = 0, opt_scorers = 0, opt_posts = 0, opt_ring = 0;
    int opt_aug_pzx = 1;               // ukf_aug.pzx: 1 paired cross-covariance product (default), 0 literal
    bool aug_attr_set[10] = {};
    double *express_tab_d = nullptr;   // sin/cos tables of the packet decoder (csrc/express.cuh)
    unsigned long long *counters_d = nullptr;
    char err[512] = "";
};

// set the handle's device for the duration of a call, restore the caller's afterwards
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

static inline int fail(Handle *h, int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(h->err, sizeof h->err, fmt, ap);
    va_end(ap);
    return code;
}

#define LSLAM_CUDA_TRY(h, expr)                                                                   \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail((h), -4 /* LSLAM_ERR_CUDA */, "%s: %s", #expr, cudaGetErrorString(e_));   \
    } while (0)

} // namespace lslam
