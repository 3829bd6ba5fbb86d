// Minimal CPU emulation of the CUDA subset used by lidar_slam_b200/csrc.
//
// TEST INFRASTRUCTURE.  It exists so the *actual kernel sources* (and the C-ABI host code
// around them) can be compiled with g++ and exercised against the oracle in the CPU-only
// test suite, where no GPU exists.  The result (tests/emu/liblslam_emu.so) is loaded only
// by tests/; the product loader (lidar_slam_b200/_lib.py) knows nothing about it and
// fails loudly when the real sm_100a library is missing.
//
// Model: one OS thread per CUDA thread, one block at a time.  __syncthreads is a barrier
// over the block's threads; warp collectives (__shfl*, __ballot, __reduce*, __any, __all,
// __syncwarp) are a barrier over the 32 threads of the warp plus an exchange array.  All
// collectives must be reached by every lane of the warp (full mask), which is how the
// kernels are written anyway.  blockDim.x must be a multiple of 32.
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#define LSLAM_EMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __restrict__
#define __align__(n) __attribute__((aligned(n)))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct alignas(16) double2 { double x, y; };
struct alignas(8) float2 { float x, y; };
struct alignas(8) int2 { int x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) uint2 { unsigned x, y; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(16) int4 { int x, y, z, w; };
static inline double2 make_double2(double x, double y) { return double2{x, y}; }
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline int2 make_int2(int x, int y) { return int2{x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

namespace emu {

// Sense-reversing barrier on a mutex/condvar (std::barrier spins too eagerly for ~1000
// threads on 8 cores).
class Barrier {
  public:
    explicit Barrier(unsigned n) : n_(n), count_(0), gen_(0) {}
    void wait() {
        std::unique_lock<std::mutex> lk(m_);
        unsigned g = gen_;
        if (++count_ == n_) { count_ = 0; ++gen_; cv_.notify_all(); }
        else cv_.wait(lk, [&] { return gen_ != g; });
    }
  private:
    std::mutex m_;
    std::condition_variable cv_;
    unsigned n_, count_, gen_;
};

struct WarpCtx {
    Barrier bar{32};
    uint64_t slot[32];
    // collectives over a sub-mask (lane groups that run divergent code): one barrier per distinct mask
    std::mutex sub_m;
    std::map<unsigned, std::unique_ptr<Barrier>> sub;
    Barrier &bar_for(unsigned mask) {
        if (mask == 0xffffffffu) return bar;
        std::lock_guard<std::mutex> lk(sub_m);
        auto it = sub.find(mask);
        if (it == sub.end()) it = sub.emplace(mask, std::make_unique<Barrier>((unsigned)__builtin_popcount(mask))).first;
        return *it->second;
    }
};

struct BlockCtx {
    explicit BlockCtx(unsigned nthreads, size_t smem_bytes)
        : bar(nthreads), warps((nthreads + 31) / 32), smem(smem_bytes + 16) {
        for (auto &w : warps) w = std::make_unique<WarpCtx>();
    }
    Barrier bar;
    std::vector<std::unique_ptr<WarpCtx>> warps;
    std::vector<unsigned char> smem;
};

struct ThreadState {
    BlockCtx *blk = nullptr;
    WarpCtx *warp = nullptr;
    int lane = 0;
};
inline thread_local ThreadState ts;

inline unsigned char *block_smem() {
    uintptr_t p = (uintptr_t)ts.blk->smem.data();
    return (unsigned char *)((p + 15) & ~(uintptr_t)15);
}

template <class T> inline uint64_t to_bits(T v) {
    static_assert(sizeof(T) <= 8, "shuffle payload too wide");
    uint64_t b = 0; std::memcpy(&b, &v, sizeof(T)); return b;
}
template <class T> inline T from_bits(uint64_t b) { T v; std::memcpy(&v, &b, sizeof(T)); return v; }

template <class T> inline T exchange(T v, int src_lane, unsigned mask = 0xffffffffu) {
    WarpCtx &w = *ts.warp;
    Barrier &b = w.bar_for(mask);
    w.slot[ts.lane] = to_bits(v);
    b.wait();
    T r = from_bits<T>(w.slot[src_lane & 31]);
    b.wait();
    return r;
}

// every lane's value to every lane (full warp): the operand exchange of an emulated warp-wide MMA
template <class T> inline void gather_all(T v, T *out) {
    WarpCtx &w = *ts.warp;
    w.slot[ts.lane] = to_bits(v);
    w.bar.wait();
    for (int l = 0; l < 32; ++l) out[l] = from_bits<T>(w.slot[l]);
    w.bar.wait();
}

} // namespace emu

inline thread_local dim3 threadIdx, blockIdx, blockDim, gridDim;

// ---- block / warp synchronisation and collectives ----
static inline void __syncthreads() { emu::ts.blk->bar.wait(); }
static inline void __syncwarp(unsigned mask = 0xffffffffu) { emu::ts.warp->bar_for(mask).wait(); }
template <class T> static inline T __shfl_sync(unsigned m, T v, int src) { return emu::exchange(v, src, m); }
template <class T> static inline T __shfl_xor_sync(unsigned m, T v, int x) { return emu::exchange(v, emu::ts.lane ^ x, m); }
template <class T> static inline T __shfl_down_sync(unsigned m, T v, unsigned d) {
    int src = emu::ts.lane + (int)d; return emu::exchange(v, src < 32 ? src : emu::ts.lane, m);
}
template <class T> static inline T __shfl_up_sync(unsigned m, T v, unsigned d) {
    int src = emu::ts.lane - (int)d; return emu::exchange(v, src >= 0 ? src : emu::ts.lane, m);
}
static inline unsigned __ballot_sync(unsigned mask, int pred) {
    emu::WarpCtx &w = *emu::ts.warp;
    emu::Barrier &b = w.bar_for(mask);
    w.slot[emu::ts.lane] = pred ? 1 : 0;
    b.wait();
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) if ((mask >> l) & 1u) r |= (unsigned)(w.slot[l] & 1) << l;
    b.wait();
    return r;
}
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == 0xffffffffu; }
template <class T, class F> static inline T emu_reduce(T v, F f) {
    emu::WarpCtx &w = *emu::ts.warp;
    w.slot[emu::ts.lane] = emu::to_bits(v);
    w.bar.wait();
    T r = emu::from_bits<T>(w.slot[0]);
    for (int l = 1; l < 32; ++l) r = f(r, emu::from_bits<T>(w.slot[l]));
    w.bar.wait();
    return r;
}
static inline int __reduce_max_sync(unsigned, int v) { return emu_reduce(v, [](int a, int b) { return a > b ? a : b; }); }
static inline int __reduce_min_sync(unsigned, int v) { return emu_reduce(v, [](int a, int b) { return a < b ? a : b; }); }
static inline int __reduce_add_sync(unsigned, int v) { return emu_reduce(v, [](int a, int b) { return a + b; }); }
static inline unsigned __reduce_max_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a > b ? a : b; }); }
static inline unsigned __reduce_min_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a < b ? a : b; }); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a + b; }); }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a | b; }); }
static inline unsigned __reduce_xor_sync(unsigned, unsigned v) { return emu_reduce(v, [](unsigned a, unsigned b) { return a ^ b; }); }

// ---- scalar intrinsics ----
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline unsigned __fns(unsigned mask, unsigned base, int offset) {   // offset > 0 only
    int cnt = 0;
    for (unsigned b = base; b < 32; ++b) if ((mask >> b) & 1u) { if (++cnt == offset) return b; }
    return 0xffffffffu;
}
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline unsigned __brev(unsigned v) { unsigned r = 0; for (int i = 0; i < 32; ++i) r |= ((v >> i) & 1u) << (31 - i); return r; }
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) {
    s &= 31; return s ? (hi << s) | (lo >> (32 - s)) : hi;
}
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
static inline int __double2hiint(double d) { uint64_t b; std::memcpy(&b, &d, 8); return (int)(b >> 32); }
static inline int __double2loint(double d) { uint64_t b; std::memcpy(&b, &d, 8); return (int)(b & 0xffffffffu); }
static inline double __hiloint2double(int hi, int lo) {
    uint64_t b = ((uint64_t)(unsigned)hi << 32) | (unsigned)lo; double d; std::memcpy(&d, &b, 8); return d;
}
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
static inline float __double2float_ru(double d) { float f = (float)d; return (double)f < d ? std::nextafterf(f, INFINITY) : f; }
static inline unsigned __float_as_uint(float f) { unsigned i; std::memcpy(&i, &f, 4); return i; }
static inline float __uint_as_float(unsigned i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline long long __double_as_longlong(double d) { long long i; std::memcpy(&i, &d, 8); return i; }
static inline double __longlong_as_double(long long i) { double d; std::memcpy(&d, &i, 8); return d; }
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
static inline double rsqrt(double x) { return 1.0 / sqrt(x); }
// sincos(double, double*, double*) comes from glibc (_GNU_SOURCE is on for g++)
// round-to-nearest IEEE ops (the emulation TU is compiled with -ffp-contract=off)
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline double __dsqrt_rn(double a) { return sqrt(a); }
static inline double __fma_rn(double a, double b, double c) { return fma(a, b, c); }
template <class T> static inline T __ldg(const T *p) { return *p; }

// ---- atomics (on ordinary memory) ----
static inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned atomicAdd(unsigned *p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline int atomicMax(int *p, int v) {
    int old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
static inline unsigned long long atomicMax(unsigned long long *p, unsigned long long v) {
    unsigned long long old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
static inline unsigned atomicOr(unsigned *p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
static inline int atomicOr(int *p, int v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }

// ---- runtime API stubs: "device memory" is host memory ----
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef struct { std::chrono::steady_clock::time_point t; } *cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount, cudaDevAttrMaxSharedMemoryPerBlockOptin, cudaDevAttrMaxSharedMemoryPerMultiprocessor,
                      cudaDevAttrComputeCapabilityMajor, cudaDevAttrComputeCapabilityMinor };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return cudaSuccess; }
static inline cudaError_t cudaDeviceGetAttribute(int *v, cudaDeviceAttr a, int) {
    switch (a) {
    case cudaDevAttrMultiProcessorCount: *v = 4; break;          // keeps emulated grids small
    case cudaDevAttrMaxSharedMemoryPerBlockOptin: *v = 232448; break;
    case cudaDevAttrMaxSharedMemoryPerMultiprocessor: *v = 233472; break;
    case cudaDevAttrComputeCapabilityMajor: *v = 10; break;
    case cudaDevAttrComputeCapabilityMinor: *v = 0; break;
    }
    return cudaSuccess;
}
template <class K> static inline cudaError_t cudaFuncSetAttribute(K, cudaFuncAttribute, int) { return cudaSuccess; }
// two CTAs "resident" per SM: persistent kernels then walk several tiles per CTA in the emulation as well
template <class K> static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int *n, K, int, size_t) { *n = 2; return cudaSuccess; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = std::calloc(1, n ? n : 1); return *p ? cudaSuccess : 2; }
static inline cudaError_t cudaFree(void *p) { std::free(p); return cudaSuccess; }
// in the emulation a __device__ array IS host memory, so its name is its address
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = (void *)&sym; return cudaSuccess; }
static inline cudaError_t cudaMemset(void *p, int v, size_t n) { std::memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t = nullptr) { std::memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t hgt, cudaMemcpyKind, cudaStream_t = nullptr) {
    for (size_t r = 0; r < hgt; ++r) std::memcpy((char *)d + r * dp, (const char *)s + r * sp, w);
    return cudaSuccess;
}
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline const char *cudaGetErrorString(cudaError_t e) { return e ? "emulated error" : "no error"; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new std::remove_pointer<cudaEvent_t>::type; return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->t = std::chrono::steady_clock::now(); return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) {
    *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count(); return cudaSuccess;
}

namespace emu {

template <class K, class... A>
inline void launch(K kernel, dim3 grid, dim3 block, size_t smem, A... args) {
    if (block.x % 32 != 0 || block.y != 1 || block.z != 1) {
        std::fprintf(stderr, "cuda_emu: only 1-D blocks with blockDim.x %% 32 == 0\n");
        std::abort();
    }
    for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
    for (unsigned b = 0; b < grid.x; ++b) {
        BlockCtx ctx(block.x, smem);
        std::vector<std::thread> th;
        th.reserve(block.x);
        for (unsigned t = 0; t < block.x; ++t) {
            th.emplace_back([&, t, b, by, bz] {
                threadIdx = dim3(t); blockIdx = dim3(b, by, bz); blockDim = block; gridDim = grid;
                ts.blk = &ctx; ts.warp = ctx.warps[t / 32].get(); ts.lane = (int)(t % 32);
                kernel(args...);
            });
        }
        for (auto &x : th) x.join();
    }
}

} // namespace emu
