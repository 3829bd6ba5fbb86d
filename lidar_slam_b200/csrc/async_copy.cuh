// async_copy.cuh - mbarrier + TMA bulk copy (cp.async.bulk, SASS UBLKCP) wrappers for sm_100a.
//
// A `Mbar` lives in shared memory.  Protocol used by the kernels:
//   mbar_init(bar, n)            n arrivals complete a phase (one thread, then fence + block sync)
//   mbar_arrive(bar)             one arrival (release at CTA scope)
//   mbar_expect_tx(bar, bytes)   one arrival that also arms `bytes` of asynchronous copy traffic
//   bulk_g2s(dst, src, bytes, bar)  global -> shared bulk copy by the TMA engine, completes on bar
//   mbar_wait(bar, parity)       spin until the phase with this parity has completed (acquire)
// The CPU emulation used by tests/emu implements the same protocol on std::atomic.
#pragma once
#include <stdint.h>

#ifndef LSLAM_EMU
#include <cuda.h>
#include <cuda_runtime.h>

namespace lslam {

typedef uint64_t Mbar;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(Mbar *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the initialised barriers visible to the async proxy (TMA) before first use
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive(Mbar *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(Mbar *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(Mbar *bar, unsigned parity)
{
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded spin: a protocol bug must end in an error the host sees, never in a hung GPU.
__device__ __forceinline__ void mbar_wait(Mbar *bar, unsigned parity)
{
    for (unsigned tries = 0; !mbar_try_wait(bar, parity); ++tries) {
        if (tries > (1u << 24)) __trap();
    }
}
// The same for a thread that is far ahead of what it waits for (a TMA producer waiting for a stage to be
// released): sleep between polls, so the spin does not take issue slots from the warps doing the math.
__device__ __forceinline__ void mbar_wait_relaxed(Mbar *bar, unsigned parity, unsigned ns = 256)
{
    for (unsigned tries = 0; !mbar_try_wait(bar, parity); ++tries) {
        __nanosleep(ns);
        if (tries > (1u << 22)) __trap();
    }
}
// generic-proxy writes to shared memory (e.g. a counter reset) ordered before later async-proxy traffic
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, Mbar *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- tiled TMA (cp.async.bulk.tensor, SASS UTMALDG): a box of a 3-D tensor global -> shared, addressed
// through a tensor map the host encodes (tmap_encode_3d in lslam.cu); coordinates outside the tensor are
// zero-filled by the copy engine and still count towards the barrier's byte total.
typedef CUtensorMap Tmap;
__device__ __forceinline__ void tma_load_3d(void *dst, const Tmap *map, int c0, int c1, int c2, Mbar *bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
// ---- per-thread asynchronous 16-byte copies global -> shared (cp.async, SASS LDGSTS): no register staging, so a
// thread can have any number of them in flight; cp_async_wait_all() blocks until this thread's copies have landed.
__device__ __forceinline__ void cp_async16(void *dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
#define LSLAM_GRID_CONSTANT __grid_constant__
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

} // namespace lslam

#else // ---------------------------------------------------------------- CPU emulation (tests/emu)
#include <atomic>
#include <cstring>
#include <new>
#include <thread>

namespace lslam {

struct Mbar {
    std::atomic<uint64_t> state;     // [63:32] completed phases, [31:0] unused
    std::atomic<int64_t> pending;    // arrivals still missing in the current phase
    std::atomic<int64_t> tx;         // armed - completed bytes
    int64_t count;
};

inline void mbar_init(Mbar *bar, unsigned count)
{
    new (bar) Mbar();
    bar->state.store(0); bar->pending.store(count); bar->tx.store(0); bar->count = count;
}
inline void mbar_fence_init() {}
inline void fence_proxy_async() {}
inline void mbar_emu_check(Mbar *bar)
{
    // the thread that brings (pending, tx) to (0, 0) completes the phase; a spurious double check is
    // prevented by resetting `pending` with a compare-exchange
    int64_t zero = 0;
    if (bar->tx.load() == 0 && bar->pending.compare_exchange_strong(zero, bar->count)) bar->state.fetch_add(1ull << 32);
}
inline void mbar_arrive(Mbar *bar) { bar->pending.fetch_sub(1); mbar_emu_check(bar); }
inline void mbar_expect_tx(Mbar *bar, unsigned bytes) { bar->tx.fetch_add(bytes); bar->pending.fetch_sub(1); mbar_emu_check(bar); }
inline bool mbar_try_wait(Mbar *bar, unsigned parity) { return ((bar->state.load() >> 32) & 1u) != parity; }
inline void mbar_wait(Mbar *bar, unsigned parity) { while (!mbar_try_wait(bar, parity)) std::this_thread::yield(); }
inline void mbar_wait_relaxed(Mbar *bar, unsigned parity, unsigned = 0) { mbar_wait(bar, parity); }
inline void bulk_g2s(void *dst, const void *src, unsigned bytes, Mbar *bar)
{
    std::memcpy(dst, src, bytes);
    bar->tx.fetch_sub(bytes);
    mbar_emu_check(bar);
}

// emulated tensor map: fp64 elements, dim[0] fastest; strides in elements
struct Tmap {
    const double *base;
    long long dim[3], stride[2];
    int box[3];
};
inline void tma_load_3d(void *dst, const Tmap *m, int c0, int c1, int c2, Mbar *bar)
{
    double *d = static_cast<double *>(dst);
    for (int z = 0; z < m->box[2]; ++z)
        for (int y = 0; y < m->box[1]; ++y)
            for (int x = 0; x < m->box[0]; ++x) {
                const long long gx = c0 + x, gy = c1 + y, gz = c2 + z;
                const bool in = gx >= 0 && gx < m->dim[0] && gy >= 0 && gy < m->dim[1] && gz >= 0 && gz < m->dim[2];
                *d++ = in ? m->base[gz * m->stride[1] + gy * m->stride[0] + gx] : 0.0;
            }
    bar->tx.fetch_sub((int64_t)sizeof(double) * m->box[0] * m->box[1] * m->box[2]);
    mbar_emu_check(bar);
}
inline void cp_async16(void *dst, const void *src) { std::memcpy(dst, src, 16); }
inline void cp_async_wait_all() {}
#define LSLAM_GRID_CONSTANT
inline void prefetch_l2(const void *) {}

} // namespace lslam
#endif
