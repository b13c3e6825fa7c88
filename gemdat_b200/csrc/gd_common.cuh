// Shared plumbing for libgemdat_b200.so: error reporting, launch helpers and the
// sm_100a PTX wrappers (mbarrier + TMA bulk copy) the streaming kernels use.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/gemdat_b200.h"

#define GD_NOSITE (-1)

void gd_set_error(const char *fmt, ...);

#define GD_CUDA_CHECK(expr)                                                                  \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            gd_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,   \
                         __LINE__);                                                          \
            return GD_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

#define GD_REQUIRE(cond, code, ...)                                                          \
    do {                                                                                     \
        if (!(cond)) {                                                                       \
            gd_set_error(__VA_ARGS__);                                                       \
            return (code);                                                                   \
        }                                                                                    \
    } while (0)

#define GD_LAUNCH_CHECK()                                                                    \
    do {                                                                                     \
        cudaError_t _e = cudaGetLastError();                                                 \
        if (_e != cudaSuccess) {                                                             \
            gd_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),         \
                         __FILE__, __LINE__);                                                \
            return GD_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

struct GdDeviceInfo {
    int sm_count;
    int max_smem_optin;
};
// cached per process; sm_count < 0 when no device is usable
const GdDeviceInfo &gd_device_info();

static inline int64_t gd_ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Frame-chunk geometry shared by K1 (fused event count), K2 (events, fills) and the
// scans: the frame axis is cut into chunks of f frames, f a power of two in [32, 256]
// chosen so that a 10^9 atom-frame shard gives ~4*10^6 (chunk, atom) segments.
struct GdChunks {
    int64_t n_frames;
    int32_t n_atoms;
    int32_t f;         // frames per chunk
    int64_t n_chunks;  // ceil(n_frames / f)
};
static inline GdChunks gd_chunks(int64_t n_frames, int32_t n_atoms) {
    GdChunks g;
    g.n_frames = n_frames;
    g.n_atoms = n_atoms;
    // GD_CHUNK_SHIFT (default 20) moves the size at which chunks grow (A/B runs)
    static const int shift = [] {
        const char *e = getenv("GD_CHUNK_SHIFT");
        return e && atoi(e) > 0 ? atoi(e) : 20;
    }();
    const int64_t want = (n_frames * (int64_t)n_atoms) >> shift;
    int32_t f = 32;
    while (f < want && f < 256) f <<= 1;
    g.f = f;
    g.n_chunks = (n_frames + f - 1) / f;
    if (g.n_chunks < 1) g.n_chunks = 1;
    return g;
}

#ifdef __CUDACC__

// ---------------------------------------------------------------- mbarrier / TMA (sm_90+)
__device__ __forceinline__ uint32_t gd_smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void gd_mbar_init(uint64_t *bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(gd_smem_u32(bar)), "r"(arrivals)
                 : "memory");
}

// make barrier initialisation visible to the async (TMA) proxy
__device__ __forceinline__ void gd_fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void gd_mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(gd_smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void gd_mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(gd_smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool gd_mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        "  .reg .pred p;\n"
        "  mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "  selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(gd_smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void gd_mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!gd_mbar_try_wait(bar, parity)) {
    }
}

// 1-D TMA bulk copy global -> shared, completion reported on an mbarrier.
// dst/src 16-byte aligned, bytes a non-zero multiple of 16.  SASS: UBLKCP.
__device__ __forceinline__ void gd_tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                               uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(gd_smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(gd_smem_u32(bar))
        : "memory");
}

// 1-D bulk prefetch global -> L2 (no destination, no completion to wait for): src 16-byte
// aligned, bytes a multiple of 16.  SASS: UBLKPF.
__device__ __forceinline__ void gd_prefetch_l2(const void *gmem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem_src), "r"(bytes) : "memory");
}

// 1-D TMA bulk copy shared -> global (bulk async-group completion).
__device__ __forceinline__ void gd_tma_store_1d(void *gmem_dst, const void *smem_src,
                                                uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(gd_smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void gd_tma_store_commit() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void gd_tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void gd_tma_store_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// generic-proxy smem writes -> visible to the async proxy (before a TMA store)
__device__ __forceinline__ void gd_fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ int gd_warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// single block of 1024 threads: v[0] = 0, v[1..n] -> inclusive scan in place, i.e. v[0..n]
// become exclusive offsets with the total in v[n] (tables of per-atom / per-block counts)
template <int kUnused = 0>
__global__ void __launch_bounds__(1024) gd_k_scan_offsets(int64_t *__restrict__ v, int64_t n) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t running_s;
    if (threadIdx.x == 0) {
        running_s = 0;
        v[0] = 0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t idx = base + threadIdx.x;
        const int64_t x = idx < n ? v[idx + 1] : 0;
        int64_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int64_t warp_base = 0;
        for (int w = 0; w < warp; ++w) warp_base += warp_tot[w];
        const int64_t running = running_s;
        __syncthreads();
        if (idx < n) v[idx + 1] = running + warp_base + incl;
        if (threadIdx.x == 1023) running_s = running + warp_base + incl;
        __syncthreads();
    }
}

#endif  // __CUDACC__
