// Collective jumps: the pair search of reference
// src/gemdat/collective.py:61-106 on the device.
//
//   gd_collective_sort   jump rows ordered by (stop time, start time), stable   (:68)
//   gd_collective_close  [n_sites, n_sites] bytes: minimum-image distance < max_dist (:83-88)
//   gd_collective_count  one warp per jump i walks its partners j > i 32 at a time;
//                        the first j with start_j - stop_i > max_steps ends the walk of i
//                        (:76-77, a break on a non-monotonic key, kept as is), a j with
//                        start_i - stop_j > max_steps or on the same atom is skipped
//                        (:78-81); counts -> exclusive scan -> number of pairs, and the
//                        number of jumps in no pair (:100-101)
//   gd_collective_fill   same walk, writes the (i, j) rows in the reference's emission
//                        order (i ascending, j ascending within i)
//
// The walk reads 24 B per (i, j) candidate from L2 (the sorted table of a window stays
// resident) and four bytes of the close matrix; it is latency/L2 bound, not HBM bound.
#include "gd_common.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace {

constexpr int kRow = 6;  // atom, start site, destination site, start time, stop time, residence

__global__ void k_sort_keys(const int32_t *__restrict__ jumps, int64_t n, uint64_t *__restrict__ keys,
                            int32_t *__restrict__ idx) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint32_t start = (uint32_t)jumps[r * kRow + 3];
    const uint32_t stop = (uint32_t)jumps[r * kRow + 4];
    keys[r] = ((uint64_t)stop << 32) | start;
    idx[r] = (int32_t)r;
}

__global__ void k_gather_rows(const int32_t *__restrict__ jumps, const int32_t *__restrict__ idx, int64_t n,
                              int32_t *__restrict__ sorted) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= n * 3) return;
    const int64_t r = t / 3;
    const int c = (int)(t - r * 3);
    const int2 *src = reinterpret_cast<const int2 *>(jumps + (int64_t)idx[r] * kRow);
    reinterpret_cast<int2 *>(sorted + r * kRow)[c] = src[c];
}

__global__ void k_close(const double *__restrict__ dist, int64_t n, double max_dist, uint8_t *__restrict__ close) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t < n) close[t] = dist[t] < max_dist ? 1 : 0;
}

struct JumpRow {
    int32_t atom, s, d, start, stop;
};

__device__ __forceinline__ JumpRow load_row(const int32_t *__restrict__ ev, int64_t r) {
    const int2 *p = reinterpret_cast<const int2 *>(ev + r * kRow);
    const int2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    JumpRow o;
    o.atom = a.x;
    o.s = a.y;
    o.d = b.x;
    o.start = b.y;
    o.stop = c.x;
    return o;
}

// One warp per jump i.  kFill=false: counts[i] = pairs (i, *), hit[] marks every jump in a
// pair.  kFill=true: offsets[i] is where the pairs of i start in `pairs`.
template <bool kFill>
__global__ void __launch_bounds__(256)
k_collective(const int32_t *__restrict__ ev, int64_t n, const uint8_t *__restrict__ close, int32_t n_sites,
             int32_t max_steps, int64_t *__restrict__ counts_or_offsets, int32_t *__restrict__ pairs,
             uint8_t *__restrict__ hit) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (gridDim.x * (int64_t)blockDim.x) >> 5;
    // jumps late in the table have short walks: hand them out interleaved
    for (int64_t i = warp0; i < n - 1; i += n_warps) {
        const JumpRow a = load_row(ev, i);
        const uint8_t *cs = close + (int64_t)a.s * n_sites;
        const uint8_t *cd = close + (int64_t)a.d * n_sites;
        int64_t cnt = 0;
        int64_t off = kFill ? counts_or_offsets[i] : 0;
        for (int64_t base = i + 1; base < n; base += 32) {
            const int64_t j = base + lane;
            const bool valid = j < n;
            JumpRow b;
            if (valid) b = load_row(ev, j);
            const bool brk = valid && (b.start - a.stop > max_steps);
            const unsigned bmask = __ballot_sync(0xffffffffu, brk);
            const int limit = bmask ? __ffs(bmask) - 1 : 32;
            bool ok = valid && lane < limit && !(a.start - b.stop > max_steps) && a.atom != b.atom;
            if (ok) ok = (__ldg(cs + b.s) | __ldg(cs + b.d) | __ldg(cd + b.s) | __ldg(cd + b.d)) != 0;
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (kFill) {
                if (ok) {
                    const int64_t p = off + __popc(m & ((1u << lane) - 1u));
                    reinterpret_cast<int2 *>(pairs)[p] = make_int2((int32_t)i, (int32_t)j);
                }
                off += __popc(m);
            } else {
                cnt += __popc(m);
                if (ok) hit[j] = 1;
            }
            if (bmask) break;
        }
        if (!kFill && lane == 0) {
            counts_or_offsets[i] = cnt;
            if (cnt) hit[i] = 1;
        }
    }
    if (!kFill && warp0 == 0 && lane == 0 && n > 0) counts_or_offsets[n - 1] = 0;
}

__global__ void k_sum_hits(const uint8_t *__restrict__ hit, int64_t n, unsigned long long *__restrict__ total) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    unsigned c = 0;
    for (; t < n; t += gridDim.x * (int64_t)blockDim.x) c += hit[t];
    for (int o = 16; o; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(total, (unsigned long long)c);
}

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// work layout: [offsets int64 (n+1)] [hit u8 n] [total u64] [counts int64 n] [keys, keys_alt u64 n]
// [idx, idx_alt i32 n] [cub temp]
struct WorkLayout {
    size_t offsets, hit, total, counts, keys, keys_alt, idx, idx_alt, temp, temp_bytes, bytes;
};

WorkLayout work_layout(int64_t n) {
    WorkLayout w;
    size_t sort_tmp = 0, scan_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (const uint64_t *)nullptr, (uint64_t *)nullptr,
                                    (const int32_t *)nullptr, (int32_t *)nullptr, (int64_t)n);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp, (const int64_t *)nullptr, (int64_t *)nullptr, (int64_t)(n + 1));
    size_t at = 0;
    auto take = [&](size_t bytes) {
        const size_t here = at;
        at += align256(bytes);
        return here;
    };
    const size_t m = (size_t)(n > 0 ? n : 1);
    w.offsets = take((m + 1) * 8);
    w.hit = take(m);
    w.total = take(8);
    w.counts = take((m + 1) * 8);
    w.keys = take(m * 8);
    w.keys_alt = take(m * 8);
    w.idx = take(m * 4);
    w.idx_alt = take(m * 4);
    w.temp_bytes = sort_tmp > scan_tmp ? sort_tmp : scan_tmp;
    w.temp = take(w.temp_bytes);
    w.bytes = at;
    return w;
}

}  // namespace

extern "C" int64_t gd_collective_work_bytes(int64_t n_jumps) {
    if (n_jumps < 0) return -1;
    return (int64_t)work_layout(n_jumps).bytes;
}

extern "C" gd_status_t gd_collective_sort(const int32_t *d_jumps, int64_t n_jumps, void *d_work,
                                          int32_t *d_sorted, void *stream) {
    GD_REQUIRE(n_jumps >= 0 && n_jumps < INT32_MAX, GD_ERR_INVALID, "gd_collective_sort: bad n_jumps");
    if (n_jumps == 0) return GD_OK;
    GD_REQUIRE(d_jumps && d_work && d_sorted, GD_ERR_INVALID, "gd_collective_sort: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const WorkLayout w = work_layout(n_jumps);
    char *base = (char *)d_work;
    uint64_t *keys = (uint64_t *)(base + w.keys), *keys_alt = (uint64_t *)(base + w.keys_alt);
    int32_t *idx = (int32_t *)(base + w.idx), *idx_alt = (int32_t *)(base + w.idx_alt);
    const int tb = 256;
    k_sort_keys<<<(unsigned)gd_ceil_div(n_jumps, tb), tb, 0, st>>>(d_jumps, n_jumps, keys, idx);
    GD_LAUNCH_CHECK();
    size_t temp_bytes = w.temp_bytes;
    // LSD radix sort: stable, so rows with equal (stop, start) keep their incoming order
    GD_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(base + w.temp, temp_bytes, keys, keys_alt, idx, idx_alt,
                                                  (int64_t)n_jumps, 0, 64, st));
    k_gather_rows<<<(unsigned)gd_ceil_div(n_jumps * 3, tb), tb, 0, st>>>(d_jumps, idx_alt, n_jumps, d_sorted);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_collective_close(const double *d_dist, int32_t n_sites, double max_dist,
                                           uint8_t *d_close, void *stream) {
    GD_REQUIRE(n_sites >= 0 && (int64_t)n_sites * n_sites < ((int64_t)1 << 34), GD_ERR_INVALID,
               "gd_collective_close: bad n_sites");
    if (n_sites == 0) return GD_OK;
    GD_REQUIRE(d_dist && d_close, GD_ERR_INVALID, "gd_collective_close: null pointer");
    const int64_t n = (int64_t)n_sites * n_sites;
    k_close<<<(unsigned)gd_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(d_dist, n, max_dist, d_close);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

static int collective_grid(int64_t n_jumps) {
    const GdDeviceInfo &dev = gd_device_info();
    const int64_t want = gd_ceil_div(n_jumps, 8);  // 8 warps per CTA
    const int64_t cap = (int64_t)(dev.sm_count > 0 ? dev.sm_count : 148) * 8;
    return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

extern "C" gd_status_t gd_collective_count(const int32_t *d_sorted, int64_t n_jumps, const uint8_t *d_close,
                                           int32_t n_sites, int32_t max_steps, void *d_work,
                                           int64_t *h_n_pairs, int64_t *h_n_solo, void *stream) {
    GD_REQUIRE(n_jumps >= 0 && n_jumps < INT32_MAX && n_sites >= 0, GD_ERR_INVALID, "gd_collective_count: bad size");
    GD_REQUIRE(h_n_pairs && h_n_solo, GD_ERR_INVALID, "gd_collective_count: null result pointer");
    *h_n_pairs = 0;
    *h_n_solo = n_jumps;
    if (n_jumps == 0) return GD_OK;
    GD_REQUIRE(d_sorted && d_close && d_work, GD_ERR_INVALID, "gd_collective_count: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const WorkLayout w = work_layout(n_jumps);
    char *base = (char *)d_work;
    int64_t *offsets = (int64_t *)(base + w.offsets), *counts = (int64_t *)(base + w.counts);
    uint8_t *hit = (uint8_t *)(base + w.hit);
    unsigned long long *total = (unsigned long long *)(base + w.total);
    GD_CUDA_CHECK(cudaMemsetAsync(base + w.hit, 0, (w.counts - w.hit) + (size_t)(n_jumps + 1) * 8, st));
    k_collective<false><<<collective_grid(n_jumps), 256, 0, st>>>(d_sorted, n_jumps, d_close, n_sites, max_steps,
                                                                  counts, nullptr, hit);
    GD_LAUNCH_CHECK();
    size_t temp_bytes = w.temp_bytes;
    GD_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(base + w.temp, temp_bytes, counts, offsets, (int64_t)(n_jumps + 1), st));
    k_sum_hits<<<collective_grid(n_jumps), 256, 0, st>>>(hit, n_jumps, total);
    GD_LAUNCH_CHECK();
    int64_t n_pairs = 0;
    unsigned long long n_hit = 0;
    GD_CUDA_CHECK(cudaMemcpyAsync(&n_pairs, offsets + n_jumps, 8, cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaMemcpyAsync(&n_hit, total, 8, cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));
    *h_n_pairs = n_pairs;
    *h_n_solo = n_jumps - (int64_t)n_hit;
    return GD_OK;
}

extern "C" gd_status_t gd_collective_fill(const int32_t *d_sorted, int64_t n_jumps, const uint8_t *d_close,
                                          int32_t n_sites, int32_t max_steps, const void *d_work,
                                          int32_t *d_pairs, void *stream) {
    GD_REQUIRE(n_jumps >= 0 && n_jumps < INT32_MAX && n_sites >= 0, GD_ERR_INVALID, "gd_collective_fill: bad size");
    if (n_jumps < 2) return GD_OK;
    GD_REQUIRE(d_sorted && d_close && d_work && d_pairs, GD_ERR_INVALID, "gd_collective_fill: null pointer");
    const WorkLayout w = work_layout(n_jumps);
    int64_t *offsets = (int64_t *)((char *)d_work + w.offsets);
    k_collective<true><<<collective_grid(n_jumps), 256, 0, (cudaStream_t)stream>>>(
        d_sorted, n_jumps, d_close, n_sites, max_steps, offsets, d_pairs, nullptr);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
