// N1 -- reductions over the state arrays and the event table.
//
// Replaces, for the device-resident tables of K1/K2:
//   Transitions.occupancy            (reference src/gemdat/transitions.py:280-306:
//                                     np.unique(states, return_counts=True) over [T, A])
//   Transitions.residence_time       (:339-373: Python loop over atoms, np.diff/np.nonzero
//                                     run-length encoding of every column)
//   Transitions.split / _split_transitions_events (:375-410, :837-879: boolean masks
//                                     over the event DataFrame per part)
//
// occupancy is a histogram of the int32 states over n_sites + 1 bins (4 B read per
// atom-frame, HBM bound): per-CTA shared-memory bins, NOSITE counted in registers (many
// atoms of a warp are in transit at the same time, every other bin is hit by one atom at a
// time because a site holds one atom).  residence_time is a function of the OUTER-change
// rows of the event table alone -- a boundary of the reference's np.diff is exactly such a
// row -- so it costs O(events), not O(frames): one thread per row looks up the next outer
// change of the same atom; rows come out in (atom, time) order by count -> scan -> fill.
// split uses that an atom's rows are sorted by time: the rows of one part are a contiguous
// sub-range per atom, found by binary search.
#include "gd_common.cuh"

namespace {

// ---------------------------------------------------------------- occupancy

constexpr int kOccThreads = 256;
constexpr int kOccMaxSmemBins = 48 * 1024;  // 192 KB of uint32

__device__ __forceinline__ void occ_add(int32_t s, int32_t n_sites, unsigned int *hist,
                                        unsigned long long *counts, unsigned int &nosite) {
    if (s == GD_NOSITE) {
        ++nosite;
    } else if ((uint32_t)s < (uint32_t)n_sites) {
        if (hist)
            atomicAdd(hist + s + 1, 1u);
        else
            atomicAdd(counts + s + 1, 1ull);
    }
}

__global__ void __launch_bounds__(kOccThreads)
    k_occupancy(const int32_t *__restrict__ states, int64_t n, int32_t n_sites, int use_smem,
                int aligned16, unsigned long long *__restrict__ counts) {
    extern __shared__ unsigned int s_hist[];
    unsigned int *hist = use_smem ? s_hist : nullptr;
    if (hist) {
        for (int i = threadIdx.x; i <= n_sites; i += kOccThreads) hist[i] = 0u;
        __syncthreads();
    }
    unsigned int nosite = 0;
    const int64_t gtid = (int64_t)blockIdx.x * kOccThreads + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * kOccThreads;
    int64_t done = 0;
    if (aligned16) {
        const int64_t n4 = n >> 2;
        const int4 *s4 = reinterpret_cast<const int4 *>(states);
        for (int64_t i = gtid; i < n4; i += stride) {
            const int4 v = __ldcs(s4 + i);
            occ_add(v.x, n_sites, hist, counts, nosite);
            occ_add(v.y, n_sites, hist, counts, nosite);
            occ_add(v.z, n_sites, hist, counts, nosite);
            occ_add(v.w, n_sites, hist, counts, nosite);
        }
        done = n4 << 2;
    }
    for (int64_t i = done + gtid; i < n; i += stride)
        occ_add(__ldcs(states + i), n_sites, hist, counts, nosite);
    nosite = (unsigned int)gd_warp_sum((int)nosite);
    if ((threadIdx.x & 31) == 0 && nosite) {
        if (hist)
            atomicAdd(hist, nosite);
        else
            atomicAdd(counts, (unsigned long long)nosite);
    }
    if (hist) {
        __syncthreads();
        for (int i = threadIdx.x; i <= n_sites; i += kOccThreads) {
            const unsigned int v = hist[i];
            if (v) atomicAdd(counts + i, (unsigned long long)v);
        }
    }
}

// ---------------------------------------------------------------- residence time

constexpr int kResThreads = 256;

// Row i is an outer change when start site != destination site.  It opens a visit of its
// destination site that ends at the next outer change of the same atom; the visit is
// reported when that next change exists (the last visit of an atom is dropped,
// transitions.py:362) and the site is not NOSITE (:365-366).  The visit before an atom's
// first change is never reported either (:362, boundaries[:-1] / boundaries[1:]).
__device__ __forceinline__ bool residence_of(const int32_t *__restrict__ ev, int64_t n_events,
                                             int64_t i, int32_t &atom, int32_t &site,
                                             int32_t &dt) {
    const int32_t *r = ev + i * 6;
    atom = r[0];
    site = r[2];
    if (r[1] == site || site == GD_NOSITE) return false;
    const int32_t t0 = r[5];
    for (int64_t j = i + 1; j < n_events; ++j) {
        const int32_t *q = ev + j * 6;
        if (q[0] != atom) return false;
        if (q[1] != q[2]) {
            dt = q[5] - t0;
            return true;
        }
    }
    return false;
}

// block-wide exclusive scan of one flag per thread; returns the thread's offset, total in *total
__device__ __forceinline__ int block_excl_scan(int v, int *total) {
    __shared__ int warp_tot[kResThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int base = 0, all = 0;
#pragma unroll
    for (int w = 0; w < kResThreads / 32; ++w) {
        const int t = warp_tot[w];
        if (w < warp) base += t;
        all += t;
    }
    __syncthreads();
    *total = all;
    return base + incl - v;
}

template <bool kFill>
__global__ void __launch_bounds__(kResThreads)
    k_residence(const int32_t *__restrict__ ev, int64_t n_events, int64_t *__restrict__ block_off,
                int32_t *__restrict__ rows) {
    const int64_t i = (int64_t)blockIdx.x * kResThreads + threadIdx.x;
    int32_t atom = 0, site = 0, dt = 0;
    const bool emit = i < n_events && residence_of(ev, n_events, i, atom, site, dt);
    int total;
    const int off = block_excl_scan(emit ? 1 : 0, &total);
    if (!kFill) {
        if (threadIdx.x == 0) block_off[blockIdx.x + 1] = total;
    } else if (emit) {
        int32_t *o = rows + (block_off[blockIdx.x] + off) * 3;
        o[0] = atom;
        o[1] = site;
        o[2] = dt;
    }
}

// ---------------------------------------------------------------- split

// first row of the atom's block [r0, r1) whose time is >= bound
__device__ __forceinline__ int64_t first_at_or_after(const int32_t *__restrict__ ev, int64_t r0,
                                                     int64_t r1, int64_t bound) {
    while (r0 < r1) {
        const int64_t mid = (r0 + r1) >> 1;
        if ((int64_t)ev[mid * 6 + 5] < bound)
            r0 = mid + 1;
        else
            r1 = mid;
    }
    return r0;
}

// thread per (part, atom): the part's rows of the atom are [lo, hi); counts go to
// part_off[p, a + 1] (scanned per part afterwards), lo to src_lo[p, a]
__global__ void __launch_bounds__(256)
    k_split_count(const int32_t *__restrict__ ev, const int64_t *__restrict__ atom_off,
                  int32_t n_atoms, const int64_t *__restrict__ bins, int32_t n_parts,
                  int64_t *__restrict__ part_off, int64_t *__restrict__ src_lo) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= (int64_t)n_parts * n_atoms) return;
    const int32_t p = (int32_t)(k / n_atoms), a = (int32_t)(k - (int64_t)p * n_atoms);
    const int64_t r0 = atom_off[a], r1 = atom_off[a + 1];
    const int64_t lo = first_at_or_after(ev, r0, r1, bins[p]);
    const int64_t hi = first_at_or_after(ev, lo, r1, bins[p + 1]);
    src_lo[k] = lo;
    int64_t *row = part_off + (int64_t)p * (n_atoms + 1);
    row[a + 1] = hi - lo;
}

// one block per part: exclusive scan over atoms of its row of part_off
__global__ void __launch_bounds__(1024)
    k_split_scan(int64_t *__restrict__ part_off, int32_t n_atoms) {
    int64_t *v = part_off + (int64_t)blockIdx.x * (n_atoms + 1);
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t running_s;
    if (threadIdx.x == 0) {
        running_s = 0;
        v[0] = 0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n_atoms; base += 1024) {
        const int idx = base + threadIdx.x;
        const int64_t x = idx < n_atoms ? v[idx + 1] : 0;
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
        if (idx < n_atoms) v[idx + 1] = running + warp_base + incl;
        if (threadIdx.x == 1023) running_s = running + warp_base + incl;
        __syncthreads();
    }
}

// block per (atom, part): copy the rows, shifting the time column to the part's first frame;
// the source row numbers are the index labels the reference's boolean mask keeps
__global__ void __launch_bounds__(128)
    k_split_fill(const int32_t *__restrict__ ev, int32_t n_atoms, const int64_t *__restrict__ bins,
                 const int64_t *__restrict__ part_off, const int64_t *__restrict__ src_lo,
                 const int64_t *__restrict__ part_base, int32_t *__restrict__ rows,
                 int64_t *__restrict__ index) {
    const int32_t a = blockIdx.x, p = blockIdx.y;
    const int64_t *po = part_off + (int64_t)p * (n_atoms + 1);
    const int64_t n = po[a + 1] - po[a];
    if (n <= 0) return;
    const int64_t lo = src_lo[(int64_t)p * n_atoms + a];
    const int64_t dst = part_base[p] + po[a];
    const int32_t shift = (int32_t)bins[p];
    for (int64_t k = threadIdx.x; k < n * 6; k += blockDim.x) {
        int32_t v = ev[lo * 6 + k];
        if (k % 6 == 5) v -= shift;
        rows[dst * 6 + k] = v;
    }
    if (index)
        for (int64_t k = threadIdx.x; k < n; k += blockDim.x) index[dst + k] = lo + k;
}

}  // namespace

extern "C" gd_status_t gd_occupancy(const int32_t *d_states, int64_t n, int32_t n_sites,
                                    unsigned long long *d_counts, void *stream) {
    GD_REQUIRE(n >= 0 && n_sites > 0 && d_counts && (n == 0 || d_states), GD_ERR_INVALID,
               "gd_occupancy: bad argument");
    if (n == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    const int use_smem = n_sites + 1 <= kOccMaxSmemBins;
    const size_t smem = use_smem ? (size_t)(n_sites + 1) * sizeof(unsigned int) : 0;
    if (smem > 48 * 1024)
        GD_CUDA_CHECK(cudaFuncSetAttribute(k_occupancy, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem));
    // resident CTAs per SM by shared memory; every CTA's uint32 bins hold < 2^31 elements
    int per_sm = use_smem ? (int)((size_t)200 * 1024 / (smem + 1024)) : 8;
    per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
    int64_t blocks = (int64_t)dev.sm_count * per_sm;
    const int64_t need = gd_ceil_div(gd_ceil_div(n, 4), kOccThreads);
    if (blocks > need) blocks = need;
    const int64_t min_blocks = gd_ceil_div(n, (int64_t)1 << 30);
    if (blocks < min_blocks) blocks = min_blocks;
    const int aligned16 = (reinterpret_cast<uintptr_t>(d_states) & 15) == 0;
    k_occupancy<<<(unsigned)blocks, kOccThreads, smem, (cudaStream_t)stream>>>(
        d_states, n, n_sites, use_smem, aligned16, d_counts);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" int64_t gd_residence_work_bytes(int64_t n_events) {
    if (n_events < 0) n_events = 0;
    return (gd_ceil_div(n_events, kResThreads) + 2) * (int64_t)sizeof(int64_t);
}

extern "C" gd_status_t gd_residence_count(const int32_t *d_events, int64_t n_events, void *d_work,
                                          int64_t *h_n_rows, void *stream) {
    GD_REQUIRE(n_events >= 0 && d_work && h_n_rows && (n_events == 0 || d_events), GD_ERR_INVALID,
               "gd_residence_count: bad argument");
    *h_n_rows = 0;
    if (n_events == 0) return GD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t *block_off = static_cast<int64_t *>(d_work);
    const int64_t blocks = gd_ceil_div(n_events, kResThreads);
    GD_REQUIRE(blocks < ((int64_t)1 << 31), GD_ERR_INVALID, "gd_residence_count: too many events");
    k_residence<false><<<(unsigned)blocks, kResThreads, 0, st>>>(d_events, n_events, block_off,
                                                                 nullptr);
    GD_LAUNCH_CHECK();
    gd_k_scan_offsets<<<1, 1024, 0, st>>>(block_off, blocks);
    GD_LAUNCH_CHECK();
    GD_CUDA_CHECK(cudaMemcpyAsync(h_n_rows, block_off + blocks, sizeof(int64_t),
                                  cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));
    return GD_OK;
}

extern "C" gd_status_t gd_residence_fill(const int32_t *d_events, int64_t n_events,
                                         const void *d_work, int32_t *d_rows, void *stream) {
    GD_REQUIRE(n_events >= 0 && d_work && (n_events == 0 || (d_events && d_rows)), GD_ERR_INVALID,
               "gd_residence_fill: bad argument");
    if (n_events == 0) return GD_OK;
    const int64_t blocks = gd_ceil_div(n_events, kResThreads);
    k_residence<true><<<(unsigned)blocks, kResThreads, 0, (cudaStream_t)stream>>>(
        d_events, n_events, const_cast<int64_t *>(static_cast<const int64_t *>(d_work)), d_rows);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_split_count(const int32_t *d_events,
                                             const int64_t *d_atom_offsets, int32_t n_atoms,
                                             const int64_t *d_bins, int32_t n_parts,
                                             int64_t *d_part_offsets, int64_t *d_src_lo,
                                             void *stream) {
    GD_REQUIRE(d_atom_offsets && d_bins && d_part_offsets && d_src_lo && n_atoms > 0 && n_parts > 0,
               GD_ERR_INVALID, "gd_events_split_count: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t items = (int64_t)n_parts * n_atoms;
    k_split_count<<<(unsigned)gd_ceil_div(items, 256), 256, 0, st>>>(
        d_events, d_atom_offsets, n_atoms, d_bins, n_parts, d_part_offsets, d_src_lo);
    GD_LAUNCH_CHECK();
    k_split_scan<<<(unsigned)n_parts, 1024, 0, st>>>(d_part_offsets, n_atoms);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_split_fill(const int32_t *d_events, int32_t n_atoms,
                                            const int64_t *d_bins, int32_t n_parts,
                                            const int64_t *d_part_offsets, const int64_t *d_src_lo,
                                            const int64_t *d_part_base, int32_t *d_rows,
                                            int64_t *d_index, void *stream) {
    GD_REQUIRE(d_events && d_bins && d_part_offsets && d_src_lo && d_part_base && d_rows &&
                   n_atoms > 0 && n_parts > 0 && n_parts <= 65535,
               GD_ERR_INVALID, "gd_events_split_fill: bad argument");
    dim3 grid((unsigned)n_atoms, (unsigned)n_parts);
    k_split_fill<<<grid, 128, 0, (cudaStream_t)stream>>>(d_events, n_atoms, d_bins, d_part_offsets,
                                                         d_src_lo, d_part_base, d_rows, d_index);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
