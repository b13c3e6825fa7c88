// K3 -- transition events -> jumps, plus the (start, destination) count matrix.
//
// Replaces _generic_transitions_to_jumps (reference src/gemdat/jumps.py:24-119: pandas
// groupby + DataFrame.iterrows, >= 25 us per event) and _calculate_transitions_matrix
// (src/gemdat/transitions.py:812-834: np.unique(axis=0) + fancy assignment).
//
// The conversion is a per-atom sequential state machine over that atom's events
// (already contiguous and time-ordered in the K2 table), so atoms run in parallel:
// one warp per atom streams the rows with coalesced 128-bit loads into shared
// memory and lane 0 steps the machine.  Two phases (count, fill) give the exact
// output size without any allocation in the library.
#include "gd_common.cuh"

namespace {

struct Jump {
    int32_t atom, start_site, dest_site, start_time, stop_time, residence;
};

// The state machine of jumps.py:45-105, on int32 rows
// (atom, start site, destination site, start inner, destination inner, time).
struct JumpMachine {
    bool have_from = false;
    int32_t from_start_site = 0, from_dest_site = 0, from_start_time = 0;
    bool have_cand = false;
    int32_t c_atom = 0, c_start_site = 0, c_dest_site = 0, c_dest_inner = 0, c_start_time = 0,
            c_stop_time = 0;
    int64_t minimal_residence;

    __device__ explicit JumpMachine(int64_t mr) : minimal_residence(mr) {}

    template <typename Emit>
    __device__ __forceinline__ void close_candidate(int64_t until, Emit &&emit) {
        const int64_t residence = until - ((int64_t)c_stop_time - 1);
        if (residence >= minimal_residence && c_dest_inner != GD_NOSITE &&
            c_start_site != c_dest_site) {  // jumps.py:113 start != destination
            emit(Jump{c_atom, c_start_site, c_dest_site, c_start_time, c_stop_time,
                      (int32_t)residence});
        }
        have_cand = false;
    }

    template <typename Emit>
    __device__ __forceinline__ void step(int32_t atom, int32_t s0, int32_t s1, int32_t i1,
                                         int32_t t, Emit &&emit) {
        if (have_cand) {
            if (s1 == c_dest_site) {
                if (i1 != GD_NOSITE) c_dest_inner = i1;
            } else {
                close_candidate(t, emit);
            }
        }
        if (s0 != GD_NOSITE && s0 != s1) {
            have_from = true;
            from_start_site = s0;
            from_dest_site = s1;
            from_start_time = t;
        }
        if (have_from) {
            if (s1 == from_start_site) {
                have_from = false;
                have_cand = false;
            } else if (i1 != GD_NOSITE || s1 != from_dest_site) {
                have_cand = true;
                c_atom = atom;
                c_start_site = from_start_site;
                c_dest_site = s1;
                c_dest_inner = i1;
                c_start_time = from_start_time;
                c_stop_time = t + 1;
                have_from = false;
            }
        }
    }
};

constexpr int kWarpsPerBlock = 4;
constexpr int kRowsPerBatch = 128;  // 128 rows * 24 B = 3 KiB per warp

template <bool kFill>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    k_jumps(const int32_t *__restrict__ events, const int64_t *__restrict__ atom_offsets,
            int32_t n_atoms, int64_t n_frames_total, int64_t minimal_residence,
            int64_t *__restrict__ counts, int32_t *__restrict__ jumps) {
    __shared__ int32_t rows[kWarpsPerBlock][kRowsPerBatch * 6];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int32_t a = blockIdx.x * kWarpsPerBlock + warp;
    if (a >= n_atoms) return;
    const int64_t r0 = atom_offsets[a], r1 = atom_offsets[a + 1];
    int64_t out = kFill ? counts[a] : 0;  // counts hold exclusive offsets in the fill phase
    int64_t n_out = 0;
    JumpMachine m(minimal_residence);
    auto emit = [&](const Jump &j) {
        if (kFill) {
            int32_t *o = jumps + (out + n_out) * 6;
            o[0] = j.atom;
            o[1] = j.start_site;
            o[2] = j.dest_site;
            o[3] = j.start_time;
            o[4] = j.stop_time;
            o[5] = j.residence;
        }
        ++n_out;
    };
    for (int64_t base = r0; base < r1; base += kRowsPerBatch) {
        const int n_rows = (int)min((int64_t)kRowsPerBatch, r1 - base);
        const int32_t *src = events + base * 6;
        for (int k = lane; k < n_rows * 6; k += 32) rows[warp][k] = __ldg(src + k);
        __syncwarp();
        if (lane == 0) {
            for (int r = 0; r < n_rows; ++r) {
                const int32_t *e = &rows[warp][r * 6];
                m.step(e[0], e[1], e[2], e[4], e[5], emit);
            }
        }
        __syncwarp();
    }
    if (lane == 0) {
        if (m.have_cand) m.close_candidate(n_frames_total, emit);  // jumps.py:97-105
        if (!kFill) counts[a] = n_out;
    }
}

// single block exclusive scan, counts[0..n) -> offsets, counts[n] = total
__global__ void __launch_bounds__(1024) k_scan_counts(int64_t *__restrict__ c, int32_t n) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t running_s;
    if (threadIdx.x == 0) running_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n; base += 1024) {
        const int idx = base + threadIdx.x;
        const int64_t v = idx < n ? c[idx] : 0;
        int64_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int64_t warp_base = 0;
        for (int w = 0; w < warp; ++w) warp_base += warp_tot[w];
        const int64_t running = running_s;
        __syncthreads();
        if (idx < n) c[idx] = running + warp_base + incl - v;
        if (threadIdx.x == 1023) running_s = running + warp_base + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) c[n] = running_s;
}

// ---------------------------------------------------------------- pair matrix

__global__ void __launch_bounds__(256)
    k_pair_count(const int32_t *__restrict__ rows, int64_t n_rows, int32_t stride, int32_t cs,
                 int32_t cd, int32_t n_sites, unsigned long long *__restrict__ scratch) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const int64_t w = (int64_t)n_sites + 1;
    for (; i < n_rows; i += step) {
        const int32_t s = rows[i * stride + cs], d = rows[i * stride + cd];
        if (s < -1 || s >= n_sites || d < -1 || d >= n_sites) continue;
        atomicAdd(scratch + (int64_t)(s + 1) * w + (d + 1), 1ULL);
    }
}

// numpy semantics of `m[start_idx, stop_idx] = counts` with unique rows sorted
// ascending: index -1 aliases n-1 and the lexicographically last source pair wins.
__global__ void __launch_bounds__(256)
    k_pair_finalize(const unsigned long long *__restrict__ scratch, int32_t n_sites,
                    long long *__restrict__ matrix) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_sites;
    if (idx >= n * n) return;
    const int64_t i = idx / n, j = idx - i * n;
    const int64_t w = n + 1;
    unsigned long long v = scratch[(i + 1) * w + (j + 1)];
    if (v == 0 && j == n - 1) v = scratch[(i + 1) * w + 0];
    if (v == 0 && i == n - 1) v = scratch[0 * w + (j + 1)];
    if (v == 0 && i == n - 1 && j == n - 1) v = scratch[0];
    matrix[idx] = (long long)v;
}

}  // namespace

extern "C" gd_status_t gd_jumps_count(const int32_t *d_events, const int64_t *d_atom_offsets,
                                      int32_t n_atoms, int64_t n_frames_total,
                                      int64_t minimal_residence, int64_t *d_jump_counts,
                                      int64_t *h_n_jumps, void *stream) {
    GD_REQUIRE(d_events && d_atom_offsets && d_jump_counts && h_n_jumps, GD_ERR_INVALID,
               "gd_jumps_count: null argument");
    GD_REQUIRE(n_atoms > 0, GD_ERR_INVALID, "gd_jumps_count: n_atoms must be positive");
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)gd_ceil_div(n_atoms, kWarpsPerBlock);
    k_jumps<false><<<blocks, kWarpsPerBlock * 32, 0, st>>>(d_events, d_atom_offsets, n_atoms,
                                                          n_frames_total, minimal_residence,
                                                          d_jump_counts, nullptr);
    GD_LAUNCH_CHECK();
    k_scan_counts<<<1, 1024, 0, st>>>(d_jump_counts, n_atoms);
    GD_LAUNCH_CHECK();
    GD_CUDA_CHECK(cudaMemcpyAsync(h_n_jumps, d_jump_counts + n_atoms, sizeof(int64_t),
                                  cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));
    return GD_OK;
}

extern "C" gd_status_t gd_jumps_fill(const int32_t *d_events, const int64_t *d_atom_offsets,
                                     int32_t n_atoms, int64_t n_frames_total,
                                     int64_t minimal_residence, const int64_t *d_jump_counts,
                                     int32_t *d_jumps, void *stream) {
    GD_REQUIRE(d_events && d_atom_offsets && d_jump_counts && d_jumps, GD_ERR_INVALID,
               "gd_jumps_fill: null argument");
    GD_REQUIRE(n_atoms > 0, GD_ERR_INVALID, "gd_jumps_fill: n_atoms must be positive");
    const unsigned blocks = (unsigned)gd_ceil_div(n_atoms, kWarpsPerBlock);
    k_jumps<true><<<blocks, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        d_events, d_atom_offsets, n_atoms, n_frames_total, minimal_residence,
        const_cast<int64_t *>(d_jump_counts), d_jumps);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_pair_matrix(const int32_t *d_rows, int64_t n_rows, int32_t row_stride,
                                      int32_t col_start, int32_t col_dest, int32_t n_sites,
                                      unsigned long long *d_scratch, long long *d_matrix,
                                      void *stream) {
    GD_REQUIRE(d_scratch && d_matrix && n_sites > 0 && n_rows >= 0, GD_ERR_INVALID,
               "gd_pair_matrix: bad argument");
    GD_REQUIRE(n_rows == 0 || d_rows, GD_ERR_INVALID, "gd_pair_matrix: null rows");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows > 0) {
        int64_t blocks = gd_ceil_div(n_rows, 256);
        if (blocks > 4096) blocks = 4096;
        k_pair_count<<<(unsigned)blocks, 256, 0, st>>>(d_rows, n_rows, row_stride, col_start,
                                                      col_dest, n_sites, d_scratch);
        GD_LAUNCH_CHECK();
    }
    const int64_t cells = (int64_t)n_sites * n_sites;
    k_pair_finalize<<<(unsigned)gd_ceil_div(cells, 256), 256, 0, st>>>(d_scratch, n_sites, d_matrix);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
