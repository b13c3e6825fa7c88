// Per-frame averages over atoms and the drift correction built on them: the callers on the
// input side of the path (SURVEY 8a rows a1-a3 feed on their output).
//
// Replaces
//   Trajectory.center_of_mass           (reference src/gemdat/trajectory.py:732-749:
//                                        np.average(base + cumulative_displacements, axis=1, weights=masses))
//   Trajectory.drift                    (:751-791: np.mean(displacements of the fixed species, axis=1))
//   Trajectory.apply_drift_correction   (:793-830: coords = displacements - drift in displacement
//                                        form; positions are then base + np.cumsum(coords) mod 1)
//
// All three are one pass over [T, N, 3] float64 with a reduction over the atoms of a frame (a
// frame's atoms are contiguous: one CTA per frame, coalesced) or an elementwise pass: HBM bound.
//   k_frame_mean     24 B read per atom-frame                    -> [T, 3]
//   k_drift          24 B read per selected atom-frame (the row before comes from L2) -> [T, 3]
//   k_cumsum3_*      the [T, 3] drift summed along T (block scans + carries, three small launches)
//   k_drift_correct  24 B read + 24 B write per atom-frame: mod(base + cum - D, 1)
// The reference accumulates (d_t - drift_t) per atom sequentially in float64; here the cumulative
// displacement comes from K0 (x_t - x_0 - integer winding, exact up to one rounding) and the
// cumulative drift is subtracted from it: equal up to a few ulp of the running sum (the tests
// state the tolerance and compare modulo 1).
#include "gd_common.cuh"

namespace {

constexpr int kThreads = 128;

// fixed-order block reduction of three values (deterministic)
__device__ __forceinline__ void block_sum3(double &x, double &y, double &z) {
    __shared__ double s[3][kThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x += __shfl_xor_sync(0xffffffffu, x, o);
        y += __shfl_xor_sync(0xffffffffu, y, o);
        z += __shfl_xor_sync(0xffffffffu, z, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        s[0][warp] = x;
        s[1][warp] = y;
        s[2][warp] = z;
    }
    __syncthreads();
    x = y = z = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        x += s[0][w];
        y += s[1][w];
        z += s[2][w];
    }
    __syncthreads();
}

// out[t, c] = sum_a w[a] (base[a, c] + x[t, a, c]) / sum_a w[a]     (w == NULL: plain mean)
__global__ void __launch_bounds__(kThreads)
    k_frame_mean(const double *__restrict__ x, const double *__restrict__ base,
                 const double *__restrict__ w, int64_t n_frames, int32_t n_atoms,
                 double *__restrict__ out) {
    double wsum = 0.0, dummy0 = 0.0, dummy1 = 0.0;
    for (int a = threadIdx.x; a < n_atoms; a += kThreads) wsum += w ? w[a] : 1.0;
    block_sum3(wsum, dummy0, dummy1);
    for (int64_t t = blockIdx.x; t < n_frames; t += gridDim.x) {
        const double *row = x + t * (int64_t)n_atoms * 3;
        double sx = 0.0, sy = 0.0, sz = 0.0;
        for (int a = threadIdx.x; a < n_atoms; a += kThreads) {
            const double wa = w ? w[a] : 1.0;
            double px = __ldcs(row + 3 * a), py = __ldcs(row + 3 * a + 1), pz = __ldcs(row + 3 * a + 2);
            if (base) {
                px += base[3 * a];
                py += base[3 * a + 1];
                pz += base[3 * a + 2];
            }
            sx += wa * px;
            sy += wa * py;
            sz += wa * pz;
        }
        block_sum3(sx, sy, sz);
        if (threadIdx.x == 0) {
            out[3 * t] = sx / wsum;
            out[3 * t + 1] = sy / wsum;
            out[3 * t + 2] = sz / wsum;
        }
    }
}

__device__ __forceinline__ double wrapped_step(double now, double before) {
    const double d = now - before;
    return d - rint(d);  // np.around: half to even
}

// out[t, c] = mean over the selected atoms of the wrapped step x[t] - x[t-1]; the step of frame 0
// is 0 (pymatgen to_displacements) unless the caller passes the row before the shard
__global__ void __launch_bounds__(kThreads)
    k_drift(const double *__restrict__ pos, int64_t n_frames, int32_t n_atoms,
            const int32_t *__restrict__ sel, int32_t n_sel, const double *__restrict__ prev_row,
            double *__restrict__ out) {
    const int64_t stride = (int64_t)n_atoms * 3;
    for (int64_t t = blockIdx.x; t < n_frames; t += gridDim.x) {
        const double *row = pos + t * stride;
        const double *before = t > 0 ? row - stride : prev_row;
        double sx = 0.0, sy = 0.0, sz = 0.0;
        if (before != nullptr) {
            for (int k = threadIdx.x; k < n_sel; k += kThreads) {
                const int a = sel ? sel[k] : k;
                sx += wrapped_step(row[3 * a], before[3 * a]);
                sy += wrapped_step(row[3 * a + 1], before[3 * a + 1]);
                sz += wrapped_step(row[3 * a + 2], before[3 * a + 2]);
            }
        }
        block_sum3(sx, sy, sz);
        if (threadIdx.x == 0) {
            out[3 * t] = sx / (double)n_sel;
            out[3 * t + 1] = sy / (double)n_sel;
            out[3 * t + 2] = sz / (double)n_sel;
        }
    }
}

// Inclusive sum of v[T, 3] along T in three small launches (all loads coalesced):
//   k_cumsum3_local  every CTA scans its own block of 1024 frames into `out`, block total -> totals
//   k_cumsum3_carry  one CTA turns the block totals into carries (exclusive scan, tile by tile)
//   k_cumsum3_add    every CTA adds its carry
constexpr int kScanBlock = 1024;

// inclusive scan of one value per thread over the CTA (1024 threads), three channels at once;
// returns the CTA total in tot[]
__device__ __forceinline__ void block_scan3(double incl[3], double tot[3]) {
    __shared__ double warp_tot[3][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double up = __shfl_up_sync(0xffffffffu, incl[c], o);
            if (lane >= o) incl[c] += up;
        }
    }
    if (lane == 31) {
#pragma unroll
        for (int c = 0; c < 3; ++c) warp_tot[c][warp] = incl[c];
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        double before = 0.0, all = 0.0;
        for (int w = 0; w < 32; ++w) {
            const double x = warp_tot[c][w];
            if (w < warp) before += x;
            all += x;
        }
        incl[c] += before;
        tot[c] = all;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kScanBlock)
    k_cumsum3_local(const double *__restrict__ v, int64_t n_frames, double *__restrict__ out,
                    double *__restrict__ totals) {
    const int64_t t = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
    double incl[3], tot[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) incl[c] = t < n_frames ? v[3 * t + c] : 0.0;
    block_scan3(incl, tot);
    if (t < n_frames) {
#pragma unroll
        for (int c = 0; c < 3; ++c) out[3 * t + c] = incl[c];
    }
    if (threadIdx.x == 0) {
#pragma unroll
        for (int c = 0; c < 3; ++c) totals[3 * (int64_t)blockIdx.x + c] = tot[c];
    }
}

// totals[nb, 3] -> exclusive sums in place (one CTA, tiles of 1024 blocks with a running carry)
__global__ void __launch_bounds__(kScanBlock) k_cumsum3_carry(double *__restrict__ totals, int64_t nb) {
    __shared__ double carry[3];
    __shared__ double s_incl[3][kScanBlock];
    if (threadIdx.x < 3) carry[threadIdx.x] = 0.0;
    __syncthreads();
    for (int64_t base = 0; base < nb; base += kScanBlock) {
        const int64_t b = base + threadIdx.x;
        double incl[3], tot[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) incl[c] = b < nb ? totals[3 * b + c] : 0.0;
        block_scan3(incl, tot);
#pragma unroll
        for (int c = 0; c < 3; ++c) s_incl[c][threadIdx.x] = incl[c];
        __syncthreads();
        if (b < nb) {
#pragma unroll
            for (int c = 0; c < 3; ++c)  // exclusive value: the inclusive one of the block before
                totals[3 * b + c] = threadIdx.x == 0 ? carry[c] : carry[c] + s_incl[c][threadIdx.x - 1];
        }
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) carry[c] += tot[c];
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kScanBlock)
    k_cumsum3_add(double *__restrict__ out, int64_t n_frames, const double *__restrict__ carries) {
    if (blockIdx.x == 0) return;  // carry 0
    const int64_t t = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
    if (t >= n_frames) return;
#pragma unroll
    for (int c = 0; c < 3; ++c) out[3 * t + c] = carries[3 * (int64_t)blockIdx.x + c] + out[3 * t + c];
}

// out[t, a, c] = np.mod(base[a, c] + cum[t, a, c] - D[t, c], 1)      (out may alias cum)
__global__ void __launch_bounds__(256)
    k_drift_correct(const double *cum, const double *__restrict__ base,
                    const double *__restrict__ shift, int64_t n_frames, int32_t n_atoms, double *out) {
    const int row = n_atoms * 3;
    constexpr int kAhead = 4;  // loads in flight per thread
    for (int64_t t = blockIdx.x; t < n_frames; t += gridDim.x) {
        const double s0 = shift[3 * t], s1 = shift[3 * t + 1], s2 = shift[3 * t + 2];
        const double *in = cum + t * (int64_t)row;
        double *o = out + t * (int64_t)row;
        for (int k0 = threadIdx.x; k0 < row; k0 += 256 * kAhead) {
            double v[kAhead];
#pragma unroll
            for (int u = 0; u < kAhead; ++u) {
                const int k = k0 + 256 * u;
                v[u] = k < row ? __ldcs(in + k) : 0.0;
            }
#pragma unroll
            for (int u = 0; u < kAhead; ++u) {
                const int k = k0 + 256 * u;
                if (k < row) {
                    const int c = k % 3;
                    const double x = base[k] + (v[u] - (c == 0 ? s0 : (c == 1 ? s1 : s2)));
                    __stcs(o + k, x - floor(x));
                }
            }
        }
    }
}

inline unsigned frame_grid(int64_t n_frames) {
    const GdDeviceInfo &dev = gd_device_info();
    const int64_t cap = (int64_t)(dev.sm_count > 0 ? dev.sm_count : 148) * 16;
    return (unsigned)(n_frames < cap ? n_frames : cap);
}

}  // namespace

extern "C" gd_status_t gd_frame_mean(const double *d_x, const double *d_base, const double *d_weights,
                                     int64_t n_frames, int32_t n_atoms, double *d_out, void *stream) {
    GD_REQUIRE(d_x && d_out && n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_frame_mean: bad argument");
    k_frame_mean<<<frame_grid(n_frames), kThreads, 0, (cudaStream_t)stream>>>(d_x, d_base, d_weights,
                                                                            n_frames, n_atoms, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_drift(const double *d_pos, int64_t n_frames, int32_t n_atoms,
                                const int32_t *d_selected, int32_t n_selected,
                                const double *d_prev_row, double *d_drift, void *stream) {
    GD_REQUIRE(d_pos && d_drift && n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_drift: bad argument");
    GD_REQUIRE(n_selected > 0 && (d_selected || n_selected == n_atoms), GD_ERR_INVALID,
               "gd_drift: the selection is empty (or NULL with n_selected != n_atoms)");
    k_drift<<<frame_grid(n_frames), kThreads, 0, (cudaStream_t)stream>>>(
        d_pos, n_frames, n_atoms, d_selected, n_selected, d_prev_row, d_drift);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" int64_t gd_cumsum_frames3_work_bytes(int64_t n_frames) {
    return (n_frames > 0 ? gd_ceil_div(n_frames, (int64_t)kScanBlock) : 1) * 3 * (int64_t)sizeof(double);
}

extern "C" gd_status_t gd_cumsum_frames3(const double *d_in, int64_t n_frames, void *d_work,
                                         double *d_out, void *stream) {
    GD_REQUIRE(d_in && d_out && d_work && n_frames > 0, GD_ERR_INVALID, "gd_cumsum_frames3: bad argument");
    GD_REQUIRE(d_in != d_out, GD_ERR_INVALID, "gd_cumsum_frames3: d_out must not alias d_in");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nb = gd_ceil_div(n_frames, (int64_t)kScanBlock);
    double *totals = static_cast<double *>(d_work);
    k_cumsum3_local<<<(unsigned)nb, kScanBlock, 0, st>>>(d_in, n_frames, d_out, totals);
    GD_LAUNCH_CHECK();
    if (nb > 1) {
        k_cumsum3_carry<<<1, kScanBlock, 0, st>>>(totals, nb);
        GD_LAUNCH_CHECK();
        k_cumsum3_add<<<(unsigned)nb, kScanBlock, 0, st>>>(d_out, n_frames, totals);
        GD_LAUNCH_CHECK();
    }
    return GD_OK;
}

extern "C" gd_status_t gd_drift_correct(const double *d_cumulative, const double *d_base,
                                        const double *d_cumulative_drift, int64_t n_frames,
                                        int32_t n_atoms, double *d_out, void *stream) {
    GD_REQUIRE(d_cumulative && d_base && d_cumulative_drift && d_out && n_frames > 0 && n_atoms > 0,
               GD_ERR_INVALID, "gd_drift_correct: bad argument");
    GD_REQUIRE((int64_t)n_atoms * 3 < INT32_MAX, GD_ERR_INVALID, "gd_drift_correct: too many atoms");
    k_drift_correct<<<frame_grid(n_frames), 256, 0, (cudaStream_t)stream>>>(
        d_cumulative, d_base, d_cumulative_drift, n_frames, n_atoms, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
