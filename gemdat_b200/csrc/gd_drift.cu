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
//   k_cumsum3        the [T, 3] drift summed along T (one CTA, carried tile by tile)
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

// one CTA: inclusive sum of v[T, 3] along T, tile by tile (1024 frames) with a running carry
__global__ void __launch_bounds__(1024) k_cumsum3(const double *__restrict__ v, int64_t n_frames,
                                                  double *__restrict__ out) {
    __shared__ double warp_tot[3][32];
    __shared__ double carry[3];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 3) carry[threadIdx.x] = 0.0;
    __syncthreads();
    for (int64_t base = 0; base < n_frames; base += 1024) {
        const int64_t t = base + threadIdx.x;
        double incl[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) incl[c] = t < n_frames ? v[3 * t + c] : 0.0;
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
        double before[3] = {carry[0], carry[1], carry[2]};
        for (int w = 0; w < warp; ++w) {
#pragma unroll
            for (int c = 0; c < 3; ++c) before[c] += warp_tot[c][w];
        }
        if (t < n_frames) {
#pragma unroll
            for (int c = 0; c < 3; ++c) out[3 * t + c] = before[c] + incl[c];
        }
        __syncthreads();
        if (threadIdx.x == 1023) {
#pragma unroll
            for (int c = 0; c < 3; ++c) carry[c] = before[c] + incl[c];
        }
        __syncthreads();
    }
}

// out[t, a, c] = np.mod(base[a, c] + cum[t, a, c] - D[t, c], 1)      (out may alias cum)
__global__ void __launch_bounds__(256)
    k_drift_correct(const double *cum, const double *__restrict__ base,
                    const double *__restrict__ shift, int64_t n_frames, int32_t n_atoms, double *out) {
    const int64_t row = (int64_t)n_atoms * 3;
    const int64_t n = n_frames * row;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int64_t t = i / row;
        const int64_t k = i - t * row;  // 3 a + c
        const int c = (int)(k % 3);
        const double x = base[k] + (cum[i] - shift[3 * t + c]);
        out[i] = x - floor(x);
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

extern "C" gd_status_t gd_cumsum_frames3(const double *d_in, int64_t n_frames, double *d_out,
                                         void *stream) {
    GD_REQUIRE(d_in && d_out && n_frames > 0, GD_ERR_INVALID, "gd_cumsum_frames3: bad argument");
    k_cumsum3<<<1, 1024, 0, (cudaStream_t)stream>>>(d_in, n_frames, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_drift_correct(const double *d_cumulative, const double *d_base,
                                        const double *d_cumulative_drift, int64_t n_frames,
                                        int32_t n_atoms, double *d_out, void *stream) {
    GD_REQUIRE(d_cumulative && d_base && d_cumulative_drift && d_out && n_frames > 0 && n_atoms > 0,
               GD_ERR_INVALID, "gd_drift_correct: bad argument");
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    const int64_t n = n_frames * n_atoms * 3;
    int64_t blocks = gd_ceil_div(n, 256);
    if (blocks > (int64_t)dev.sm_count * 16) blocks = (int64_t)dev.sm_count * 16;
    k_drift_correct<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        d_cumulative, d_base, d_cumulative_drift, n_frames, n_atoms, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
