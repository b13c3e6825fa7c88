// K4 -- all-lag mean squared displacement by direct summation.
//
// Replaces Trajectory.mean_squared_displacement (reference
// src/gemdat/trajectory.py:896-933), which builds S1 - 2*S2 from a length-2T complex
// numpy FFT per atom and component.  Here msd[a,k] = mean_t |r[a,t+k]-r[a,t]|^2 is
// summed as written: more accurate than the FFT form at small lags (no
// cancellation) and fp64-pipe bound instead of memory bound.
//
// Layout: positions are first transposed to SoA [atom][component][frame] so both the
// r[t] tile and the r[t+k] window stream in with unit stride.  One CTA owns
// (atom, block of 256*8 consecutive lags); a thread owns 8 consecutive lags and
// slides an 8-frame register window over the r[t+k] stream: per frame step it
// loads ONE new window element and ONE broadcast r[t] (6 LDS.64) for 8 pair
// terms (24 DADD + 24 DFMA), which keeps the kernel on the fp64 pipe rather than on
// shared-memory bandwidth.  The window is stored with one pad slot per 8 so the
// 8-double thread stride is bank-conflict free.
//
// Algorithmic work: 6 fp64 instructions (3 DADD + 3 DFMA) per (atom, t, lag) pair;
// roofline = fp64 pipe.
#include "gd_common.cuh"

namespace {

constexpr int kM = 8;             // lags per thread
constexpr int kMsdThreads = 256;  // threads per CTA
constexpr int kLagsPerBlock = kM * kMsdThreads;  // 2048
constexpr int kBT = 256;          // frames of r[t] per tile
constexpr int kWin = kBT + kLagsPerBlock;  // r[t+k] window elements per tile
__host__ __device__ constexpr int padded(int j) { return j + (j >> 3); }
constexpr int kWinPad = kWin + (kWin >> 3) + 8;

// [T, A, 3] -> [A, 3, T] through shared tiles of kTrT frames x 32 columns (column = atom*3+c):
// eight loads per thread are in flight before the first store, output runs are kTrT * 8 bytes
constexpr int kTrT = 64;
__global__ void __launch_bounds__(256)
    k_to_soa(const double *__restrict__ in, int64_t n_frames, int64_t n_cols,
             double *__restrict__ out) {
    __shared__ double tile[kTrT][33];
    const int64_t t0 = (int64_t)blockIdx.x * kTrT, c0 = (int64_t)blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    double v[kTrT / 8];
#pragma unroll
    for (int k = 0; k < kTrT / 8; ++k) {
        const int64_t t = t0 + ty + 8 * k, c = c0 + tx;
        v[k] = (t < n_frames && c < n_cols) ? __ldcs(in + t * n_cols + c) : 0.0;
    }
#pragma unroll
    for (int k = 0; k < kTrT / 8; ++k) tile[ty + 8 * k][tx] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int r = ty + 8 * k;  // column of the tile = row of the output
        const int64_t c = c0 + r;
        if (c >= n_cols) continue;
#pragma unroll
        for (int h = 0; h < kTrT / 32; ++h) {
            const int64_t t = t0 + tx + 32 * h;
            if (t < n_frames) __stcs(out + c * n_frames + t, tile[tx + 32 * h][r]);
        }
    }
}

template <bool kCheck>
__device__ __forceinline__ void msd_tile(const double *__restrict__ si, const double *__restrict__ sj,
                                         int base, int64_t t0, int64_t k0, int64_t n_frames,
                                         int n_i, double (&acc)[kM]) {
    // si: [3][kBT] tile of r[t0 + i]; sj: [3][kWinPad] window, element j <-> frame t0 + K0 + j
    // this thread's lags: k0 + m, window index for (i, m) = i + base + m
    double w[3][kM];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int m = 0; m < kM; ++m) w[c][m] = sj[c * kWinPad + padded(base + m)];
    for (int i0 = 0; i0 < n_i; i0 += kM) {
#pragma unroll
        for (int s = 0; s < kM; ++s) {
            const int i = i0 + s;
            if (s > 0 || i0 > 0) {  // slide: slot (s-1) receives element i + kM - 1
                const int j = base + i + kM - 1;
#pragma unroll
                for (int c = 0; c < 3; ++c) w[c][(s + kM - 1) % kM] = sj[c * kWinPad + padded(j)];
            }
            const double x = si[i], y = si[kBT + i], z = si[2 * kBT + i];
#pragma unroll
            for (int m = 0; m < kM; ++m) {
                const int slot = (s + m) % kM;
                const double dx = w[0][slot] - x, dy = w[1][slot] - y, dz = w[2][slot] - z;
                double term = fma(dz, dz, fma(dy, dy, dx * dx));
                if (kCheck) {
                    const bool ok = (i < n_i) && (t0 + i + k0 + m < n_frames);
                    term = ok ? term : 0.0;
                }
                acc[m] += term;
            }
        }
    }
}

__global__ void __launch_bounds__(kMsdThreads, 2)
    k_msd(const double *__restrict__ soa, int64_t n_frames, int32_t n_atoms, int64_t lag0,
          int64_t n_lags, double *__restrict__ msd) {
    extern __shared__ __align__(16) double sm[];
    double *si = sm;             // [3][kBT]
    double *sj = sm + 3 * kBT;   // [3][kWinPad]
    const int32_t a = blockIdx.y;
    const int64_t K0 = lag0 + (int64_t)blockIdx.x * kLagsPerBlock;  // first lag of this CTA
    const int base = threadIdx.x * kM;
    const int64_t k0 = K0 + base;  // first lag of this thread
    const double *r = soa + (int64_t)a * 3 * n_frames;
    double acc[kM];
#pragma unroll
    for (int m = 0; m < kM; ++m) acc[m] = 0.0;

    const int64_t n_t = n_frames - K0;  // frames t with at least one valid pair in this CTA
    for (int64_t t0 = 0; t0 < n_t; t0 += kBT) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < 3 * kBT; idx += kMsdThreads) {
            const int c = idx / kBT, i = idx - c * kBT;
            const int64_t t = t0 + i;
            si[idx] = t < n_frames ? r[(int64_t)c * n_frames + t] : 0.0;
        }
        for (int idx = threadIdx.x; idx < 3 * kWin; idx += kMsdThreads) {
            const int c = idx / kWin, j = idx - c * kWin;
            const int64_t t = t0 + K0 + j;
            sj[c * kWinPad + padded(j)] = t < n_frames ? r[(int64_t)c * n_frames + t] : 0.0;
        }
        __syncthreads();
        const int64_t left = n_t - t0;
        const int n_i = (int)(left < kBT ? left : kBT);
        // the whole tile is in range iff its last frame paired with the CTA's last lag is
        const bool full = (n_i == kBT) && (t0 + kBT - 1 + K0 + kLagsPerBlock - 1 < n_frames);
        if (full)
            msd_tile<false>(si, sj, base, t0, k0, n_frames, kBT, acc);
        else
            msd_tile<true>(si, sj, base, t0, k0, n_frames, n_i, acc);
    }
#pragma unroll
    for (int m = 0; m < kM; ++m) {
        const int64_t k = k0 + m;
        const int64_t j = k - lag0;
        if (j < n_lags && k < n_frames) {
            msd[(int64_t)a * n_lags + j] = k == 0 ? 0.0 : acc[m] / (double)(n_frames - k);
        }
    }
}

}  // namespace

extern "C" gd_status_t gd_transpose_soa(const double *d_in, int64_t n_frames, int32_t n_atoms,
                                        double *d_out, void *stream) {
    GD_REQUIRE(d_in && d_out && n_frames > 0 && n_atoms > 0, GD_ERR_INVALID,
               "gd_transpose_soa: bad argument");
    const int64_t n_cols = (int64_t)n_atoms * 3;
    dim3 grid((unsigned)gd_ceil_div(n_frames, kTrT), (unsigned)gd_ceil_div(n_cols, 32));
    GD_REQUIRE(grid.y <= 65535, GD_ERR_INVALID, "gd_transpose_soa: too many atoms");
    k_to_soa<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, n_frames, n_cols, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_transpose_f64(const double *d_in, int64_t n_rows, int64_t n_cols,
                                        double *d_out, void *stream) {
    GD_REQUIRE(d_in && d_out && n_rows > 0 && n_cols > 0, GD_ERR_INVALID,
               "gd_transpose_f64: bad argument");
    dim3 grid((unsigned)gd_ceil_div(n_rows, kTrT), (unsigned)gd_ceil_div(n_cols, 32));
    GD_REQUIRE(grid.y <= 65535, GD_ERR_INVALID, "gd_transpose_f64: too many columns");
    k_to_soa<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, n_rows, n_cols, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_msd_direct(const double *d_soa, int64_t n_frames, int32_t n_atoms,
                                     int64_t lag0, int64_t n_lags, double *d_msd, void *stream) {
    GD_REQUIRE(d_soa && d_msd, GD_ERR_INVALID, "gd_msd_direct: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0 && n_atoms <= 65535, GD_ERR_INVALID,
               "gd_msd_direct: n_atoms must be in [1, 65535]");
    GD_REQUIRE(lag0 >= 0 && n_lags > 0 && lag0 + n_lags <= n_frames, GD_ERR_INVALID,
               "gd_msd_direct: lag range must lie inside [0, n_frames)");
    const size_t smem = (size_t)(3 * kBT + 3 * kWinPad) * sizeof(double);
    GD_CUDA_CHECK(cudaFuncSetAttribute(k_msd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)gd_ceil_div(n_lags, kLagsPerBlock), (unsigned)n_atoms);
    k_msd<<<grid, kMsdThreads, smem, (cudaStream_t)stream>>>(d_soa, n_frames, n_atoms, lag0, n_lags,
                                                            d_msd);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
