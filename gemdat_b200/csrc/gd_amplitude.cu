// a4 -- vibration amplitudes from the speed array.
//
// Replaces TrajectoryMetrics.amplitudes (reference src/gemdat/metrics.py:209-231: a Python
// loop over atoms with np.sign / np.roll / np.array_split / np.sum inside) and the
// mean / standard deviation of vibration_amplitude (:190-207).
//
// The reference cuts every atom's speed row after each sign flip except the first and the
// last one (`splits[1:-1] + 1`; a flip at index t means sign(s[t]) != sign(s[t+1]), the last
// element is compared with the first through np.roll) and sums the pieces.  With
//   S_i = sum of the elements after flip i-1 up to and including flip i   (i = 0 .. n-1)
//   R   = sum of the elements after the last flip
// the amplitudes of an atom are [S_0 + S_1, S_2, ..., S_{n-2}, S_{n-1} + R] (n - 1 values;
// one value, the whole row, when n < 2).  So this is the ordered compaction K2 does for
// events with a floating-point payload: the speed array is kept TIME-major [T, A] (lane =
// atom, a warp reads 32 neighbouring atoms of one frame: coalesced), one thread owns one
// (frame chunk, atom) segment:
//   summary : flips per segment + the open sum after its last flip
//   carry   : per atom over the chunks: rank of the segment's first flip, sum carried in
//   fill    : S_i written at atom_offset + i - 1 (S_0 kept aside), in time order
//   fix-up  : S_0 and R added to the atom's first / last amplitude
// Sums run left to right inside a piece (numpy's pairwise np.sum groups differently beyond
// 8 elements: agreement is to a few ulp, the tests state the tolerance).
//
// Algorithmic traffic: 8 B read per atom-frame per pass (two passes) + 8 B per amplitude.
#include "gd_common.cuh"

namespace {

typedef GdChunks Geometry;

__device__ __forceinline__ int sgn(double x) { return (x > 0.0) - (x < 0.0); }

struct AmpWork {
    double *carry;    // [n_chunks * n_atoms] open sum after the last flip -> sum carried in
    double *s0;       // [n_atoms]
    double *rest;     // [n_atoms]
    int32_t *rank;    // [n_chunks * n_atoms] flips in the segment -> rank of its first flip
    int32_t *n_flips; // [n_atoms]
};

__host__ __device__ inline int64_t align16(int64_t x) { return (x + 15) & ~(int64_t)15; }

inline AmpWork carve(void *work, const Geometry &g) {
    const int64_t items = g.n_chunks * g.n_atoms;
    char *p = static_cast<char *>(work);
    AmpWork w;
    w.carry = reinterpret_cast<double *>(p);
    p += align16(items * 8);
    w.s0 = reinterpret_cast<double *>(p);
    p += align16((int64_t)g.n_atoms * 8);
    w.rest = reinterpret_cast<double *>(p);
    p += align16((int64_t)g.n_atoms * 8);
    w.rank = reinterpret_cast<int32_t *>(p);
    p += align16(items * 4);
    w.n_flips = reinterpret_cast<int32_t *>(p);
    return w;
}

// One thread walks one (chunk, atom) segment of the time-major speed array.  kFill = false:
// count flips and leave the open sum; kFill = true: write the finished sums.
template <bool kFill>
__global__ void __launch_bounds__(256)
    k_amp_walk(const double *__restrict__ speed, Geometry g, AmpWork w,
               const int64_t *__restrict__ atom_off, double *__restrict__ out) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= g.n_chunks * g.n_atoms) return;
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
    const int64_t A = g.n_atoms;
    const int64_t t0 = c * g.f, t1 = min(g.n_frames, t0 + g.f);
    const double *q = speed + t0 * A + a;
    double acc = kFill ? w.carry[item] : 0.0;
    int rank = kFill ? w.rank[item] : 0;
    const int64_t base = kFill ? atom_off[a] : 0;
    double cur = __ldcs(q);
    int s_cur = sgn(cur);
    constexpr int kAhead = 8;
    auto step = [&](double nxt) {
        const int s_nxt = sgn(nxt);
        acc += cur;
        if (s_cur != s_nxt) {  // flip at this element: it closes the piece
            if (kFill) {
                if (rank == 0)
                    w.s0[a] = acc;
                else
                    out[base + rank - 1] = acc;
            }
            ++rank;
            acc = 0.0;
        }
        cur = nxt;
        s_cur = s_nxt;
    };
    int64_t t = t0;
    // elements t0 .. t1-2 are compared with their successor inside the segment
    for (; t + kAhead < t1; t += kAhead) {
        double v[kAhead];
#pragma unroll
        for (int j = 0; j < kAhead; ++j) v[j] = __ldcs(q + (int64_t)(j + 1) * A);
        q += kAhead * A;
#pragma unroll
        for (int j = 0; j < kAhead; ++j) step(v[j]);
    }
    for (; t + 1 < t1; ++t) {
        q += A;
        step(__ldcs(q));
    }
    // last element of the segment: its successor is the first element of the next chunk, or
    // (np.roll) the atom's first element when the segment ends the row
    step(t1 < g.n_frames ? speed[t1 * A + a] : speed[a]);
    if (!kFill) {
        w.rank[item] = rank;
        w.carry[item] = acc;
    }
}

// warp per atom: exclusive scan over the chunks of (flips, open sum) with the segmented-sum
// operator  (n1, s1) o (n2, s2) = (n1 + n2, n2 > 0 ? s2 : s1 + s2), 32 chunks per step;
// amplitude count of the atom -> amp_count[a + 1]
__global__ void __launch_bounds__(128)
    k_amp_carry(Geometry g, AmpWork w, int64_t *__restrict__ amp_count) {
    const int32_t a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (a >= g.n_atoms) return;
    int32_t run = 0;
    double carry = 0.0;
    for (int64_t c0 = 0; c0 < g.n_chunks; c0 += 32) {
        const int64_t c = c0 + lane;
        const bool on = c < g.n_chunks;
        const int64_t k = c * g.n_atoms + a;
        const int32_t n = on ? w.rank[k] : 0;
        const double open = on ? w.carry[k] : 0.0;
        int32_t in = n;
        double is = open;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t un = __shfl_up_sync(0xffffffffu, in, o);
            const double us = __shfl_up_sync(0xffffffffu, is, o);
            if (lane >= o) {
                is = in > 0 ? is : us + is;
                in += un;
            }
        }
        // exclusive values: the inclusive ones of the lane before, after the running carry
        int32_t en = __shfl_up_sync(0xffffffffu, in, 1);
        double es = __shfl_up_sync(0xffffffffu, is, 1);
        if (lane == 0) {
            en = 0;
            es = 0.0;
        }
        if (on) {
            w.rank[k] = run + en;
            w.carry[k] = en > 0 ? es : carry + es;
        }
        const int32_t tn = __shfl_sync(0xffffffffu, in, 31);
        const double ts = __shfl_sync(0xffffffffu, is, 31);
        carry = tn > 0 ? ts : carry + ts;
        run += tn;
    }
    if (lane == 0) {
        w.n_flips[a] = run;
        w.rest[a] = carry;
        amp_count[a + 1] = run >= 2 ? run - 1 : 1;
    }
}

__global__ void __launch_bounds__(128)
    k_amp_fixup(int32_t n_atoms, AmpWork w, const int64_t *__restrict__ atom_off,
                double *__restrict__ out) {
    const int32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_atoms) return;
    const int32_t n = w.n_flips[a];
    const int64_t o = atom_off[a];
    if (n >= 2) {
        out[o] = w.s0[a] + out[o];
        out[o + n - 2] += w.rest[a];
    } else if (n == 1) {
        out[o] = w.s0[a] + w.rest[a];
    } else {
        out[o] = w.rest[a];
    }
}

// ---------------------------------------------------------------- mean / standard deviation

constexpr int kRedThreads = 256;

__device__ __forceinline__ double block_sum(double v) {
    __shared__ double warp_tot[kRedThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < kRedThreads / 32; ++k) s += warp_tot[k];
    __syncthreads();
    return s;
}

// pass 0: per-block sums of x; pass 1: per-block sums of (x - mean)^2 with mean = *mean_in
__global__ void __launch_bounds__(kRedThreads)
    k_moment(const double *__restrict__ x, int64_t n, const double *__restrict__ mean_in,
             double *__restrict__ partial) {
    const double mu = mean_in ? *mean_in : 0.0;
    const int64_t stride = (int64_t)gridDim.x * kRedThreads;
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * kRedThreads + threadIdx.x; i < n; i += stride) {
        const double v = x[i];
        if (mean_in) {
            const double d = v - mu;
            acc = fma(d, d, acc);
        } else {
            acc += v;
        }
    }
    acc = block_sum(acc);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// single block: out[0] = (sum of the partials) / n   (fixed order: deterministic)
__global__ void __launch_bounds__(kRedThreads)
    k_moment_final(const double *__restrict__ partial, int n_partial, int64_t n, int take_sqrt,
                   double *__restrict__ out) {
    double acc = 0.0;
    for (int i = threadIdx.x; i < n_partial; i += kRedThreads) acc += partial[i];
    acc = block_sum(acc);
    if (threadIdx.x == 0) {
        const double m = acc / (double)n;
        out[0] = take_sqrt ? sqrt(m) : m;
    }
}

// ---------------------------------------------------------------- spectral moments (N2)
//
// meanfreq of utils.py:128-160 from the one-sided spectrum X[atom, k] (k = 0 .. n/2, complex128,
// one row per atom -- the layout a batched real-to-complex transform leaves): with
// P_k = w_k |X_k|^2 (w = 2 except for DC and, for even n, Nyquist: scipy.signal.periodogram's
// one-sided scaling) the mean frequency is sum_k f_k P_k / sum_k P_k, f_k = k fs / n -- the
// common factors 1/(fs n) and (f_1 - f_0) of the reference cancel in the ratio.  The DC bin is
// skipped: after the constant detrend of the periodogram it holds |sum(x - mean)|^2, rounding
// noise.  One pass over the spectrum, one CTA per atom, fixed-order reduction (deterministic).
__global__ void __launch_bounds__(kRedThreads)
    k_spec_meanfreq(const double2 *__restrict__ spec, int64_t n_bins, int64_t n, double scale,
                    double *__restrict__ out) {
    const double2 *row = spec + (int64_t)blockIdx.x * n_bins;
    double m0 = 0.0, m1 = 0.0;
    for (int64_t k = 1 + threadIdx.x; k < n_bins; k += kRedThreads) {
        const double2 x = row[k];
        const double w = (2 * k == n) ? 1.0 : 2.0;
        const double pwr = w * fma(x.x, x.x, x.y * x.y);
        m0 += pwr;
        m1 = fma((double)k, pwr, m1);
    }
    m0 = block_sum(m0);
    m1 = block_sum(m1);
    if (threadIdx.x == 0) out[blockIdx.x] = (m1 / m0) * scale;
}

}  // namespace

extern "C" gd_status_t gd_meanfreq(const double *d_spectrum, int64_t n_samples, int32_t n_atoms,
                                   double sampling_frequency, double *d_meanfreq, void *stream) {
    GD_REQUIRE(d_spectrum && d_meanfreq && n_samples > 0 && n_atoms > 0, GD_ERR_INVALID,
               "gd_meanfreq: bad argument");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_spectrum) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_meanfreq: the spectrum must be 16-byte aligned");
    const int64_t n_bins = n_samples / 2 + 1;
    k_spec_meanfreq<<<(unsigned)n_atoms, kRedThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const double2 *>(d_spectrum), n_bins, n_samples,
        sampling_frequency / (double)n_samples, d_meanfreq);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" int64_t gd_amplitudes_work_bytes(int64_t n_frames, int32_t n_atoms) {
    if (n_frames <= 0 || n_atoms <= 0) return 64;
    const Geometry g = gd_chunks(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    return align16(items * 8) + 2 * align16((int64_t)n_atoms * 8) + align16(items * 4) +
           align16((int64_t)n_atoms * 4) + 64;
}

extern "C" gd_status_t gd_amplitudes_count(const double *d_speed_tm, int64_t n_frames,
                                           int32_t n_atoms, void *d_work,
                                           int64_t *d_atom_offsets, int64_t *h_n_amplitudes,
                                           void *stream) {
    GD_REQUIRE(d_speed_tm && d_work && d_atom_offsets, GD_ERR_INVALID,
               "gd_amplitudes_count: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_amplitudes_count: empty input");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_work) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_amplitudes_count: d_work must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const Geometry g = gd_chunks(n_frames, n_atoms);
    const AmpWork w = carve(d_work, g);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_amp_walk<false><<<(unsigned)gd_ceil_div(items, 256), 256, 0, st>>>(d_speed_tm, g, w, nullptr,
                                                                         nullptr);
    GD_LAUNCH_CHECK();
    k_amp_carry<<<(unsigned)gd_ceil_div((int64_t)n_atoms * 32, 128), 128, 0, st>>>(g, w, d_atom_offsets);
    GD_LAUNCH_CHECK();
    gd_k_scan_offsets<<<1, 1024, 0, st>>>(d_atom_offsets, (int64_t)n_atoms);
    GD_LAUNCH_CHECK();
    if (h_n_amplitudes) {
        GD_CUDA_CHECK(cudaMemcpyAsync(h_n_amplitudes, d_atom_offsets + n_atoms, sizeof(int64_t),
                                      cudaMemcpyDeviceToHost, st));
        GD_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    return GD_OK;
}

extern "C" gd_status_t gd_amplitudes_fill(const double *d_speed_tm, int64_t n_frames,
                                          int32_t n_atoms, const void *d_work,
                                          const int64_t *d_atom_offsets, double *d_amplitudes,
                                          void *stream) {
    GD_REQUIRE(d_speed_tm && d_work && d_atom_offsets && d_amplitudes, GD_ERR_INVALID,
               "gd_amplitudes_fill: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_amplitudes_fill: empty input");
    cudaStream_t st = (cudaStream_t)stream;
    const Geometry g = gd_chunks(n_frames, n_atoms);
    const AmpWork w = carve(const_cast<void *>(d_work), g);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_amp_walk<true><<<(unsigned)gd_ceil_div(items, 256), 256, 0, st>>>(d_speed_tm, g, w,
                                                                        d_atom_offsets, d_amplitudes);
    GD_LAUNCH_CHECK();
    k_amp_fixup<<<(unsigned)gd_ceil_div(n_atoms, 128), 128, 0, st>>>(n_atoms, w, d_atom_offsets,
                                                                     d_amplitudes);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" int64_t gd_mean_std_work_bytes(void) { return (1024 + 8) * (int64_t)sizeof(double); }

extern "C" gd_status_t gd_mean_std(const double *d_x, int64_t n, void *d_work, double *h_mean,
                                   double *h_std, void *stream) {
    GD_REQUIRE(d_x && d_work && h_mean && h_std && n > 0, GD_ERR_INVALID, "gd_mean_std: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    double *partial = static_cast<double *>(d_work);
    double *res = partial + 1024;  // [0] mean, [1] std
    int blocks = (int)(gd_ceil_div(n, kRedThreads * 8) < 1024 ? gd_ceil_div(n, kRedThreads * 8) : 1024);
    if (blocks < 1) blocks = 1;
    k_moment<<<blocks, kRedThreads, 0, st>>>(d_x, n, nullptr, partial);
    GD_LAUNCH_CHECK();
    k_moment_final<<<1, kRedThreads, 0, st>>>(partial, blocks, n, 0, res);
    GD_LAUNCH_CHECK();
    k_moment<<<blocks, kRedThreads, 0, st>>>(d_x, n, res, partial);
    GD_LAUNCH_CHECK();
    k_moment_final<<<1, kRedThreads, 0, st>>>(partial, blocks, n, 1, res + 1);
    GD_LAUNCH_CHECK();
    double h[2];
    GD_CUDA_CHECK(cudaMemcpyAsync(h, res, sizeof(h), cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));
    *h_mean = h[0];
    *h_std = h[1];
    return GD_OK;
}
