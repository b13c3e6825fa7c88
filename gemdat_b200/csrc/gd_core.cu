// Error reporting, device info and the small utility entry points.
#include <stdarg.h>
#include <string.h>

#include "gd_common.cuh"

static thread_local char g_err[512] = "";

void gd_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *gd_last_error(void) { return g_err; }
extern "C" int gd_version(void) { return 100; }

const GdDeviceInfo &gd_device_info() {
    static GdDeviceInfo info = [] {
        GdDeviceInfo d{-1, 0};
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return d;
        int sm = 0, smem = 0;
        if (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return d;
        if (cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) !=
            cudaSuccess)
            return d;
        d.sm_count = sm;
        d.max_smem_optin = smem;
        return d;
    }();
    return info;
}

extern "C" int gd_sm_count(void) { return gd_device_info().sm_count; }

// ------------------------------------------------------------------ int32 -> int64
__global__ void __launch_bounds__(256) k_widen(const int32_t *__restrict__ in,
                                               long long *__restrict__ out, int64_t n, int aligned16) {
    // 4 elements per thread: one 16-byte load, two 16-byte stores (frame slices of a state
    // array or rows of a part's event table may start off a 16-byte boundary: element-wise then)
    int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t n4 = aligned16 ? n >> 2 : 0;
    const int4 *in4 = reinterpret_cast<const int4 *>(in);
    longlong2 *out2 = reinterpret_cast<longlong2 *>(out);
    for (int64_t i = i4; i < n4; i += stride) {
        int4 v = __ldg(in4 + i);
        out2[2 * i] = make_longlong2(v.x, v.y);
        out2[2 * i + 1] = make_longlong2(v.z, v.w);
    }
    for (int64_t i = (n4 << 2) + i4; i < n; i += stride) out[i] = in[i];
}

extern "C" gd_status_t gd_widen_i32_i64(const int32_t *d_in, int64_t *d_out, int64_t n,
                                        void *stream) {
    GD_REQUIRE(n >= 0 && (n == 0 || (d_in && d_out)), GD_ERR_INVALID, "gd_widen_i32_i64: bad args");
    if (n == 0) return GD_OK;
    const int aligned16 = (reinterpret_cast<uintptr_t>(d_in) & 15) == 0 &&
                          (reinterpret_cast<uintptr_t>(d_out) & 15) == 0;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    int64_t blocks = gd_ceil_div(gd_ceil_div(n, 4), 256);
    int64_t cap = (int64_t)dev.sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    k_widen<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        d_in, reinterpret_cast<long long *>(d_out), n, aligned16);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

// ------------------------------------------------------------------ wrap to [0, 1)
// Trajectory.to_positions ends with np.mod(coords, 1) (reference src/gemdat/trajectory.py:121-127).
// d_out == NULL: only report whether any coordinate lies outside [0, 1).
__global__ void __launch_bounds__(256) k_wrap_unit(const double *__restrict__ in,
                                                   double *__restrict__ out, int64_t n,
                                                   int32_t *__restrict__ flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    bool outside = false;
    for (; i < n; i += stride) {
        const double x = in[i];
        const bool bad = !(x >= 0.0 && x < 1.0);
        outside |= bad;
        if (out) out[i] = bad ? x - floor(x) : x;  // np.mod(x, 1): fmod, +1 when negative
    }
    if (outside) *flag = 1;
}

extern "C" gd_status_t gd_wrap_unit(const double *d_in, double *d_out, int64_t n,
                                    int32_t *d_flag, void *stream) {
    GD_REQUIRE(n >= 0 && d_flag && (n == 0 || d_in), GD_ERR_INVALID, "gd_wrap_unit: bad args");
    if (n == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    int64_t blocks = gd_ceil_div(n, 256);
    const int64_t cap = (int64_t)dev.sm_count * 16;
    if (blocks > cap) blocks = cap;
    k_wrap_unit<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n, d_flag);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

// ------------------------------------------------------------------ LUT remap
__global__ void __launch_bounds__(256) k_remap(int32_t *__restrict__ s, int64_t n,
                                               const int32_t *__restrict__ lut, int32_t n_sites) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int32_t v = s[i];
        if (v >= 0 && v < n_sites) {
            int32_t w = __ldg(lut + v);
            if (w != v) s[i] = w;
        }
    }
}

extern "C" gd_status_t gd_remap_states(int32_t *d_states, int64_t n, const int32_t *d_lut,
                                       int32_t n_sites, void *stream) {
    GD_REQUIRE(n >= 0 && d_states && d_lut && n_sites > 0, GD_ERR_INVALID,
               "gd_remap_states: bad args");
    if (n == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    int64_t blocks = gd_ceil_div(n, 256);
    int64_t cap = (int64_t)dev.sm_count * 8;
    if (blocks > cap) blocks = cap;
    k_remap<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_states, n, d_lut, n_sites);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

// ------------------------------------------------------------------ fp64 peak probe
__global__ void __launch_bounds__(256) k_fp64_peak(int64_t iters, double *sink) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    double a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int64_t i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456) sink[0] = s;  // keep the chain alive
}

extern "C" gd_status_t gd_fp64_peak(int64_t iters, double *h_dfma_per_s, void *stream) {
    GD_REQUIRE(iters > 0 && h_dfma_per_s, GD_ERR_INVALID, "gd_fp64_peak: bad args");
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    cudaStream_t st = (cudaStream_t)stream;
    double *sink = nullptr;
    GD_CUDA_CHECK(cudaMalloc(&sink, sizeof(double)));
    cudaEvent_t e0, e1;
    GD_CUDA_CHECK(cudaEventCreate(&e0));
    GD_CUDA_CHECK(cudaEventCreate(&e1));
    int blocks = dev.sm_count * 8;
    k_fp64_peak<<<blocks, 256, 0, st>>>(iters / 8 + 1, sink);  // warm-up
    GD_CUDA_CHECK(cudaEventRecord(e0, st));
    k_fp64_peak<<<blocks, 256, 0, st>>>(iters, sink);
    GD_CUDA_CHECK(cudaEventRecord(e1, st));
    GD_CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0.f;
    GD_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    *h_dfma_per_s = (double)blocks * 256.0 * 8.0 * (double)iters / (ms * 1e-3);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    return GD_OK;
}
