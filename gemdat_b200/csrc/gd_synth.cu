// Synthetic trajectory generator (bench / tests).  The recipe is specified in
// gemdat_b200/synth.py, which replays it bit-identically in numpy for any frame
// window: only exactly rounded IEEE operations and integer hashing are used.
#include "gd_common.cuh"

namespace {

__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

// Irwin-Hall(4) integer in [-131070, 131070]
__device__ __forceinline__ int noise_knot(uint64_t seed, int32_t a, int c, int64_t k) {
    const uint64_t h = splitmix64(seed ^ (((uint64_t)a * 3ULL + (uint64_t)c) * 0xD1B54A32D192ED03ULL +
                                          (uint64_t)k * 0x8CB92BA72F3D8DD7ULL));
    const int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) +
                  (int)((h >> 48) & 0xFFFF);
    return s - 131070;
}

struct SynthParams {
    const double *centre;
    const double *delta;
    int32_t n_atoms, n_epochs, epoch_len, knot_len;
    double scale[3];
    uint64_t seed;
    int64_t frame0, n_frames;
    double *pos;
};

__global__ void __launch_bounds__(256) k_synth(const __grid_constant__ SynthParams p) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = p.n_frames * p.n_atoms;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; idx < total; idx += stride) {
        const int64_t tl = idx / p.n_atoms;
        const int32_t a = (int32_t)(idx - tl * p.n_atoms);
        const int64_t t = p.frame0 + tl;
        const int64_t e = t / p.epoch_len;
        const int phi = (int)(t - e * p.epoch_len);
        const int w8 = max(0, phi - (p.epoch_len - 8));
        const double w = (double)w8 * 0.125;
        const int64_t k = t / p.knot_len;
        const int q = (int)(t - k * p.knot_len);
        const double *cen = p.centre + ((int64_t)a * p.n_epochs + e) * 3;
        const double *del = p.delta + ((int64_t)a * p.n_epochs + e) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int g0 = noise_knot(p.seed, a, c, k), g1 = noise_knot(p.seed, a, c, k + 1);
            const int nint = g0 * p.knot_len + (g1 - g0) * q;
            double x = __dadd_rn(cen[c], __dmul_rn(w, del[c]));
            x = __dadd_rn(x, __dmul_rn((double)nint, p.scale[c]));
            x = __dsub_rn(x, floor(x));
            if (x >= 1.0) x = 0.0;
            p.pos[idx * 3 + c] = x;
        }
    }
}

}  // namespace

extern "C" gd_status_t gd_synth_positions(const double *d_centre, const double *d_delta,
                                          int32_t n_atoms, int32_t n_epochs, int32_t epoch_len,
                                          int32_t knot_len, const double *h_noise_scale,
                                          uint64_t seed, int64_t frame0, int64_t n_frames,
                                          double *d_pos_frac, void *stream) {
    GD_REQUIRE(d_centre && d_delta && h_noise_scale && d_pos_frac, GD_ERR_INVALID,
               "gd_synth_positions: null argument");
    GD_REQUIRE(n_atoms > 0 && n_epochs > 0 && epoch_len >= 16 && knot_len > 0 &&
                   epoch_len % knot_len == 0 && frame0 >= 0 && n_frames >= 0,
               GD_ERR_INVALID, "gd_synth_positions: bad sizes");
    GD_REQUIRE((frame0 + n_frames + epoch_len - 1) / epoch_len <= n_epochs, GD_ERR_INVALID,
               "gd_synth_positions: schedule has too few epochs");
    if (n_frames == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    SynthParams p;
    p.centre = d_centre;
    p.delta = d_delta;
    p.n_atoms = n_atoms;
    p.n_epochs = n_epochs;
    p.epoch_len = epoch_len;
    p.knot_len = knot_len;
    for (int c = 0; c < 3; ++c) p.scale[c] = h_noise_scale[c];
    p.seed = seed;
    p.frame0 = frame0;
    p.n_frames = n_frames;
    p.pos = d_pos_frac;
    int64_t blocks = gd_ceil_div(n_frames * n_atoms, 256);
    const int64_t cap = (int64_t)dev.sm_count * 16;
    if (blocks > cap) blocks = cap;
    k_synth<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
