// K5 -- density histogram of fractional positions.
//
// Replaces trajectory_to_volume (reference src/gemdat/volume.py:449-509): three
// np.digitize calls against np.linspace(0, 1, n)[1:] edges, then an
// np.unique(axis=0, return_counts=True) sort of all T*A rows.
//
// Bin index: the number of edges e_j = fl(j * fl(1/(n-1))), j = 1..n-1 (last one
// exactly 1.0), that are <= x -- evaluated exactly as np.digitize would, by
// correcting the guess floor(x*(n-1)) against the two neighbouring edges.
#include "gd_common.cuh"

namespace {

struct Axis {
    int n_edges;   // n
    double step;   // fl(1 / (n - 1))
};

__device__ __forceinline__ int bin_of(const Axis &ax, double x) {
    const int last = ax.n_edges - 2;  // highest voxel index
    if (last <= 0) return 0;
    // guess floor(x * (n - 1)); the exact answer -- the number of edges <= x, np.digitize --
    // differs from it by at most one (the scaled coordinate is off by a few ulp, a bin is one
    // unit wide), so ONE comparison with each neighbouring exactly-computed edge settles it
    int g = min((int)(x * (double)(ax.n_edges - 1)), last);
    const double lo = __dmul_rn((double)g, ax.step);  // edge(g); 0 for g = 0
    const double hi = g >= last ? 1.0 : __dmul_rn((double)(g + 1), ax.step);
    return g + (int)(hi <= x) - (int)(lo > x);
}

struct DensityParams {
    Axis ax[3];
    int64_t n;
};

// v1: one global RED per atom-frame (grid mostly L2 resident)
__global__ void __launch_bounds__(256)
    k_density(const double *__restrict__ pos, const __grid_constant__ DensityParams p,
              unsigned int *__restrict__ grid, int *__restrict__ range_flag) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const int64_t ny = p.ax[1].n_edges - 1, nz = p.ax[2].n_edges - 1;
    bool bad = false;
    for (; i < p.n; i += step) {
        const double x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
        if (!(x >= 0.0 && x < 1.0 && y >= 0.0 && y < 1.0 && z >= 0.0 && z < 1.0)) {
            bad = true;
            continue;
        }
        const int64_t v = ((int64_t)bin_of(p.ax[0], x) * ny + bin_of(p.ax[1], y)) * nz +
                          bin_of(p.ax[2], z);
        atomicAdd(grid + v, 1u);
    }
    if (bad) *range_flag = 1;
}

// v2: strip walker.  One warp owns 32 neighbouring atoms (lane = atom) and walks a chunk of
// frames, so a lane sees ONE atom's consecutive frames: an atom moves a fraction of a 0.1 A
// voxel per MD step, so consecutive frames mostly fall into the same voxel and are counted in
// registers (run-length); one RED per RUN leaves the SM instead of one per atom-frame
// (measured on the C3 recipe: 0.44 runs per atom-frame; a shared-memory window or hash per
// (atom, 256 frames) would only get to 0.37 distinct voxels per atom-frame).
// A frame row of the strip is 768 contiguous bytes.  kSmem: the whole grid fits the CTA's shared
// memory (coarse grids, small cells): runs are counted into a CTA-private copy of the grid
// with shared-memory atomics and the copy is added to the global grid once at the end.
constexpr int kDenWarps = 4;
constexpr int kDenAhead = 4;  // frame rows fetched ahead per lane
constexpr int kDenMaxSmemVoxels = 48 * 1024;

struct DensityFrames {
    Axis ax[3];
    int64_t n_frames;
    int32_t n_atoms;
    int32_t n_strips;
    int32_t chunk;       // frames per warp
    int64_t n_chunks;
    int32_t n_voxels;    // kSmem only
};

template <bool kSmem>
__global__ void __launch_bounds__(kDenWarps * 32)
    k_density_runs(const double *__restrict__ pos, const __grid_constant__ DensityFrames p,
                   unsigned int *__restrict__ grid, int *__restrict__ range_flag) {
    extern __shared__ unsigned int s_grid[];
    if (kSmem) {
        for (int i = threadIdx.x; i < p.n_voxels; i += blockDim.x) s_grid[i] = 0u;
        __syncthreads();
    }
    unsigned int *target = kSmem ? s_grid : grid;
    const int lane = threadIdx.x & 31;
    const int64_t n_items = (int64_t)p.n_strips * p.n_chunks;
    const int ny = p.ax[1].n_edges - 1, nz = p.ax[2].n_edges - 1;
    const int64_t A3 = (int64_t)p.n_atoms * 3;
    bool bad = false;
    // strip-major item order: the CTAs resident together walk the same atoms' neighbourhoods
    for (int64_t item = (int64_t)blockIdx.x * kDenWarps + (threadIdx.x >> 5); item < n_items;
         item += (int64_t)gridDim.x * kDenWarps) {
        const int32_t strip = (int32_t)(item / p.n_chunks);
        const int64_t chunk = item - (int64_t)strip * p.n_chunks;
        const int32_t a = strip * 32 + lane;
        const bool on = a < p.n_atoms;
        const int64_t t0 = chunk * p.chunk;
        const int n_own = (int)min((int64_t)p.chunk, p.n_frames - t0);
        const double *q = pos + t0 * A3 + 3 * (int64_t)(on ? a : p.n_atoms - 1);
        int32_t run_voxel = -1;
        unsigned int run = 0;
        auto count = [&](double x, double y, double z) {
            if (!(x >= 0.0 && x < 1.0 && y >= 0.0 && y < 1.0 && z >= 0.0 && z < 1.0)) {
                bad = true;
                return;
            }
            const int32_t v = (bin_of(p.ax[0], x) * ny + bin_of(p.ax[1], y)) * nz + bin_of(p.ax[2], z);
            if (v == run_voxel) {
                ++run;
            } else {
                if (run) atomicAdd(target + run_voxel, run);
                run_voxel = v;
                run = 1;
            }
        };
        int k = 0;
        for (; k + kDenAhead <= n_own; k += kDenAhead) {
            double x[kDenAhead], y[kDenAhead], z[kDenAhead];
#pragma unroll
            for (int j = 0; j < kDenAhead; ++j) {
                const double *r = q + (int64_t)j * A3;
                x[j] = __ldcs(r);
                y[j] = __ldcs(r + 1);
                z[j] = __ldcs(r + 2);
            }
            q += (int64_t)kDenAhead * A3;
            if (on) {
#pragma unroll
                for (int j = 0; j < kDenAhead; ++j) count(x[j], y[j], z[j]);
            }
        }
        for (; k < n_own; ++k) {
            const double x = __ldcs(q), y = __ldcs(q + 1), z = __ldcs(q + 2);
            q += A3;
            if (on) count(x, y, z);
        }
        if (run) atomicAdd(target + run_voxel, run);
    }
    if (bad) *range_flag = 1;
    if (kSmem) {
        __syncthreads();
        for (int i = threadIdx.x; i < p.n_voxels; i += blockDim.x) {
            const unsigned int v = s_grid[i];
            if (v) atomicAdd(grid + i, v);
        }
    }
}

}  // namespace

extern "C" gd_status_t gd_density_frames(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                         const int32_t *h_nedges, uint32_t *d_grid,
                                         int32_t *d_range_flag, void *stream) {
    GD_REQUIRE(d_pos_frac && h_nedges && d_grid && d_range_flag, GD_ERR_INVALID,
               "gd_density_frames: null argument");
    GD_REQUIRE(n_frames >= 0 && n_atoms > 0, GD_ERR_INVALID, "gd_density_frames: bad size");
    DensityFrames p;
    int64_t voxels = 1;
    for (int k = 0; k < 3; ++k) {
        GD_REQUIRE(h_nedges[k] >= 2, GD_ERR_INVALID,
                   "gd_density_frames: need at least 2 edges per axis");
        p.ax[k].n_edges = h_nedges[k];
        p.ax[k].step = 1.0 / (double)(h_nedges[k] - 1);
        voxels *= h_nedges[k] - 1;
    }
    GD_REQUIRE(voxels < ((int64_t)1 << 31), GD_ERR_INVALID, "gd_density_frames: grid too large");
    if (n_frames == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    p.n_frames = n_frames;
    p.n_atoms = n_atoms;
    p.n_strips = (n_atoms + 31) / 32;
    // frames per warp: long enough to amortise the run bookkeeping, short enough to give every
    // SM several waves of warps
    int64_t chunk = 512;
    while (chunk > 32 && (int64_t)p.n_strips * gd_ceil_div(n_frames, chunk) < (int64_t)dev.sm_count * 64)
        chunk >>= 1;
    p.chunk = (int32_t)chunk;
    p.n_chunks = gd_ceil_div(n_frames, chunk);
    p.n_voxels = (int32_t)voxels;
    const int64_t items = (int64_t)p.n_strips * p.n_chunks;
    cudaStream_t st = (cudaStream_t)stream;
    if (voxels <= kDenMaxSmemVoxels) {
        const size_t smem = (size_t)voxels * sizeof(unsigned int);
        if (smem > 48 * 1024)
            GD_CUDA_CHECK(cudaFuncSetAttribute(k_density_runs<true>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = (int)((size_t)200 * 1024 / (smem + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        int64_t blocks = (int64_t)dev.sm_count * per_sm;
        if (blocks > gd_ceil_div(items, kDenWarps)) blocks = gd_ceil_div(items, kDenWarps);
        k_density_runs<true><<<(unsigned)blocks, kDenWarps * 32, smem, st>>>(d_pos_frac, p, d_grid,
                                                                            d_range_flag);
    } else {
        int64_t blocks = gd_ceil_div(items, kDenWarps);
        const int64_t cap = (int64_t)dev.sm_count * 64;
        if (blocks > cap) blocks = cap;
        k_density_runs<false><<<(unsigned)blocks, kDenWarps * 32, 0, st>>>(d_pos_frac, p, d_grid,
                                                                          d_range_flag);
    }
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_density(const double *d_pos_frac, int64_t n_atom_frames,
                                  const int32_t *h_nedges, uint32_t *d_grid, int32_t *d_range_flag,
                                  void *stream) {
    GD_REQUIRE(d_pos_frac && h_nedges && d_grid && d_range_flag, GD_ERR_INVALID,
               "gd_density: null argument");
    GD_REQUIRE(n_atom_frames >= 0, GD_ERR_INVALID, "gd_density: negative size");
    DensityParams p;
    for (int k = 0; k < 3; ++k) {
        GD_REQUIRE(h_nedges[k] >= 2, GD_ERR_INVALID, "gd_density: need at least 2 edges per axis");
        p.ax[k].n_edges = h_nedges[k];
        p.ax[k].step = 1.0 / (double)(h_nedges[k] - 1);
    }
    p.n = n_atom_frames;
    if (n_atom_frames == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    int64_t blocks = gd_ceil_div(n_atom_frames, 256);
    const int64_t cap = (int64_t)dev.sm_count * 16;
    if (blocks > cap) blocks = cap;
    k_density<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_pos_frac, p, d_grid,
                                                                 d_range_flag);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
