// K5 -- density histogram of fractional positions.
//
// Replaces trajectory_to_volume (reference src/gemdat/volume.py:449-509): three
// np.digitize calls against np.linspace(0, 1, n)[1:] edges, then an
// np.unique(axis=0, return_counts=True) sort of all T*A rows.
//
// Bin index: the number of edges e_j = fl(j * fl(1/(n-1))), j = 1..n-1 (last one
// exactly 1.0), that are <= x -- evaluated exactly as np.digitize would, by
// correcting the guess floor(x*(n-1)) against the neighbouring edges.
#include "gd_common.cuh"

namespace {

struct Axis {
    int n_edges;   // n
    double step;   // fl(1 / (n - 1))
};

__device__ __forceinline__ double edge(const Axis &ax, int j) {
    return j >= ax.n_edges - 1 ? 1.0 : __dmul_rn((double)j, ax.step);
}

__device__ __forceinline__ int bin_of(const Axis &ax, double x) {
    const int last = ax.n_edges - 2;  // highest voxel index
    if (last <= 0) return 0;
    int g = (int)(x * (double)(ax.n_edges - 1));
    g = min(max(g, 0), last);
    while (g < last && edge(ax, g + 1) <= x) ++g;
    while (g > 0 && edge(ax, g) > x) --g;
    return g;
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

}  // namespace

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
