// K0 -- unwrapped displacement path.
//
// Replaces pymatgen Trajectory.to_displacements behind Trajectory.displacements
// (reference src/gemdat/trajectory.py:182-193), cumulative_displacements (:705-713: a
// sequential float64 np.cumsum over time), distances_from_base_position (:715-730: a
// Python loop over frames) and _lengths (:40-60).
//
// The reference wraps every frame-to-frame step with d - round(d) and sums the
// steps.  Mathematically cum[t] = (x[t] - x[0]) - W[t] with W[t] the running sum
// of the integers round(x[tau] - x[tau-1]); W is an exact integer prefix sum, so
// the time axis parallelises without changing any floating-point result by more
// than the rounding of two subtractions (the reference itself accumulates ~sqrt(T)
// roundings; agreement is ~1e-13 absolute, see tests).  Three passes over chunks of
// F frames: per-(chunk, atom) winding totals -> per-atom exclusive scan over chunks
// -> apply.
//
// Algorithmic traffic: 24 B read (count) + 24 B read + 24/48/8 B written (apply)
// per atom-frame, depending on the outputs requested.
#include <string.h>

#include "gd_common.cuh"
#include "gd_pbc.cuh"

namespace {

__host__ __device__ inline int64_t unwrap_chunk_frames(int64_t n_frames, int32_t n_atoms) {
    int64_t want = (n_frames * (int64_t)n_atoms) >> 17;
    int64_t f = 8;
    while (f < want && f < 512) f <<= 1;
    return f;
}

struct UGeom {
    int64_t n_frames;
    int32_t n_atoms;
    int64_t f;
    int64_t n_chunks;
};

struct Mat9 {
    double m[9];
    double g[9];  // metric tensor M M^T
};

__device__ __forceinline__ int winding(double cur, double prev) {
    // np.around == rint (half to even) on the rounded difference
    return (int)rint(__dsub_rn(cur, prev));
}

// pass 1: winding total of every (chunk, atom, component)
__global__ void __launch_bounds__(256)
    k_unwrap_count(const double *__restrict__ pos, UGeom g, const double *__restrict__ prev_row,
                   int32_t *__restrict__ work) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= g.n_chunks * g.n_atoms) return;
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
    const int64_t t0 = c * g.f, t1 = min(g.n_frames, t0 + g.f);
    const int64_t A3 = (int64_t)g.n_atoms * 3;
    double px, py, pz;
    int wx = 0, wy = 0, wz = 0;
    int64_t t = t0;
    if (t0 == 0) {
        if (prev_row) {
            px = prev_row[3 * a];
            py = prev_row[3 * a + 1];
            pz = prev_row[3 * a + 2];
        } else {  // d[0] = 0
            px = pos[3 * a];
            py = pos[3 * a + 1];
            pz = pos[3 * a + 2];
            t = 1;
        }
    } else {
        const double *q = pos + (t0 - 1) * A3 + 3 * a;
        px = q[0];
        py = q[1];
        pz = q[2];
    }
#pragma unroll 4
    for (; t < t1; ++t) {
        const double *q = pos + t * A3 + 3 * a;
        const double x = q[0], y = q[1], z = q[2];
        wx += winding(x, px);
        wy += winding(y, py);
        wz += winding(z, pz);
        px = x;
        py = y;
        pz = z;
    }
    int32_t *w = work + (c * g.n_atoms + a) * 3;
    w[0] = wx;
    w[1] = wy;
    w[2] = wz;
}

// pass 2: per (atom, component) exclusive scan over chunks (+ incoming winding)
__global__ void __launch_bounds__(128)
    k_unwrap_scan(UGeom g, const int32_t *__restrict__ base_winding, int32_t *__restrict__ work,
                  int32_t *__restrict__ winding_out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // a*3 + comp
    const int64_t n = (int64_t)g.n_atoms * 3;
    if (k >= n) return;
    int32_t run = base_winding ? base_winding[k] : 0;
#pragma unroll 8
    for (int64_t c = 0; c < g.n_chunks; ++c) {
        const int32_t v = work[c * n + k];
        work[c * n + k] = run;
        run += v;
    }
    if (winding_out) winding_out[k] = run;
}

// pass 3: apply
__global__ void __launch_bounds__(256)
    k_unwrap_apply(const double *__restrict__ pos, UGeom g, const double *__restrict__ prev_row,
                   const double *__restrict__ base_pos, const int32_t *__restrict__ work,
                   const int32_t *__restrict__ extra_winding, const __grid_constant__ Mat9 L,
                   double *__restrict__ cum_frac, double *__restrict__ cum_cart,
                   double *__restrict__ dist, double *__restrict__ speed) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= g.n_chunks * g.n_atoms) return;
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
    const int64_t t0 = c * g.f, t1 = min(g.n_frames, t0 + g.f);
    const int64_t A3 = (int64_t)g.n_atoms * 3;
    const double *b = base_pos ? base_pos + 3 * a : pos + 3 * a;
    const double bx = b[0], by = b[1], bz = b[2];
    const int32_t *w = work + (c * g.n_atoms + a) * 3;
    int wx = w[0], wy = w[1], wz = w[2];
    if (extra_winding) {  // windings of the preceding shards (work was scanned from zero)
        wx += extra_winding[3 * a];
        wy += extra_winding[3 * a + 1];
        wz += extra_winding[3 * a + 2];
    }
    double px, py, pz;
    bool first_is_zero = false;
    if (t0 == 0) {
        if (prev_row) {
            px = prev_row[3 * a];
            py = prev_row[3 * a + 1];
            pz = prev_row[3 * a + 2];
        } else {
            px = bx;
            py = by;
            pz = bz;
            first_is_zero = true;
        }
    } else {
        const double *q = pos + (t0 - 1) * A3 + 3 * a;
        px = q[0];
        py = q[1];
        pz = q[2];
    }
    // metric-tensor length of a fractional displacement (_lengths, trajectory.py:40-60)
    auto length = [&](double cx, double cy, double cz) {
        const double gx = fma(cz, L.g[6], fma(cy, L.g[3], cx * L.g[0]));
        const double gy = fma(cz, L.g[7], fma(cy, L.g[4], cx * L.g[1]));
        const double gz = fma(cz, L.g[8], fma(cy, L.g[5], cx * L.g[2]));
        return sqrt(fmax(fma(gz, cz, fma(gy, cy, gx * cx)), 0.0));
    };
    // speed = np.diff(distance, prepend=0) (metrics.py:46-47): distance of the frame before
    // this segment, from the position before it and the winding up to it
    double pd = 0.0;
    if (speed && !first_is_zero)
        pd = length(__dsub_rn(__dsub_rn(px, bx), (double)wx), __dsub_rn(__dsub_rn(py, by), (double)wy),
                    __dsub_rn(__dsub_rn(pz, bz), (double)wz));
#pragma unroll 2
    for (int64_t t = t0; t < t1; ++t) {
        const double *q = pos + t * A3 + 3 * a;
        const double x = q[0], y = q[1], z = q[2];
        if (!(first_is_zero && t == 0)) {
            wx += winding(x, px);
            wy += winding(y, py);
            wz += winding(z, pz);
        }
        px = x;
        py = y;
        pz = z;
        const double cx = __dsub_rn(__dsub_rn(x, bx), (double)wx);
        const double cy = __dsub_rn(__dsub_rn(y, by), (double)wy);
        const double cz = __dsub_rn(__dsub_rn(z, bz), (double)wz);
        const int64_t o = t * A3 + 3 * a;
        if (cum_frac) {
            cum_frac[o] = cx;
            cum_frac[o + 1] = cy;
            cum_frac[o + 2] = cz;
        }
        if (cum_cart) {
            cum_cart[o] = fma(cz, L.m[6], fma(cy, L.m[3], cx * L.m[0]));
            cum_cart[o + 1] = fma(cz, L.m[7], fma(cy, L.m[4], cx * L.m[1]));
            cum_cart[o + 2] = fma(cz, L.m[8], fma(cy, L.m[5], cx * L.m[2]));
        }
        if (dist || speed) {
            const double d = length(cx, cy, cz);
            if (dist) dist[(int64_t)a * g.n_frames + t] = d;
            if (speed) speed[t * (int64_t)g.n_atoms + a] = d - pd;  // time-major: lane = atom
            pd = d;
        }
    }
}

// ---------------------------------------------------------------- minimum-image distances
// (pymatgen Lattice.get_all_distances: LLL-reduced basis, 27 images; see gd_pbc.cuh)

__global__ void __launch_bounds__(256)
    k_pbc_distances(const double *__restrict__ f1, int32_t n1, const double *__restrict__ f2,
                    int32_t n2, const __grid_constant__ GdPbc P, double *__restrict__ out) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)n1 * n2) return;
    const int32_t i = (int32_t)(idx / n2), j = (int32_t)(idx - (int64_t)i * n2);
    double a[3], b[3];
    gd_pbc_cart(P, f1[3 * i], f1[3 * i + 1], f1[3 * i + 2], a);
    gd_pbc_cart(P, f2[3 * j], f2[3 * j + 1], f2[3 * j + 2], b);
    out[idx] = sqrt(gd_pbc_d2(P, a, b));
}

Mat9 make_mat(const double *h_lattice) {
    Mat9 L;
    memcpy(L.m, h_lattice, sizeof(L.m));
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += h_lattice[3 * i + k] * h_lattice[3 * j + k];
            L.g[3 * i + j] = s;
        }
    return L;
}

}  // namespace

extern "C" int64_t gd_unwrap_chunks(int64_t n_frames, int32_t n_atoms) {
    if (n_frames <= 0 || n_atoms <= 0) return 1;
    return gd_ceil_div(n_frames, unwrap_chunk_frames(n_frames, n_atoms));
}

namespace {

UGeom unwrap_geometry(int64_t n_frames, int32_t n_atoms) {
    UGeom g;
    g.n_frames = n_frames;
    g.n_atoms = n_atoms;
    g.f = unwrap_chunk_frames(n_frames, n_atoms);
    g.n_chunks = gd_ceil_div(n_frames, g.f);
    return g;
}

gd_status_t launch_windings(const double *d_pos_frac, const UGeom &g, const double *d_prev_row,
                            const int32_t *d_base_winding, int32_t *d_work,
                            int32_t *d_winding_out, cudaStream_t st) {
    const int64_t items = g.n_chunks * g.n_atoms;
    k_unwrap_count<<<(unsigned)gd_ceil_div(items, 256), 256, 0, st>>>(d_pos_frac, g, d_prev_row,
                                                                    d_work);
    GD_LAUNCH_CHECK();
    k_unwrap_scan<<<(unsigned)gd_ceil_div((int64_t)g.n_atoms * 3, 128), 128, 0, st>>>(
        g, d_base_winding, d_work, d_winding_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

gd_status_t launch_apply(const double *d_pos_frac, const UGeom &g, const double *h_lattice,
                         const double *d_prev_row, const double *d_base_pos, const int32_t *d_work,
                         const int32_t *d_extra_winding, double *d_cum_frac, double *d_cum_cart,
                         double *d_dist, double *d_speed, cudaStream_t st) {
    if (!(d_cum_frac || d_cum_cart || d_dist || d_speed)) return GD_OK;
    const Mat9 L = make_mat(h_lattice);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_unwrap_apply<<<(unsigned)gd_ceil_div(items, 256), 256, 0, st>>>(
        d_pos_frac, g, d_prev_row, d_base_pos, d_work, d_extra_winding, L, d_cum_frac, d_cum_cart,
        d_dist, d_speed);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

}  // namespace

extern "C" gd_status_t gd_unwrap(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                 const double *h_lattice, const double *d_prev_row,
                                 const int32_t *d_base_winding, const double *d_base_pos,
                                 int32_t *d_work, double *d_cum_frac, double *d_cum_cart,
                                 double *d_dist, double *d_speed, int32_t *d_winding_out,
                                 void *stream) {
    GD_REQUIRE(d_pos_frac && h_lattice && d_work, GD_ERR_INVALID, "gd_unwrap: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_unwrap: empty input");
    cudaStream_t st = (cudaStream_t)stream;
    const UGeom g = unwrap_geometry(n_frames, n_atoms);
    gd_status_t rc = launch_windings(d_pos_frac, g, d_prev_row, d_base_winding, d_work,
                                     d_winding_out, st);
    if (rc != GD_OK) return rc;
    return launch_apply(d_pos_frac, g, h_lattice, d_prev_row, d_base_pos, d_work, nullptr,
                        d_cum_frac, d_cum_cart, d_dist, d_speed, st);
}

extern "C" gd_status_t gd_unwrap_windings(const double *d_pos_frac, int64_t n_frames,
                                          int32_t n_atoms, const double *d_prev_row,
                                          int32_t *d_work, int32_t *d_winding_out, void *stream) {
    GD_REQUIRE(d_pos_frac && d_work && d_winding_out, GD_ERR_INVALID,
               "gd_unwrap_windings: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_unwrap_windings: empty input");
    return launch_windings(d_pos_frac, unwrap_geometry(n_frames, n_atoms), d_prev_row, nullptr,
                           d_work, d_winding_out, (cudaStream_t)stream);
}

extern "C" gd_status_t gd_unwrap_apply(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                       const double *h_lattice, const double *d_prev_row,
                                       const int32_t *d_base_winding, const double *d_base_pos,
                                       const int32_t *d_work, double *d_cum_frac,
                                       double *d_cum_cart, double *d_dist, double *d_speed,
                                       void *stream) {
    GD_REQUIRE(d_pos_frac && h_lattice && d_work, GD_ERR_INVALID, "gd_unwrap_apply: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_unwrap_apply: empty input");
    return launch_apply(d_pos_frac, unwrap_geometry(n_frames, n_atoms), h_lattice, d_prev_row,
                        d_base_pos, d_work, d_base_winding, d_cum_frac, d_cum_cart, d_dist, d_speed,
                        (cudaStream_t)stream);
}

extern "C" gd_status_t gd_lll_reduce_host(const double *h_lattice, double *h_lll, double *h_mapping) {
    GD_REQUIRE(h_lattice && h_lll && h_mapping, GD_ERR_INVALID, "gd_lll_reduce_host: null argument");
    gd_lll_reduce(h_lattice, h_lll, h_mapping);
    return GD_OK;
}

extern "C" gd_status_t gd_pbc_distances(const double *h_lattice, const double *d_frac1, int32_t n1,
                                        const double *d_frac2, int32_t n2, double *d_out,
                                        void *stream) {
    GD_REQUIRE(h_lattice && d_frac1 && d_frac2 && d_out && n1 > 0 && n2 > 0, GD_ERR_INVALID,
               "gd_pbc_distances: bad argument");
    const GdPbc P = gd_pbc_build(h_lattice);
    const int64_t n = (int64_t)n1 * n2;
    k_pbc_distances<<<(unsigned)gd_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
        d_frac1, n1, d_frac2, n2, P, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
