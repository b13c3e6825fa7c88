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
#include <stdlib.h>
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

// pass 2: per (atom, component) exclusive scan over chunks (+ incoming winding); one warp per
// column, 32 chunks per step
__global__ void __launch_bounds__(128)
    k_unwrap_scan(UGeom g, const int32_t *__restrict__ base_winding, int32_t *__restrict__ work,
                  int32_t *__restrict__ winding_out) {
    const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // a*3 + comp
    const int lane = threadIdx.x & 31;
    const int64_t n = (int64_t)g.n_atoms * 3;
    if (k >= n) return;
    int32_t run = base_winding ? base_winding[k] : 0;
    for (int64_t c0 = 0; c0 < g.n_chunks; c0 += 32) {
        const int64_t c = c0 + lane;
        const bool on = c < g.n_chunks;
        const int32_t v = on ? work[c * n + k] : 0;
        int32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        if (on) work[c * n + k] = run + incl - v;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (winding_out && lane == 0) winding_out[k] = run;
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

// ---------------------------------------------------------------- fused: one pass over HBM
//
// The three passes above stream the input twice (winding totals, then apply).  The fused kernel
// reads it from HBM once: a CTA takes a tile of kTileF frames x 32 atoms (lane = atom), sums
// the tile's windings (pass 1), learns the winding before the tile by a DECOUPLED LOOK-BACK
// along the frame axis -- every tile publishes its own total, then its inclusive prefix, as
// self-describing 32-bit words (value << 2 | state), and a tile walks back over its
// predecessors' words until it meets a prefix -- and applies it while the tile is still in L2
// (pass 2 re-reads the 96 KB it has just read).  Tiles are handed out through a ticket counter
// so that a tile only ever waits for tiles that are already running.  The prefix is an exact
// integer, so the outputs are bit-identical to the three-pass route.
constexpr int kTileWarps = 4;
constexpr int kTileWarpF = 32;                         // frames per warp
constexpr int kTileF = kTileWarps * kTileWarpF;        // frames per tile
constexpr uint32_t kStAggregate = 1u, kStPrefix = 2u;  // 0 = not published yet

struct FusedGeom {
    int64_t n_frames;
    int32_t n_atoms;
    int32_t n_strips;
    int64_t n_tchunks;  // tiles along the frame axis
    int32_t prefetch;   // rows are 16-byte aligned: bulk L2 prefetch of the tile
    int32_t dist_tile;  // the dynamic shared memory holds a 32 x 33 tile per warp: dist leaves transposed
};

__device__ __forceinline__ uint32_t st_pack(int32_t v, uint32_t state) {
    return ((uint32_t)v << 2) | state;
}
__device__ __forceinline__ int32_t st_value(uint32_t w) { return (int32_t)w >> 2; }

__global__ void __launch_bounds__(kTileWarps * 32)
    k_unwrap_fused(const double *__restrict__ pos, FusedGeom g, const double *__restrict__ prev_row,
                   const int32_t *__restrict__ base_winding, const double *__restrict__ base_pos,
                   uint32_t *__restrict__ ticket, uint32_t *__restrict__ status,
                   const __grid_constant__ Mat9 L, double *__restrict__ cum_frac,
                   double *__restrict__ cum_cart, double *__restrict__ dist,
                   double *__restrict__ speed, int32_t *__restrict__ winding_out) {
    extern __shared__ double s_dyn[];  // requested to cap the CTAs per SM; holds the dist tiles
    __shared__ uint32_t s_tile;
    __shared__ int32_t s_agg[kTileWarps][3][32];
    __shared__ int32_t s_base[3][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t tc = tile / g.n_strips;                    // position along the frame axis
    const int32_t strip = (int32_t)(tile - tc * g.n_strips);
    const int32_t a = min(strip * 32 + lane, g.n_atoms - 1);  // lanes past the end shadow the last atom
    const bool on = strip * 32 + lane < g.n_atoms;
    const int64_t A3 = (int64_t)g.n_atoms * 3;
    const int64_t t0 = tc * kTileF + (int64_t)warp * kTileWarpF;
    const int64_t t1 = min(g.n_frames, t0 + kTileWarpF);

    // The tile's rows are asked for at once (bulk prefetch into L2, one row per thread), so the
    // bytes in flight from HBM are the whole tile and not what a lane's registers can hold; the
    // loads of pass 1 then mostly find their rows in L2.
    if (g.prefetch) {
        const int64_t t = tc * kTileF + threadIdx.x;
        const int r = min(32, g.n_atoms - strip * 32);
        if (t < g.n_frames) gd_prefetch_l2(pos + t * A3 + (int64_t)strip * 96, (uint32_t)(r * 24));
    }
    // previous position of this warp's first frame
    double px = 0.0, py = 0.0, pz = 0.0;
    bool first_is_zero = false;  // global frame 0 without a halo row: d[0] = 0
    if (t0 < g.n_frames) {
        const double *q = t0 > 0 ? pos + (t0 - 1) * A3 + 3 * (int64_t)a
                                 : (prev_row ? prev_row + 3 * (int64_t)a : nullptr);
        if (q) {
            px = q[0];
            py = q[1];
            pz = q[2];
        } else {
            first_is_zero = true;
        }
    }
    // ---- pass 1: windings of this warp's frames
    int wx = 0, wy = 0, wz = 0;
    {
        double ax = px, ay = py, az = pz;
        const double *q = pos + t0 * A3 + 3 * (int64_t)a;
        constexpr int kAhead = 8;
        int64_t t = t0;
        for (; t + kAhead <= t1; t += kAhead) {
            double x[kAhead], y[kAhead], z[kAhead];
#pragma unroll
            for (int j = 0; j < kAhead; ++j) {
                x[j] = q[j * A3];
                y[j] = q[j * A3 + 1];
                z[j] = q[j * A3 + 2];
            }
            q += kAhead * A3;
#pragma unroll
            for (int j = 0; j < kAhead; ++j) {
                if (!(first_is_zero && t + j == 0)) {
                    wx += winding(x[j], ax);
                    wy += winding(y[j], ay);
                    wz += winding(z[j], az);
                }
                ax = x[j];
                ay = y[j];
                az = z[j];
            }
        }
        for (; t < t1; ++t) {
            const double x = q[0], y = q[1], z = q[2];
            q += A3;
            if (!(first_is_zero && t == 0)) {
                wx += winding(x, ax);
                wy += winding(y, ay);
                wz += winding(z, az);
            }
            ax = x;
            ay = y;
            az = z;
        }
    }
    s_agg[warp][0][lane] = wx;
    s_agg[warp][1][lane] = wy;
    s_agg[warp][2][lane] = wz;
    __syncthreads();
    // ---- look-back (warp 0; every lane follows its own three columns)
    if (warp == 0) {
        int32_t tot[3], run[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            tot[c] = 0;
#pragma unroll
            for (int w = 0; w < kTileWarps; ++w) tot[c] += s_agg[w][c][lane];
        }
        volatile uint32_t *mine = status + (tile * 3) * 32;
        if (tc > 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) mine[c * 32 + lane] = st_pack(tot[c], kStAggregate);
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            int32_t acc = 0;
            int64_t j = tc - 1;
            for (; j >= 0; --j) {
                volatile const uint32_t *theirs = status + ((j * g.n_strips + strip) * 3 + c) * 32 + lane;
                uint32_t w;
                while (((w = *theirs) & 3u) == 0u) {
                }
                acc += st_value(w);
                if ((w & 3u) == kStPrefix) break;
            }
            if (j < 0 && base_winding) acc += base_winding[3 * a + c];
            run[c] = acc;
            mine[c * 32 + lane] = st_pack(acc + tot[c], kStPrefix);
            s_base[c][lane] = acc;
        }
        if (winding_out && on && tc == g.n_tchunks - 1) {
#pragma unroll
            for (int c = 0; c < 3; ++c) winding_out[3 * a + c] = run[c] + tot[c];
        }
    }
    __syncthreads();
    if (!(cum_frac || cum_cart || dist || speed) || t0 >= g.n_frames) return;
    // ---- pass 2: apply (the tile is in L2)
    wx = s_base[0][lane];
    wy = s_base[1][lane];
    wz = s_base[2][lane];
    for (int w = 0; w < warp; ++w) {
        wx += s_agg[w][0][lane];
        wy += s_agg[w][1][lane];
        wz += s_agg[w][2][lane];
    }
    const double *b = base_pos ? base_pos + 3 * (int64_t)a : pos + 3 * (int64_t)a;
    const double bx = b[0], by = b[1], bz = b[2];
    auto length = [&](double cx, double cy, double cz) {
        const double gx = fma(cz, L.g[6], fma(cy, L.g[3], cx * L.g[0]));
        const double gy = fma(cz, L.g[7], fma(cy, L.g[4], cx * L.g[1]));
        const double gz = fma(cz, L.g[8], fma(cy, L.g[5], cx * L.g[2]));
        return sqrt(fmax(fma(gz, cz, fma(gy, cy, gx * cx)), 0.0));
    };
    double pd = 0.0;
    if (speed && !first_is_zero)
        pd = length(__dsub_rn(__dsub_rn(px, bx), (double)wx), __dsub_rn(__dsub_rn(py, by), (double)wy),
                    __dsub_rn(__dsub_rn(pz, bz), (double)wz));
    const double *q = pos + t0 * A3 + 3 * (int64_t)a;
    constexpr int kAhead2 = 4;
    double *dist_tile = (dist && g.dist_tile) ? s_dyn + warp * (32 * 33) : nullptr;
    auto apply = [&](double x, double y, double z, int64_t t) {
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
        const int64_t o = t * A3 + 3 * (int64_t)a;
        if (on) {
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
        }
        if (dist || speed) {
            const double d = length(cx, cy, cz);
            if (dist) {
                // dist is atom-major [A, T]: with lane = atom a direct store touches 32 rows per
                // instruction.  The warp's 32 x 32 values go through shared memory instead and
                // leave as 256-byte runs of one atom's row (below).
                if (dist_tile)
                    dist_tile[(int)(t - t0) * 33 + lane] = d;
                else if (on)
                    dist[(int64_t)a * g.n_frames + t] = d;
            }
            if (speed && on) speed[t * (int64_t)g.n_atoms + a] = d - pd;
            pd = d;
        }
    };
    int64_t t = t0;
    for (; t + kAhead2 <= t1; t += kAhead2) {
        double x[kAhead2], y[kAhead2], z[kAhead2];
#pragma unroll
        for (int j = 0; j < kAhead2; ++j) {
            x[j] = q[j * A3];
            y[j] = q[j * A3 + 1];
            z[j] = q[j * A3 + 2];
        }
        q += kAhead2 * A3;
#pragma unroll
        for (int j = 0; j < kAhead2; ++j) apply(x[j], y[j], z[j], t + j);
    }
    for (; t < t1; ++t) {
        apply(q[0], q[1], q[2], t);
        q += A3;
    }
    if (dist_tile) {
        __syncwarp();
        const int n_t = (int)(t1 - t0);
        const int n_a = min(32, g.n_atoms - strip * 32);
        double *row = dist + (int64_t)strip * 32 * g.n_frames + t0 + lane;
        for (int i = 0; i < n_a; ++i, row += g.n_frames)
            if (lane < n_t) __stcs(row, dist_tile[lane * 33 + i]);
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

namespace {
FusedGeom fused_geometry(int64_t n_frames, int32_t n_atoms) {
    FusedGeom g;
    g.n_frames = n_frames;
    g.n_atoms = n_atoms;
    g.n_strips = (n_atoms + 31) / 32;
    g.n_tchunks = gd_ceil_div(n_frames, kTileF);
    g.prefetch = 0;
    g.dist_tile = 0;
    return g;
}
// int32 words of d_work the fused kernel needs: ticket (padded) + 96 status words per tile
int64_t fused_work_words(const FusedGeom &g) { return 32 + g.n_tchunks * g.n_strips * 96; }
}  // namespace

extern "C" int64_t gd_unwrap_chunks(int64_t n_frames, int32_t n_atoms) {
    if (n_frames <= 0 || n_atoms <= 0) return 1;
    // d_work is sized n_chunks * n_atoms * 3 int32: enough for the three-pass route (per-chunk
    // windings) and for the fused route (ticket + per-tile status words)
    const int64_t three_pass = gd_ceil_div(n_frames, unwrap_chunk_frames(n_frames, n_atoms));
    const int64_t fused = gd_ceil_div(fused_work_words(fused_geometry(n_frames, n_atoms)), (int64_t)n_atoms * 3);
    return three_pass > fused ? three_pass : fused;
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
    k_unwrap_scan<<<(unsigned)gd_ceil_div((int64_t)g.n_atoms * 3 * 32, 128), 128, 0, st>>>(
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
    // GD_UNWRAP_FUSED=0: the three-pass route (A/B; the outputs are bit-identical)
    const char *env = getenv("GD_UNWRAP_FUSED");
    const bool fused = !(env && atoi(env) == 0) && (d_cum_frac || d_cum_cart || d_dist || d_speed) &&
                       n_frames < ((int64_t)1 << 29);
    if (fused) {
        FusedGeom fg = fused_geometry(n_frames, n_atoms);
        // GD_UNWRAP_PREFETCH=0 switches the L2 prefetch off (A/B)
        const char *pf = getenv("GD_UNWRAP_PREFETCH");
        fg.prefetch = (n_atoms % 2 == 0) && (reinterpret_cast<uintptr_t>(d_pos_frac) & 15) == 0 &&
                      !(pf && atoi(pf) == 0);
        const int64_t tiles = fg.n_tchunks * fg.n_strips;
        GD_REQUIRE(tiles < ((int64_t)1 << 31), GD_ERR_INVALID, "gd_unwrap: too many tiles");
        uint32_t *words = reinterpret_cast<uint32_t *>(d_work);
        GD_CUDA_CHECK(cudaMemsetAsync(words, 0, (size_t)fused_work_words(fg) * 4, st));
        const Mat9 L = make_mat(h_lattice);
        // resident tiles must stay in L2 between the two passes (96 KB each): the dynamic
        // shared memory request only caps the CTAs per SM (GD_UNWRAP_CTAS_PER_SM, default 5:
        // 148 x 5 x 96 KB = 71 MB of the 126 MB L2; measured 3 / 4 / 5 / 6 / 8 per SM: 3749 /
        // 5026 / 5301 / 4745 / 4365 GB/s, profiles/r2h_k0_sweep.txt)
        // (with only the 8-byte outputs -- distance, speed -- 6 per SM is the optimum:
        // 4 / 5 / 6 / 7 per SM: 3.61 / 2.99 / 2.74 / 2.74 ms per 4e8 atom-frames for the speed, and
        // 6 is the most the distance tile fits, profiles/r2zf_k0_persm.txt)
        int per_sm = (d_cum_frac || d_cum_cart) ? 5 : 6;
        if (const char *e = getenv("GD_UNWRAP_CTAS_PER_SM")) per_sm = atoi(e) > 0 ? atoi(e) : per_sm;
        const GdDeviceInfo &dev = gd_device_info();
        GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
        int smem = dev.max_smem_optin / per_sm - 2048;
        if (smem < 0) smem = 0;
        fg.dist_tile = smem >= (int)(kTileWarps * 32 * 33 * sizeof(double));
        GD_CUDA_CHECK(cudaFuncSetAttribute(k_unwrap_fused, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           smem));
        k_unwrap_fused<<<(unsigned)tiles, kTileWarps * 32, (size_t)smem, st>>>(
            d_pos_frac, fg, d_prev_row, d_base_winding, d_base_pos, words, words + 32, L, d_cum_frac,
            d_cum_cart, d_dist, d_speed, d_winding_out);
        GD_LAUNCH_CHECK();
        return GD_OK;
    }
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
