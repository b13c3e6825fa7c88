// Radial distribution histograms: reference src/gemdat/rdf.py:111-268.
//
//   gd_rdf_classes  per (frame, floating atom): the transition state it is in
//                   (_get_states_array, rdf.py:53-61, with _uniqify_labels' off-by-one kept)
//                   as an index into the caller's table of state names, plus for every state
//                   the first (frame, code) at which it occurs (the order of the reference's
//                   result dict)
//   gd_rdf_hist     per frame: minimum-image distances source atoms x target atoms
//                   (Lattice.get_all_distances: both wrapped to [0,1), 27 images, same
//                   operation order as gd_pbc_distances), np.digitize(right=True) or
//                   np.histogram binning against the caller's exact edges -- on the squared
//                   distance, through per-edge thresholds that reproduce the comparison with
//                   the rounded square root exactly -- counted per (class of the source,
//                   group of the target, bin)
//
// One CTA per (frame, slice of the sources); the frame's targets sit in shared memory wrapped
// to [0,1); one warp per source, lanes over targets, a warp-private shared-memory histogram
// flushed to the global uint64 table per (source, target tile).  Distances come from gd_pbc.cuh
// (targets staged as Cartesian coordinates of the LLL-reduced, wrapped point): 27 image sums for a
// general cell, one candidate per axis for an axis-aligned one.  Pairs beyond the last edge -- the
// bulk when the cell is larger than max_dist -- are counted in registers; the others find their
// bin from an fp32 guess corrected against exact per-edge thresholds on d2.
#include <vector>

#include "gd_common.cuh"
#include "gd_pbc.cuh"

// Tuning knob for A/B builds (tools/rdf_variants.sh).
#ifdef GD_RDF_MIN_CTAS  // resident CTAs per SM the register allocation must allow
#define GD_RDF_BOUNDS __launch_bounds__(8 * 32, GD_RDF_MIN_CTAS)
#else
// (per-frame cells live in shared memory and the compiler keeps all 27 image vectors in registers,
// 245 of them, one CTA per SM.  Capping them is worse: general cell, 2000 x 1008 x 1007 pairs:
// 45.3 ms at 1 CTA/SM, 79.7 at 2 (128 registers, spills), 116.7 at 3; profiles/r2zs_rdf_frames_time.txt)
#define GD_RDF_BOUNDS __launch_bounds__(8 * 32, kPerFrame ? GD_RDF_FRAMES_MIN_CTAS : 0)
#endif
#ifndef GD_RDF_FRAMES_MIN_CTAS
#define GD_RDF_FRAMES_MIN_CTAS 1
#endif

namespace {

constexpr int kWarps = 8;

struct RdfParams {
    const double *pos;  // [T, N, 3] fractional
    int64_t n_frames;
    int32_t n_atoms;
    const int32_t *idx1;  // [n1] source atoms
    int32_t n1;
    const int32_t *cls;  // [T, n1] class of the source in that frame, <0 = skip; NULL = 0
    const int32_t *idx2;   // [n2] target atoms
    const int32_t *group;  // [n2] group of the target; NULL = 0
    int32_t n2;
    int32_t n_group;
    const double *edges;  // [n_edges]
    int32_t n_edges;
    int32_t mode;  // 0: digitize(right=True) -> #edges < d;  1: histogram -> #edges <= d
    int32_t skip_same;
    int32_t tile;  // targets per shared-memory tile
    double inv_res;
    GdPbc pbc;
    const GdPbc *pbc_frames;   // kPerFrame: one cell per frame (the lattice changes along the run)
    unsigned long long *hist;  // [n_class, n_group, n_edges + 1]; kPerFrame: [n_frames, n_group, n_edges + 1]
};

// Smallest x >= 0 whose correctly rounded square root is >= e (strict: > e).  sqrt is monotonic,
// so "e <= fl(sqrt(d2))" is the same test as "x <= d2": the bins are found on d2 itself and the
// pair loop needs no fp64 square root.  About two doubles share one rounded root, so the walk
// from fl(e*e) is a handful of steps.
__device__ double sqrt_threshold(double e, bool strict) {
    if (!(e > 0.0)) return (strict || e < 0.0) ? (e < 0.0 ? 0.0 : __longlong_as_double(1)) : 0.0;
    double x = e * e;
    auto ok = [&](double v) { return strict ? sqrt(v) > e : sqrt(v) >= e; };
    while (x > 0.0 && ok(x)) x = __longlong_as_double(__double_as_longlong(x) - 1);
    while (!ok(x)) x = __longlong_as_double(__double_as_longlong(x) + 1);
    return x;
}

constexpr int kOverGroups = 4;  // groups whose beyond-the-last-edge pairs are counted in registers

// kOneGroup: all targets count in one table row (species RDF): no group look-ups, one register
// counter.  When the edges are k * resolution (np.arange from 0) the fp32 bin guess is off by at
// most one and a single comparison with each neighbouring threshold is exact; other edge lists
// (checked once per CTA) keep the search loops.
// kPerFrame: the cell of frame t is pbc_frames[t] (copied to shared memory once per frame) and the
// counts of frame t go to their own table hist[t] -- the caller weights them with the frame's volume
// (rdf.py:248-261); d_class then only marks sources to skip (< 0).
template <bool kOneGroup, bool kPerFrame>
__global__ void GD_RDF_BOUNDS k_rdf_hist(const __grid_constant__ RdfParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int nb = p.n_edges + 1;
    const int hist_len = p.n_group * nb;
    // targets of the tile, structure of arrays (conflict-free 8-byte reads, no index arithmetic)
    double *s_x = reinterpret_cast<double *>(smem);                  // [tile]
    double *s_y = s_x + p.tile;
    double *s_z = s_y + p.tile;
    double *s_th = s_z + p.tile;                                     // [n_edges + 2] thresholds on d2
    int32_t *s_ag = reinterpret_cast<int32_t *>(s_th + p.n_edges + 2);   // [tile] atom | group << 24
    uint32_t *s_hist = reinterpret_cast<uint32_t *>(s_ag + p.tile);  // [kWarps][hist_len]

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv_res = (float)p.inv_res;
    uint32_t *my_hist = s_hist + (size_t)warp * hist_len;
    // bin index = number of edges passed.  s_th[k + 1] = smallest d2 that passes edge k: mode 0
    // (np.digitize right=True) edge < d; mode 1 (np.histogram) edge <= d, except the LAST edge,
    // which np.histogram's closed last bin makes strict as well.  s_th[0] = -inf and
    // s_th[n_edges + 1] = +inf close the table, so the search below needs no bounds checks.
    for (int k = threadIdx.x; k < p.n_edges + 2; k += blockDim.x) {
        double th;
        if (k == 0)
            th = -1e300;
        else if (k == p.n_edges + 1)
            th = 1e300;
        else
            th = sqrt_threshold(p.edges[k - 1], p.mode == 0 || k == p.n_edges);
        s_th[k] = th;
    }
    for (int k = threadIdx.x; k < kWarps * hist_len; k += blockDim.x) s_hist[k] = 0;
    __shared__ int s_irregular;
    __shared__ GdPbc s_pbc;
    if (threadIdx.x == 0) s_irregular = 0;
    __syncthreads();
    for (int k = threadIdx.x; k < p.n_edges; k += blockDim.x)
        if (!(fabs(p.edges[k] * p.inv_res - (double)k) <= 1e-6)) s_irregular = 1;
    __syncthreads();
    const bool uniform = s_irregular == 0;
    const double th_over = s_th[p.n_edges];  // at or above: beyond the last edge
    const bool reg_over = kOneGroup || p.n_group <= kOverGroups;

    for (int64_t t = blockIdx.x; t < p.n_frames; t += gridDim.x) {
        const double *frame = p.pos + t * (int64_t)p.n_atoms * 3;
        if (kPerFrame) {
            __syncthreads();  // the frame before is done with s_pbc
            const int *src = reinterpret_cast<const int *>(p.pbc_frames + t);
            int *dst = reinterpret_cast<int *>(&s_pbc);
            for (int k = threadIdx.x; k < (int)(sizeof(GdPbc) / sizeof(int)); k += blockDim.x) dst[k] = src[k];
            // (the __syncthreads at the top of the tile loop publishes it)
        }
        const GdPbc &P = kPerFrame ? s_pbc : p.pbc;
        for (int j0 = 0; j0 < p.n2; j0 += p.tile) {
            const int nt = min(p.tile, p.n2 - j0);
            __syncthreads();
            for (int k = threadIdx.x; k < nt; k += blockDim.x) {
                const int32_t atom = p.idx2[j0 + k];
                const int32_t grp = (!kOneGroup && p.group) ? p.group[j0 + k] : 0;
                s_ag[k] = atom | (grp << 24);
                const double *q = frame + 3 * (int64_t)atom;
                double c[3];
                gd_pbc_cart(P, q[0], q[1], q[2], c);
                s_x[k] = c[0];
                s_y[k] = c[1];
                s_z[k] = c[2];
            }
            __syncthreads();
            for (int i = blockIdx.y * kWarps + warp; i < p.n1; i += gridDim.y * kWarps) {
                const int32_t cls = p.cls ? p.cls[t * p.n1 + i] : 0;
                if (cls < 0) continue;
                // a pair of an atom with itself is recognised by the atom index; -1 never matches
                const int32_t self = p.skip_same ? p.idx1[i] : -1;
                const double *q = frame + 3 * (int64_t)p.idx1[i];
                double a[3];
                gd_pbc_cart(P, q[0], q[1], q[2], a);
                // pairs beyond the last edge are the bulk (the cell is much larger than max_dist):
                // they are counted in registers, per group, and added once per (source, tile)
                uint32_t over[kOverGroups] = {0, 0, 0, 0};
                auto pair = [&](int k) {
                    const int32_t ag = s_ag[k];
                    const double b[3] = {s_x[k], s_y[k], s_z[k]};
                    const double d2 = gd_pbc_d2(P, a, b);
                    if ((ag & 0xFFFFFF) == self) return;
                    const int grp = kOneGroup ? 0 : ag >> 24;
                    if (d2 >= th_over && reg_over) {
                        if (kOneGroup) {
                            ++over[0];
                        } else {
#pragma unroll
                            for (int u = 0; u < kOverGroups; ++u) over[u] += (grp == u);
                        }
                        return;
                    }
                    // fp32 guess of the bin, corrected against the exact thresholds
                    int g = min((int)(sqrtf((float)d2) * inv_res), p.n_edges);
                    g += (int)(s_th[g + 1] <= d2) - (int)!(s_th[g] <= d2);
                    if (!uniform) {  // (NaNs end at 0)
                        while (g > 0 && !(s_th[g] <= d2)) --g;
                        while (g < p.n_edges && s_th[g + 1] <= d2) ++g;
                    }
                    atomicAdd(&my_hist[grp * nb + g], 1u);
                };
                int k = lane;
                for (; k + 32 < nt; k += 64) {  // two pairs per trip: half the loop overhead
                    pair(k);
                    pair(k + 32);
                }
                if (k < nt) pair(k);
                if (reg_over) {
#pragma unroll
                    for (int u = 0; u < (kOneGroup ? 1 : kOverGroups); ++u) {
                        uint32_t v = over[u];
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                        if (lane == 0 && v && u < p.n_group) my_hist[u * nb + p.n_edges] += v;
                    }
                }
                __syncwarp();
                unsigned long long *dst = p.hist + (size_t)(kPerFrame ? t : (int64_t)cls) * hist_len;
                for (int k2 = lane; k2 < hist_len; k2 += 32) {
                    const uint32_t n = my_hist[k2];
                    if (n) {
                        atomicAdd(dst + k2, (unsigned long long)n);
                        my_hist[k2] = 0;
                    }
                }
                __syncwarp();
            }
        }
    }
}

__global__ void k_rdf_classes(const int32_t *__restrict__ now, const int32_t *__restrict__ prev,
                              const int32_t *__restrict__ next, int64_t n_frames, int32_t n1,
                              const int32_t *__restrict__ site_label, int32_t n_sites, int32_t n_labels,
                              const int32_t *__restrict__ lut, int32_t *__restrict__ cls,
                              unsigned long long *__restrict__ first_key) {
    const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (idx >= n_frames * n1) return;
    // _uniqify_labels (rdf.py:19-27): digitize(right=True) sends -1 and 0 to "no site" and a
    // state s >= 1 to the label of site s-1
    auto lab = [&](int32_t s) { return (s >= 1 && s <= n_sites) ? site_label[s - 1] : -1; };
    const int32_t w = n_labels + 1;
    const int32_t code = ((lab(now[idx]) + 1) * w + (lab(prev[idx]) + 1)) * w + (lab(next[idx]) + 1);
    const int32_t c = lut[code];
    cls[idx] = c;
    const unsigned long long key = ((unsigned long long)(idx / n1) << 32) | (uint32_t)code;
    if (c >= 0 && first_key[c] > key) atomicMin(first_key + c, key);
}

}  // namespace

extern "C" gd_status_t gd_rdf_classes(const int32_t *d_states, const int32_t *d_states_prev,
                                      const int32_t *d_states_next, int64_t n_frames, int32_t n_floating,
                                      const int32_t *d_site_label, int32_t n_sites, int32_t n_labels,
                                      const int32_t *d_lut, int32_t *d_class,
                                      unsigned long long *d_first_key, void *stream) {
    GD_REQUIRE(n_frames >= 0 && n_floating >= 0 && n_sites >= 0 && n_labels >= 0 && n_labels < 1000, GD_ERR_INVALID,
               "gd_rdf_classes: bad size");
    if (n_frames == 0 || n_floating == 0) return GD_OK;
    GD_REQUIRE(d_states && d_states_prev && d_states_next && d_site_label && d_lut && d_class && d_first_key,
               GD_ERR_INVALID, "gd_rdf_classes: null pointer");
    const int64_t n = n_frames * n_floating;
    k_rdf_classes<<<(unsigned)gd_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
        d_states, d_states_prev, d_states_next, n_frames, n_floating, d_site_label, n_sites, n_labels, d_lut,
        d_class, d_first_key);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

namespace {

// common part of gd_rdf_hist / gd_rdf_hist_frames: shared-memory budget, grid, launch
gd_status_t rdf_launch(RdfParams &p, bool per_frame, void *stream) {
    const int64_t n_frames = p.n_frames;
    const int32_t n1 = p.n1, n2 = p.n2, n_group = p.n_group, n_edges = p.n_edges;
    const GdDeviceInfo &dev = gd_device_info();
    const size_t smem_cap = (size_t)(dev.max_smem_optin > 0 ? dev.max_smem_optin : 200 * 1024) -
                            (per_frame ? 1024 : 0);  // (room for the static copy of the frame's cell)
    const size_t fixed = (size_t)(n_edges + 2) * 8 + (size_t)kWarps * n_group * (n_edges + 1) * 4 + 64;
    GD_REQUIRE(fixed + 32 * 32 <= smem_cap, GD_ERR_CAPACITY,
               "gd_rdf_hist: %d groups x %d bins do not fit the per-warp shared-memory histograms", n_group,
               n_edges + 1);
    int64_t tile = (int64_t)((smem_cap - fixed) / 32);
    if (tile > n2) tile = n2;
    if (tile > 4096) tile = 4096;
    p.tile = (int32_t)tile;
    const size_t smem = (size_t)p.tile * 32 + fixed;
    const int sms = dev.sm_count > 0 ? dev.sm_count : 148;
    int64_t gy = gd_ceil_div((int64_t)sms * 4, n_frames);
    const int64_t gy_max = gd_ceil_div(n1, kWarps);
    if (gy > gy_max) gy = gy_max;
    if (gy < 1) gy = 1;
    int64_t gx = n_frames < (int64_t)sms * 64 ? n_frames : (int64_t)sms * 64;
    dim3 grid((unsigned)gx, (unsigned)gy);
    const bool one_group = n_group == 1;
    auto launch = [&](auto kernel) -> gd_status_t {
        GD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kernel<<<grid, kWarps * 32, smem, (cudaStream_t)stream>>>(p);
        return GD_OK;
    };
    gd_status_t rc;
    if (per_frame)
        rc = one_group ? launch(k_rdf_hist<true, true>) : launch(k_rdf_hist<false, true>);
    else
        rc = one_group ? launch(k_rdf_hist<true, false>) : launch(k_rdf_hist<false, false>);
    if (rc != GD_OK) return rc;
    GD_LAUNCH_CHECK();
    return GD_OK;
}

}  // namespace

extern "C" gd_status_t gd_rdf_hist(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                   const double *h_lattice, const int32_t *d_idx1, int32_t n1,
                                   const int32_t *d_class, int32_t n_class, const int32_t *d_idx2,
                                   const int32_t *d_group, int32_t n2, int32_t n_group,
                                   const double *d_edges, int32_t n_edges, double resolution, int32_t mode,
                                   int32_t skip_same_atom, unsigned long long *d_hist, void *stream) {
    GD_REQUIRE(n_atoms < (1 << 24) && n_group < 128, GD_ERR_INVALID,
               "gd_rdf_hist: at most 2^24 atoms and 127 groups");
    GD_REQUIRE(n_frames >= 0 && n_atoms > 0 && n1 >= 0 && n2 >= 0 && n_class > 0 && n_group > 0 && n_edges >= 2 &&
                   (mode == 0 || mode == 1) && resolution > 0.0,
               GD_ERR_INVALID, "gd_rdf_hist: bad argument");
    if (n_frames == 0 || n1 == 0 || n2 == 0) return GD_OK;
    GD_REQUIRE(d_pos_frac && h_lattice && d_idx1 && d_idx2 && d_edges && d_hist, GD_ERR_INVALID,
               "gd_rdf_hist: null pointer");
    RdfParams p;
    p.pos = d_pos_frac;
    p.n_frames = n_frames;
    p.n_atoms = n_atoms;
    p.idx1 = d_idx1;
    p.n1 = n1;
    p.cls = d_class;
    p.idx2 = d_idx2;
    p.group = d_group;
    p.n2 = n2;
    p.n_group = n_group;
    p.edges = d_edges;
    p.n_edges = n_edges;
    p.mode = mode;
    p.skip_same = skip_same_atom;
    p.inv_res = 1.0 / resolution;
    p.hist = d_hist;
    p.pbc = gd_pbc_build(h_lattice);
    p.pbc_frames = nullptr;
    return rdf_launch(p, false, stream);
}

extern "C" int64_t gd_rdf_frames_work_bytes(int64_t n_frames) {
    return (n_frames > 0 ? n_frames : 1) * (int64_t)sizeof(GdPbc);
}

extern "C" gd_status_t gd_rdf_hist_frames(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                          const double *h_lattices, void *d_work, const int32_t *d_idx1,
                                          int32_t n1, const int32_t *d_skip, const int32_t *d_idx2,
                                          const int32_t *d_group, int32_t n2, int32_t n_group,
                                          const double *d_edges, int32_t n_edges, double resolution,
                                          int32_t mode, int32_t skip_same_atom, unsigned long long *d_hist,
                                          void *stream) {
    GD_REQUIRE(n_atoms < (1 << 24) && n_group < 128, GD_ERR_INVALID,
               "gd_rdf_hist_frames: at most 2^24 atoms and 127 groups");
    GD_REQUIRE(n_frames >= 0 && n_atoms > 0 && n1 >= 0 && n2 >= 0 && n_group > 0 && n_edges >= 2 &&
                   (mode == 0 || mode == 1) && resolution > 0.0,
               GD_ERR_INVALID, "gd_rdf_hist_frames: bad argument");
    if (n_frames == 0 || n1 == 0 || n2 == 0) return GD_OK;
    GD_REQUIRE(d_pos_frac && h_lattices && d_work && d_idx1 && d_idx2 && d_edges && d_hist, GD_ERR_INVALID,
               "gd_rdf_hist_frames: null pointer");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_work) & 7) == 0, GD_ERR_ALIGNMENT,
               "gd_rdf_hist_frames: d_work must be 8-byte aligned");
    // one reduced cell per frame, built on the host (3x3 LLL each) and copied once
    std::vector<GdPbc> cells((size_t)n_frames);
    for (int64_t t = 0; t < n_frames; ++t) cells[(size_t)t] = gd_pbc_build(h_lattices + 9 * t);
    cudaStream_t st = (cudaStream_t)stream;
    GD_CUDA_CHECK(cudaMemcpyAsync(d_work, cells.data(), cells.size() * sizeof(GdPbc), cudaMemcpyHostToDevice, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));  // `cells` is freed on return
    RdfParams p;
    p.pos = d_pos_frac;
    p.n_frames = n_frames;
    p.n_atoms = n_atoms;
    p.idx1 = d_idx1;
    p.n1 = n1;
    p.cls = d_skip;
    p.idx2 = d_idx2;
    p.group = d_group;
    p.n2 = n2;
    p.n_group = n_group;
    p.edges = d_edges;
    p.n_edges = n_edges;
    p.mode = mode;
    p.skip_same = skip_same_atom;
    p.inv_res = 1.0 / resolution;
    p.hist = d_hist;
    p.pbc = cells[0];
    p.pbc_frames = static_cast<const GdPbc *>(d_work);
    return rdf_launch(p, true, stream);
}
