// K1 -- site-state assignment for every (frame, floating atom), fused with the event
// count of K2.
//
// Replaces _calculate_atom_states (reference src/gemdat/transitions.py:736-809): the
// MDAnalysis PeriodicKDTree build + query (float32 wrap, periodic images,
// scipy cKDTree ball query) and the last-write-wins scatter, for the outer and the
// inner sphere in one pass over the positions; and the change detection of
// _calculate_transition_events (:439-452) on the states while they are still in
// registers.
//
// B200 design ("strip walker")
//  * one warp owns a strip of 32 neighbouring atoms and walks it down one chunk of
//    frames: lane = atom, so the state of frame t and t+1 of an atom meet in the same
//    lane and the event count / "ever changes" flags cost no memory traffic;
//  * each frame row of a strip is 32 x 24 B = 768 contiguous bytes of the [T, A, 3]
//    float64 array; it streams HBM -> shared memory through a warp-private ring of
//    1-D TMA bulk copies (cp.async.bulk + mbarrier, SASS UBLKCP) kStages rows deep,
//    so no warp ever waits for another warp; every lane then reads its 3 doubles
//    with conflict-free LDS at a 24-byte stride;
//  * a CTA = 8 neighbouring strips of the same frame chunk (6 KB contiguous per
//    row); CTAs are ordered strip-group-major, so the CTAs resident at any time work
//    on the same few hundred atoms at different times and their cell / site records
//    stay in L1;
//  * instead of a KD-tree, a uniform cell grid over the wrapped float32 box maps a
//    position to the few sites whose sphere can reach that cell (exact sphere/box
//    test on the host, 4 B per cell); each (cell, site) entry carries 3 flag bits
//    saying on which axes a periodic image of the atom can matter, so the common
//    path evaluates one plain distance;
//  * the exact test reproduces the reference's arithmetic bit for bit:
//    float64 frac*L -> float32, float32 wrap, float32 periodic image (+L if
//    x <= cutoff else -L if L-x <= cutoff), float64 dx*dx+dy*dy+dz*dz <= r*r with
//    every operation individually rounded (no FMA).
//
// Algorithmic traffic: 24 B read + 8 B written per atom-frame.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "gd_common.cuh"

namespace {

constexpr uint32_t kMagic = 0x47534447u;  // "GDSG"
constexpr uint32_t kBlobVersion = 2;
constexpr int kMaxGroups = 64;
constexpr int kWarps = 8;  // strips per CTA
constexpr int kThreads = kWarps * 32;
constexpr int kStages = 4;        // frame rows in flight per warp
constexpr int kRowDoubles = 96;   // 32 atoms x 3
constexpr int kVisSmemMax = 24 * 1024;  // visited flags kept in shared memory up to this size
constexpr uint32_t kRankMask = 0xFFFFFu;  // 20 bits of rank, 3 bits of image flags above

struct GdGridHeader {  // 256 bytes, little endian, position independent
    uint32_t magic;
    uint32_t version;
    int32_t n_sites;  // sites that belong to a group (= rows of the site table)
    int32_t n_sites_total;
    int32_t n_groups;
    int32_t dims[3];
    int32_t n_entries;
    int32_t is_diag;
    int32_t max_count;
    float box[3];
    float cutoff32;
    float inv_cell[3];
    float pad0;
    double inv_box[3];
    double matrix[9];
    uint64_t off_sites;
    uint64_t off_cells;
    uint64_t off_entries;
    uint64_t off_groups;
    uint64_t total_bytes;
    unsigned char pad[256 - 216];
};
static_assert(sizeof(GdGridHeader) == 256, "header layout");

struct GdGroupRec {  // 48 bytes
    double r2_outer;
    double r2_inner;
    // pre-filter of the boundary-band diagnostics on the high word of d2 (monotone for
    // d2 >= 0): (unsigned)(hi(d2) - band_lo) <= band_span  <=>  d2 within ~1.1e-3 of r^2
    int32_t band_lo_o, band_span_o, band_lo_i, band_span_i;
    double r_outer;
    double r_inner;
};
static_assert(sizeof(GdGroupRec) == 48, "group record layout");

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

#ifdef __CUDA_ARCH__
#define GD_FADD(a, b) __fadd_rn((a), (b))
#define GD_FSUB(a, b) __fsub_rn((a), (b))
#define GD_FMUL(a, b) __fmul_rn((a), (b))
#else
#define GD_FADD(a, b) ((a) + (b))
#define GD_FSUB(a, b) ((a) - (b))
#define GD_FMUL(a, b) ((a) * (b))
#endif

// MDAnalysis _ortho_pbc, float32 (lib/src/calc_distances.h) [recalled]
__host__ __device__ inline float wrap_ortho(float c, float L, double invL) {
    if (c < 0.0f) {
        float once = GD_FADD(c, L);
        if (once < 0.0f) {
            int s = (int)floor((double)c * invL);
            c = GD_FSUB(c, GD_FMUL((float)s, L));
            if (c < 0.0f) c = GD_FADD(c, L);
        } else {
            c = once;
        }
    }
    if (c >= L) {
        float once = GD_FSUB(c, L);
        if (once >= L) {
            int s = (int)floor((double)c * invL);
            c = GD_FSUB(c, GD_FMUL((float)s, L));
            if (c >= L) c = GD_FSUB(c, L);
        } else {
            c = once;
        }
    }
    return c;
}

// device cell index along one axis: float32 multiply, truncate (clamped by the caller)
__host__ __device__ inline int cell_raw(float x, float inv_cell) { return (int)GD_FMUL(x, inv_cell); }

struct StatesParams {
    const double *pos;
    const double *next_row;  // [n_atoms,3] first row of the following shard, or null
    int64_t n_frames;
    int64_t n_chunks;
    int64_t n_cblocks;       // chunk blocks = ceil(n_chunks / cpg)
    int32_t n_atoms;
    int32_t f;               // frames per chunk
    int32_t n_strips;
    int32_t spg_shift;       // log2(strips per CTA)
    const float4 *sites;     // [n_sites] x, y, z, bits(natural | group << 24)
    const uint32_t *cells;   // [ncell] count << 24 | payload
    const uint32_t *entries; // rank | flags << 20, ranks ascending inside a cell
    const GdGroupRec *groups;
    int32_t *out_o;
    int32_t *out_i;
    uint8_t *visited;  // [2*n_sites_total] or null
    unsigned long long *band;
    int32_t *work;     // [n_atoms, n_chunks] event counts, or null
    uint8_t *changed;  // [2*n_atoms]
    int32_t n_sites_total;
    int32_t n_groups;
    int32_t dims[3];
    int32_t is_diag;
    int32_t vis_mode;  // visited flags: 0 none, 1 staged in shared memory, 2 global stores
    float box[3];
    float cutoff32;
    float inv_cell[3];
    double inv_box[3];
    double m[9];
    GdGroupRec g0;  // the only group when n_groups == 1 (read from the constant bank)
};

struct BandCounters {
    unsigned int o12 = 0, i12 = 0, of = 0, fi = 0;
};

// periodic image of the atom along one axis, if it has one (lib/_augment.pyx [recalled])
__device__ __forceinline__ double axis_delta_image(double d, double s, float a, float L, float r32) {
    float img;
    if (a <= r32)
        img = __fadd_rn(a, L);
    else if (__fsub_rn(L, a) <= r32)
        img = __fsub_rn(a, L);
    else
        return d;
    const double e = __dsub_rn(s, (double)img);
    return fabs(e) < fabs(d) ? e : d;
}

// rare path: bit 0/1 = outer sphere within 1e-12 / 4e-6 A, bit 2/3 = inner sphere
__device__ __noinline__ unsigned band_diag(double d2, const GdGroupRec *g) {
    unsigned r = 0;
    if (fabs(d2 - g->r2_outer) < g->r2_outer * 1e-3) {
        const double dd = fabs(sqrt(d2) - g->r_outer);
        r |= (dd < 1e-12 ? 1u : 0u) | (dd < 4e-6 ? 2u : 0u);
    }
    if (fabs(d2 - g->r2_inner) < g->r2_inner * 1e-3) {
        const double dd = fabs(sqrt(d2) - g->r_inner);
        r |= (dd < 1e-12 ? 4u : 0u) | (dd < 4e-6 ? 8u : 0u);
    }
    return r;
}

__device__ __noinline__ float wrap_slow(float c, float L, double invL) { return wrap_ortho(c, L, invL); }

// site states of one atom-frame
template <bool kOneGroup>
__device__ __forceinline__ void classify(const StatesParams &p, const GdGroupRec *s_groups,
                                         uint8_t *s_vis, double f0, double f1, double f2,
                                         int32_t &best_o, int32_t &best_i, BandCounters &bc) {
    // ---- reference staging: float64 Cartesian -> float32 -> wrap
    double cx, cy, cz;
    if (p.is_diag) {
        cx = __dmul_rn(f0, p.m[0]);
        cy = __dmul_rn(f1, p.m[4]);
        cz = __dmul_rn(f2, p.m[8]);
    } else {
        cx = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[0]), __dmul_rn(f1, p.m[3])), __dmul_rn(f2, p.m[6]));
        cy = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[1]), __dmul_rn(f1, p.m[4])), __dmul_rn(f2, p.m[7]));
        cz = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[2]), __dmul_rn(f1, p.m[5])), __dmul_rn(f2, p.m[8]));
    }
    float ax = __double2float_rn(cx), ay = __double2float_rn(cy), az = __double2float_rn(cz);
    const float Lx = p.box[0], Ly = p.box[1], Lz = p.box[2];
    if (!(ax >= 0.0f && ax < Lx)) ax = wrap_slow(ax, Lx, p.inv_box[0]);
    if (!(ay >= 0.0f && ay < Ly)) ay = wrap_slow(ay, Ly, p.inv_box[1]);
    if (!(az >= 0.0f && az < Lz)) az = wrap_slow(az, Lz, p.inv_box[2]);
    const double a0x = (double)ax, a0y = (double)ay, a0z = (double)az;

    // ---- candidate sites of this cell
    const unsigned ix = min((unsigned)cell_raw(ax, p.inv_cell[0]), (unsigned)(p.dims[0] - 1));
    const unsigned iy = min((unsigned)cell_raw(ay, p.inv_cell[1]), (unsigned)(p.dims[1] - 1));
    const unsigned iz = min((unsigned)cell_raw(az, p.inv_cell[2]), (unsigned)(p.dims[2] - 1));
    const uint32_t cell = __ldg(p.cells + ((ix * (unsigned)p.dims[1] + iy) * (unsigned)p.dims[2] + iz));
    const uint32_t count = cell >> 24, payload = cell & 0xFFFFFFu;

    best_o = GD_NOSITE;
    best_i = GD_NOSITE;
    for (uint32_t k = 0; k < count; ++k) {
        const uint32_t word = (count == 1) ? payload : __ldg(p.entries + payload + k);
        const float4 rec = __ldg(p.sites + (word & kRankMask));
        const int natgrp = __float_as_int(rec.w);
        const GdGroupRec &g = kOneGroup ? p.g0 : s_groups[natgrp >> 24];
        const double sx = (double)rec.x, sy = (double)rec.y, sz = (double)rec.z;
        double dx = __dsub_rn(sx, a0x), dy = __dsub_rn(sy, a0y), dz = __dsub_rn(sz, a0z);
        if (word >> 20) {  // a periodic image of the atom can be the closer one (rare)
            if (word & (1u << 20)) dx = axis_delta_image(dx, sx, ax, Lx, p.cutoff32);
            if (word & (1u << 21)) dy = axis_delta_image(dy, sy, ay, Ly, p.cutoff32);
            if (word & (1u << 22)) dz = axis_delta_image(dz, sz, az, Lz, p.cutoff32);
        }
        const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
        const int natural = natgrp & 0x00FFFFFF;
        const bool hit_o = d2 <= g.r2_outer, hit_i = d2 <= g.r2_inner;
        if (hit_o) best_o = natural;  // entries ascend in overwrite order: last match wins
        if (hit_i) best_i = natural;
        if (p.vis_mode == 1) {
            if (hit_o) s_vis[natural] = 1;
            if (hit_i) s_vis[p.n_sites_total + natural] = 1;
        } else if (p.vis_mode == 2) {
            if (hit_o) p.visited[natural] = 1;
            if (hit_i) p.visited[p.n_sites_total + natural] = 1;
        }
        // boundary-band diagnostics: integer pre-filter on the high word, exact test on the
        // rare path
        const int hi = __double2hiint(d2);
        if ((unsigned)(hi - g.band_lo_o) <= (unsigned)g.band_span_o ||
            (unsigned)(hi - g.band_lo_i) <= (unsigned)g.band_span_i) {
            const unsigned b = band_diag(d2, &g);
            bc.o12 += b & 1u;
            bc.of += (b >> 1) & 1u;
            bc.i12 += (b >> 2) & 1u;
            bc.fi += (b >> 3) & 1u;
        }
    }
}

template <bool kTma, bool kOneGroup>
__global__ void __launch_bounds__(kThreads, 4)
    k_site_states(const __grid_constant__ StatesParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    double *ring = reinterpret_cast<double *>(smem);  // [kWarps][kStages][96]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)kWarps * kStages * kRowDoubles * 8);
    GdGroupRec *s_groups = reinterpret_cast<GdGroupRec *>(full + kWarps * kStages);
    uint8_t *s_vis = reinterpret_cast<uint8_t *>(s_groups + kMaxGroups);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int g = tid; g < p.n_groups; g += kThreads) s_groups[g] = p.groups[g];
    const int n_vis = 2 * p.n_sites_total;
    if (p.vis_mode == 1)
        for (int i = tid; i < (n_vis + 3) / 4; i += kThreads) reinterpret_cast<uint32_t *>(s_vis)[i] = 0;
    if (kTma && lane == 0) {
        for (int s = 0; s < kStages; ++s) gd_mbar_init(&full[warp * kStages + s], 1);
        gd_fence_barrier_init();
    }
    __syncthreads();

    // ---- which (strip, chunk) this warp owns: CTAs are strip-group-major
    const int spg = 1 << p.spg_shift;
    const int64_t grp = blockIdx.x / p.n_cblocks;
    const int64_t cb = blockIdx.x - grp * p.n_cblocks;
    const int32_t strip = (int32_t)(grp << p.spg_shift) + (warp & (spg - 1));
    const int64_t chunk = cb * (kWarps >> p.spg_shift) + (warp >> p.spg_shift);
    BandCounters bc;

    if (strip < p.n_strips && chunk < p.n_chunks) {
        const int32_t a0 = strip * 32;
        const int32_t a = a0 + lane;
        const bool lane_on = a < p.n_atoms;
        const int r = min(32, p.n_atoms - a0);  // atoms in this strip
        const int64_t t0 = chunk * p.f;
        const int64_t t1 = min(p.n_frames, t0 + (int64_t)p.f);
        const int n_own = (int)(t1 - t0);
        // halo row for the event at the chunk's last frame
        const bool halo = p.work != nullptr && (t1 < p.n_frames || p.next_row != nullptr);
        const int n_rows = n_own + (halo ? 1 : 0);
        const int64_t A3 = (int64_t)p.n_atoms * 3;
        const double *row0 = p.pos + t0 * A3 + (int64_t)a0 * 3;
        const double *halo_row = (t1 < p.n_frames) ? p.pos + t1 * A3 + (int64_t)a0 * 3
                                                   : (p.next_row ? p.next_row + (int64_t)a0 * 3 : nullptr);
        auto row_ptr = [&](int k) { return k < n_own ? row0 + (int64_t)k * A3 : halo_row; };

        // bytes of a row that the bulk copy moves (multiple of 16); an odd strip leaves the
        // last atom's z to a plain load
        const uint32_t row_bytes = (uint32_t)(r * 24) & ~15u;
        const int covered = (int)(row_bytes >> 3);  // doubles in shared memory
        double *my_ring = ring + (size_t)warp * kStages * kRowDoubles;
        uint64_t *my_full = full + warp * kStages;

        if (kTma && lane == 0) {
            for (int s = 0; s < kStages && s < n_rows; ++s) {
                gd_mbar_arrive_expect_tx(&my_full[s], row_bytes);
                gd_tma_load_1d(my_ring + s * kRowDoubles, row_ptr(s), row_bytes, &my_full[s]);
            }
        }
        // register prefetch of the non-TMA variant (odd n_atoms: rows are only 8-byte aligned)
        double n0 = 0.0, n1 = 0.0, n2 = 0.0;
        if (!kTma && lane_on) {
            const double *q = row_ptr(0) + 3 * lane;
            n0 = __ldg(q);
            n1 = __ldg(q + 1);
            n2 = __ldg(q + 2);
        }

        int32_t prev_o = GD_NOSITE, prev_i = GD_NOSITE;
        int cnt = 0;
        int chg = 0;  // bit 0: outer state changed somewhere, bit 1: inner
        const bool tail_lane = 3 * lane + 2 >= covered;  // z of an odd strip's last atom
        const double *my_row = my_ring + 3 * lane;
        int slot = 0;
        uint32_t phase = 0;
        int32_t *po = p.out_o + t0 * p.n_atoms + a;
        int32_t *pi = p.out_i + t0 * p.n_atoms + a;

        for (int k = 0; k < n_rows; ++k) {
            double f0 = 0.0, f1 = 0.0, f2 = 0.0;
            if (kTma) {
                gd_mbar_wait(&my_full[slot], phase);
                if (lane_on) {
                    const double *sp = my_row + slot * kRowDoubles;
                    f0 = sp[0];
                    f1 = sp[1];
                    f2 = tail_lane ? __ldg(row_ptr(k) + 3 * lane + 2) : sp[2];
                }
                // The slot is refilled by the async proxy (TMA); the reads above are
                // generic-proxy LDS.  Order them before the refill (observed on B200
                // without the fence: whole warps read the data of a later row).
                gd_fence_proxy_async();
                __syncwarp();
                if (lane == 0 && k + kStages < n_rows) {
                    gd_mbar_arrive_expect_tx(&my_full[slot], row_bytes);
                    gd_tma_load_1d(my_ring + slot * kRowDoubles, row_ptr(k + kStages), row_bytes,
                                   &my_full[slot]);
                }
                if (++slot == kStages) {
                    slot = 0;
                    phase ^= 1u;
                }
            } else {
                f0 = n0;
                f1 = n1;
                f2 = n2;
                if (lane_on && k + 1 < n_rows) {
                    const double *q = row_ptr(k + 1) + 3 * lane;
                    n0 = __ldg(q);
                    n1 = __ldg(q + 1);
                    n2 = __ldg(q + 2);
                }
            }
            if (!lane_on) continue;

            int32_t so, si;
            classify<kOneGroup>(p, s_groups, s_vis, f0, f1, f2, so, si, bc);

            if (k > 0) {
                const int d_o = so != prev_o, d_i = si != prev_i;
                cnt += d_o | d_i;
                chg |= d_o | (d_i << 1);
            }
            prev_o = so;
            prev_i = si;
            if (k < n_own) {
                *po = so;
                *pi = si;
                po += p.n_atoms;
                pi += p.n_atoms;
            }
        }
        if (p.work && lane_on) {
            p.work[(int64_t)a * p.n_chunks + chunk] = cnt;
            if (chg & 1) p.changed[a] = 1;
            if (chg & 2) p.changed[p.n_atoms + a] = 1;
        }
    }

    if (p.band) {
        const int o12 = gd_warp_sum((int)bc.o12), i12 = gd_warp_sum((int)bc.i12);
        const int of = gd_warp_sum((int)bc.of), fi = gd_warp_sum((int)bc.fi);
        if (lane == 0) {
            if (o12) atomicAdd(p.band + 0, (unsigned long long)o12);
            if (i12) atomicAdd(p.band + 1, (unsigned long long)i12);
            if (of) atomicAdd(p.band + 2, (unsigned long long)of);
            if (fi) atomicAdd(p.band + 3, (unsigned long long)fi);
        }
    }
    if (p.vis_mode == 1) {
        __syncthreads();
        for (int i = tid; i < n_vis; i += kThreads)
            if (s_vis[i]) p.visited[i] = 1;
    }
}

}  // namespace

// ------------------------------------------------------------------------------------ host build

extern "C" gd_status_t gd_sitegrid_build(const double *h_lattice, const double *h_sites_frac,
                                         const int32_t *h_site_group, int32_t n_sites,
                                         const double *h_group_radius, const double *h_group_inner,
                                         int32_t n_groups, void *h_blob, size_t cap,
                                         size_t *needed) {
    GD_REQUIRE(h_lattice && h_sites_frac && h_site_group && h_group_radius && h_group_inner && needed,
               GD_ERR_INVALID, "gd_sitegrid_build: null argument");
    GD_REQUIRE(n_sites > 0 && n_sites <= (int32_t)kRankMask, GD_ERR_INVALID,
               "gd_sitegrid_build: n_sites must be in [1, 2^20)");
    GD_REQUIRE(n_groups > 0 && n_groups <= kMaxGroups, GD_ERR_INVALID,
               "gd_sitegrid_build: n_groups must be in [1, %d]", kMaxGroups);

    // pymatgen Lattice.parameters: lengths = row norms, angles from dot products
    const double *m = h_lattice;
    double len[3];
    for (int i = 0; i < 3; ++i)
        len[i] = sqrt(m[3 * i] * m[3 * i] + m[3 * i + 1] * m[3 * i + 1] + m[3 * i + 2] * m[3 * i + 2]);
    for (int i = 0; i < 3; ++i) {
        const int j = (i + 1) % 3, k = (i + 2) % 3;
        double c = (m[3 * j] * m[3 * k] + m[3 * j + 1] * m[3 * k + 1] + m[3 * j + 2] * m[3 * k + 2]) /
                   (len[j] * len[k]);
        c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
        const float ang = (float)(acos(c) * 180.0 / M_PI);
        GD_REQUIRE(ang == 90.0f, GD_ERR_TRICLINIC,
                   "lattice angle %d is %.9g deg: only orthorhombic cells are supported "
                   "(the reference is ill-defined for triclinic cells, see DESIGN.md)",
                   i, (double)ang);
    }
    bool is_diag = true;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            if (i != j && m[3 * i + j] != 0.0) is_diag = false;

    double cutoff = h_group_radius[0];
    for (int g = 1; g < n_groups; ++g) cutoff = std::max(cutoff, h_group_radius[g]);
    GD_REQUIRE(cutoff > 0.0, GD_ERR_INVALID, "gd_sitegrid_build: radii must be positive");
    std::vector<GdGroupRec> groups(n_groups);
    for (int g = 0; g < n_groups; ++g) {
        const double ro = h_group_radius[g] * 1.0;
        const double ri = h_group_radius[g] * h_group_inner[g];
        GD_REQUIRE(!(cutoff < ri), GD_ERR_RADIUS_GT_CUTOFF,
                   "Set cutoff greater or equal to the radius.");
        GdGroupRec rec;
        rec.r2_outer = ro * ro;
        rec.r2_inner = ri * ri;
        // band_diag applies the exact filter |d2 - r2| < r2 * 1e-3; a window of 1.1e-3
        // widened by one unit of the high word is a superset of it
        auto hi_word = [](double v) {
            int64_t bits;
            memcpy(&bits, &v, 8);
            return (int32_t)(bits >> 32);
        };
        rec.band_lo_o = hi_word(rec.r2_outer * (1.0 - 1.1e-3)) - 1;
        rec.band_span_o = hi_word(rec.r2_outer * (1.0 + 1.1e-3)) + 1 - rec.band_lo_o;
        rec.band_lo_i = hi_word(rec.r2_inner * (1.0 - 1.1e-3)) - 1;
        rec.band_span_i = hi_word(rec.r2_inner * (1.0 + 1.1e-3)) + 1 - rec.band_lo_i;
        rec.r_outer = ro;
        rec.r_inner = ri;
        groups[g] = rec;
    }

    float box[3];
    double inv_box[3];
    for (int i = 0; i < 3; ++i) {
        box[i] = (float)len[i];
        inv_box[i] = 1.0 / (double)box[i];
    }
    const float r32 = (float)cutoff;

    // rank order = overwrite order of the reference: groups in dict order, ascending
    // site index inside a group
    for (int32_t s = 0; s < n_sites; ++s)
        GD_REQUIRE(h_site_group[s] >= -1 && h_site_group[s] < n_groups, GD_ERR_INVALID,
                   "gd_sitegrid_build: site %d has group %d", s, h_site_group[s]);
    std::vector<int32_t> order;
    order.reserve(n_sites);
    for (int g = 0; g < n_groups; ++g)
        for (int32_t s = 0; s < n_sites; ++s)
            if (h_site_group[s] == g) order.push_back(s);
    const int32_t n_ranked = (int32_t)order.size();

    // float32 staging of the site centres (float64 frac @ M -> float32 -> wrap)
    std::vector<float> sx((size_t)std::max(n_ranked, 1) * 3);
    for (int32_t e = 0; e < n_ranked; ++e) {
        const double *f = h_sites_frac + 3 * (size_t)order[e];
        for (int j = 0; j < 3; ++j) {
            volatile double t0 = f[0] * m[j];
            volatile double t1 = f[1] * m[3 + j];
            volatile double t2 = f[2] * m[6 + j];
            volatile double acc = t0 + t1;
            const double cart = acc + t2;
            sx[3 * e + j] = wrap_ortho((float)cart, box[j], inv_box[j]);
        }
    }

    // grid resolution: cells of ~k * cutoff (k = 0.75; GD_GRID_K overrides for tuning),
    // at most 256 cells per axis / 2^22 cells in total
    double kfac = 0.75;
    if (const char *env = getenv("GD_GRID_K")) {
        const double v = atof(env);
        if (v > 0.05 && v < 100.0) kfac = v;
    }
    int dims[3];
    double target = std::max(kfac * cutoff, 1e-3);
    for (;;) {
        int64_t total = 1;
        for (int j = 0; j < 3; ++j) {
            int d = (int)floor(len[j] / target);
            dims[j] = std::min(std::max(d, 1), 256);
            total *= dims[j];
        }
        if (total <= (int64_t(1) << 22)) break;
        target *= 1.26;
    }
    float inv_cell[3];
    for (int j = 0; j < 3; ++j) inv_cell[j] = (float)((double)dims[j] / (double)box[j]);
    const int64_t ncell = (int64_t)dims[0] * dims[1] * dims[2];

    // exact float32 interval of every cell along every axis: cell i of axis j holds the
    // wrapped coordinates x with lo[i] <= x <= hi[i], evaluated with the device's own
    // formula min(trunc(fl32(x * inv_cell)), dim - 1), which is monotone in x
    std::vector<float> edge[3];
    for (int j = 0; j < 3; ++j) {
        edge[j].assign(dims[j] + 1, 0.0f);
        for (int i = 1; i < dims[j]; ++i) {
            float g = (float)((double)i / (double)inv_cell[j]);
            while (g > 0.0f && cell_raw(g, inv_cell[j]) >= i) g = nextafterf(g, -INFINITY);
            while (cell_raw(g, inv_cell[j]) < i) g = nextafterf(g, INFINITY);
            edge[j][i] = g;  // smallest float that lands in cell >= i
        }
        edge[j][dims[j]] = box[j];
    }

    // (cell, site) entries by an exact sphere / box test (periodic), with the axes on
    // which an image of the atom can be the closer one flagged.  `margin` covers every
    // float32 rounding of the staged coordinates and images (<= 4e-6 A).
    const double margin = 2e-3;
    const double eps = 1e-4;
    struct AxisHit {
        int cell;
        double dist;
        bool flag;
    };
    std::vector<AxisHit> hits[3];
    std::vector<uint32_t> e_cell, e_word;
    e_cell.reserve((size_t)n_ranked * 32);
    e_word.reserve((size_t)n_ranked * 32);
    for (int32_t e = 0; e < n_ranked; ++e) {
        const GdGroupRec &g = groups[h_site_group[order[e]]];
        const double reach = std::max(g.r_outer, g.r_inner) + margin;
        for (int j = 0; j < 3; ++j) {
            hits[j].clear();
            const double s = (double)sx[3 * e + j];
            const double L = (double)box[j];
            for (int i = 0; i < dims[j]; ++i) {
                const double lo = (double)edge[j][i], hi = (double)edge[j][i + 1];
                double best = std::max(0.0, std::max(lo - s, s - hi));
                bool flag = false;
                if (lo <= (double)r32 + eps) {  // atoms here may carry the +L image
                    const double d = std::max(0.0, lo + L - s - eps);
                    if (d <= reach) {
                        flag = true;
                        best = std::min(best, d);
                    }
                }
                if (hi >= L - (double)r32 - eps) {  // ... or the -L image
                    const double d = std::max(0.0, s - (hi - L) - eps);
                    if (d <= reach) {
                        flag = true;
                        best = std::min(best, d);
                    }
                }
                if (best <= reach) hits[j].push_back(AxisHit{i, best, flag});
            }
        }
        const double reach2 = reach * reach;
        for (const AxisHit &hx : hits[0])
            for (const AxisHit &hy : hits[1]) {
                const double dxy = hx.dist * hx.dist + hy.dist * hy.dist;
                if (dxy > reach2) continue;
                for (const AxisHit &hz : hits[2]) {
                    if (dxy + hz.dist * hz.dist > reach2) continue;
                    const uint32_t flags = (hx.flag ? 1u : 0u) | (hy.flag ? 2u : 0u) | (hz.flag ? 4u : 0u);
                    e_cell.push_back((uint32_t)(((int64_t)hx.cell * dims[1] + hy.cell) * dims[2] + hz.cell));
                    e_word.push_back((uint32_t)e | (flags << 20));
                }
            }
    }
    const size_t n_entries = e_cell.size();
    GD_REQUIRE(n_entries < (size_t(1) << 24), GD_ERR_INVALID,
               "gd_sitegrid_build: %zu (cell, site) entries exceed the 2^24 limit", n_entries);

    std::vector<uint32_t> count(ncell, 0);
    for (size_t k = 0; k < n_entries; ++k) ++count[e_cell[k]];
    std::vector<uint32_t> start(ncell + 1, 0);
    uint32_t maxc = 0;
    for (int64_t c = 0; c < ncell; ++c) {
        start[c + 1] = start[c] + count[c];
        maxc = std::max(maxc, count[c]);
    }
    GD_REQUIRE(maxc <= 255, GD_ERR_INVALID,
               "gd_sitegrid_build: %u candidate sites in one cell exceed the limit of 255", maxc);

    const size_t off_sites = sizeof(GdGridHeader);
    const size_t off_groups = align256(off_sites + (size_t)n_ranked * 16);
    const size_t off_cells = align256(off_groups + (size_t)n_groups * sizeof(GdGroupRec));
    const size_t off_entries = align256(off_cells + (size_t)ncell * 4);
    const size_t total = align256(off_entries + std::max<size_t>(n_entries, 1) * 4);
    *needed = total;
    if (!h_blob || cap < total) {
        gd_set_error("gd_sitegrid_build: blob needs %zu bytes", total);
        return GD_ERR_CAPACITY;
    }
    memset(h_blob, 0, total);
    unsigned char *base = static_cast<unsigned char *>(h_blob);
    GdGridHeader *h = reinterpret_cast<GdGridHeader *>(base);
    h->magic = kMagic;
    h->version = kBlobVersion;
    h->n_sites = n_ranked;
    h->n_sites_total = n_sites;
    h->n_groups = n_groups;
    for (int j = 0; j < 3; ++j) {
        h->dims[j] = dims[j];
        h->box[j] = box[j];
        h->inv_cell[j] = inv_cell[j];
        h->inv_box[j] = inv_box[j];
    }
    h->n_entries = (int32_t)n_entries;
    h->is_diag = is_diag ? 1 : 0;
    h->max_count = (int32_t)maxc;
    h->cutoff32 = r32;
    for (int j = 0; j < 9; ++j) h->matrix[j] = m[j];
    h->off_sites = off_sites;
    h->off_groups = off_groups;
    h->off_cells = off_cells;
    h->off_entries = off_entries;
    h->total_bytes = total;

    float *srec = reinterpret_cast<float *>(base + off_sites);
    for (int32_t e = 0; e < n_ranked; ++e) {
        srec[4 * e + 0] = sx[3 * e + 0];
        srec[4 * e + 1] = sx[3 * e + 1];
        srec[4 * e + 2] = sx[3 * e + 2];
        const int32_t packed = order[e] | (h_site_group[order[e]] << 24);
        memcpy(&srec[4 * e + 3], &packed, 4);
    }
    memcpy(base + off_groups, groups.data(), (size_t)n_groups * sizeof(GdGroupRec));

    // counting sort by cell; sites were visited in ascending rank so every list ascends
    uint32_t *entries = reinterpret_cast<uint32_t *>(base + off_entries);
    std::vector<uint32_t> cursor(start.begin(), start.end() - 1);
    for (size_t k = 0; k < n_entries; ++k) entries[cursor[e_cell[k]]++] = e_word[k];
    uint32_t *cells = reinterpret_cast<uint32_t *>(base + off_cells);
    for (int64_t c = 0; c < ncell; ++c)
        cells[c] = (count[c] << 24) | (count[c] == 1 ? entries[start[c]] : start[c]);
    return GD_OK;
}

// ------------------------------------------------------------------------------------ launch

extern "C" gd_status_t gd_site_states(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                      const double *d_next_pos_row, const void *d_blob,
                                      const void *h_blob, int32_t *d_states_outer,
                                      int32_t *d_states_inner, uint8_t *d_visited,
                                      unsigned long long *d_band_counts, int32_t *d_event_work,
                                      uint8_t *d_changed, void *stream) {
    GD_REQUIRE(d_pos_frac && d_blob && h_blob && d_states_outer && d_states_inner, GD_ERR_INVALID,
               "gd_site_states: null argument");
    GD_REQUIRE(n_frames >= 0 && n_atoms > 0, GD_ERR_INVALID, "gd_site_states: bad size");
    GD_REQUIRE((d_event_work == nullptr) == (d_changed == nullptr), GD_ERR_INVALID,
               "gd_site_states: d_event_work and d_changed must both be given or both be NULL");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_pos_frac) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states: d_pos_frac must be 16-byte aligned");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_next_pos_row) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states: d_next_pos_row must be 16-byte aligned");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_blob) & 255) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states: d_blob must be 256-byte aligned");
    const GdGridHeader *h = static_cast<const GdGridHeader *>(h_blob);
    GD_REQUIRE(h->magic == kMagic && h->version == kBlobVersion, GD_ERR_INVALID,
               "gd_site_states: h_blob is not a site grid");
    if (n_frames == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");

    const unsigned char *base = static_cast<const unsigned char *>(d_blob);
    const GdChunks geo = gd_chunks(n_frames, n_atoms);
    StatesParams p;
    memset(&p, 0, sizeof(p));
    p.pos = d_pos_frac;
    p.next_row = d_next_pos_row;
    p.n_frames = n_frames;
    p.n_atoms = n_atoms;
    p.f = geo.f;
    p.n_chunks = geo.n_chunks;
    p.n_strips = (n_atoms + 31) / 32;
    int spg_shift = 0;
    while ((1 << spg_shift) < kWarps && (1 << spg_shift) < p.n_strips) ++spg_shift;
    p.spg_shift = spg_shift;
    const int cpg = kWarps >> spg_shift;
    p.n_cblocks = gd_ceil_div(geo.n_chunks, cpg);
    const int64_t n_groups_of_strips = gd_ceil_div(p.n_strips, 1 << spg_shift);
    const int64_t grid = n_groups_of_strips * p.n_cblocks;
    GD_REQUIRE(grid < (int64_t(1) << 31), GD_ERR_INVALID, "gd_site_states: shard too large");
    p.sites = reinterpret_cast<const float4 *>(base + h->off_sites);
    p.cells = reinterpret_cast<const uint32_t *>(base + h->off_cells);
    p.entries = reinterpret_cast<const uint32_t *>(base + h->off_entries);
    p.groups = reinterpret_cast<const GdGroupRec *>(base + h->off_groups);
    p.out_o = d_states_outer;
    p.out_i = d_states_inner;
    p.visited = d_visited;
    p.band = d_band_counts;
    p.work = d_event_work;
    p.changed = d_changed;
    p.n_sites_total = h->n_sites_total;
    p.n_groups = h->n_groups;
    p.is_diag = h->is_diag;
    p.cutoff32 = h->cutoff32;
    for (int j = 0; j < 3; ++j) {
        p.dims[j] = h->dims[j];
        p.box[j] = h->box[j];
        p.inv_cell[j] = h->inv_cell[j];
        p.inv_box[j] = h->inv_box[j];
    }
    for (int j = 0; j < 9; ++j) p.m[j] = h->matrix[j];

    const size_t vis_bytes = 2 * (size_t)h->n_sites_total;
    p.vis_mode = d_visited == nullptr ? 0 : (vis_bytes <= (size_t)kVisSmemMax ? 1 : 2);
    p.g0 = *reinterpret_cast<const GdGroupRec *>(static_cast<const unsigned char *>(h_blob) + h->off_groups);
    const size_t smem = (size_t)kWarps * kStages * kRowDoubles * 8 + kWarps * kStages * sizeof(uint64_t) +
                        kMaxGroups * sizeof(GdGroupRec) +
                        (p.vis_mode == 1 ? ((vis_bytes + 15) & ~size_t(15)) : 0);
    // rows of an odd-sized atom axis start on 8-byte boundaries only: no bulk copies then
    const bool tma = (n_atoms % 2) == 0;
    const bool one = h->n_groups == 1;
    void (*kern)(StatesParams) =
        tma ? (one ? k_site_states<true, true> : k_site_states<true, false>)
            : (one ? k_site_states<false, true> : k_site_states<false, false>);
    GD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
