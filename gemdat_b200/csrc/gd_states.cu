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
constexpr uint32_t kBlobVersion = 4;
constexpr int kMaxGroups = 64;
// tuning knobs (tools/k1_variants.sh builds A/B libraries with -D overrides)
#ifndef GD_K1_WARPS
#define GD_K1_WARPS 4
#endif
#ifndef GD_K1_STAGES
#define GD_K1_STAGES 4
#endif
#ifndef GD_K1_IMAGE_FILTER
#define GD_K1_IMAGE_FILTER 1
#endif
#ifndef GD_K1_EXPECT_HIT
#define GD_K1_EXPECT_HIT 0
#endif
#if GD_K1_EXPECT_HIT
#define GD_K1_LIKELY(x) __builtin_expect(!!(x), 1)
#else
#define GD_K1_LIKELY(x) (x)
#endif
#ifndef GD_K1_MIN_CTAS
#define GD_K1_MIN_CTAS 6
#endif
constexpr int kWarps = GD_K1_WARPS;  // strips per CTA
constexpr int kThreads = kWarps * 32;
constexpr int kStages = GD_K1_STAGES;  // frame rows in flight per warp
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
    uint64_t off_excl;
    unsigned char pad[256 - 224];
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

constexpr uint32_t kSiteImgShift = 20;  // site record w: natural (20 bits) | image axes << 20 | group << 24

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
    const float4 *sites;     // [n_sites] in rank (overwrite) order: x, y, z, bits(natural | img << 20 | group << 24)
    const float *excl2;      // [n_sites] squared exclusive radius of the site (< 0: none)
    const uint32_t *cells;   // [ncell] count << 24 | payload
    const uint32_t *entries; // rank | flags << 20, ranks ascending inside a cell
    const GdGroupRec *groups;
    int32_t *out_o;
    int32_t *out_i;
    uint8_t *visited;  // [2*n_sites_total] or null
    unsigned long long *band;
    int32_t *work;     // [n_atoms, n_chunks] event counts, or null
    uint8_t *changed;  // [2*n_atoms]
    int32_t *stage;    // [n_chunks, n_atoms, stage_cap, 6] event rows as they are found, or null
    int32_t *overflow; // set when a (chunk, atom) segment holds more than stage_cap events
    int32_t stage_cap;
    int32_t n_sites_total;
    int32_t n_groups;
    int32_t dims[3];
    int32_t vis_mode;  // visited flags: 0 none, 1 staged in shared memory, 2 global stores
    float box[3];
    float cutoff32;
    float inv_cell[3];
    double inv_box[3];
    double qbox[3];  // box / 4: below this |d| on an axis the atom's periodic image cannot be closer
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

// What the walk knows about one atom-frame besides its two states.
struct Staged {
    float ax, ay, az;  // float32-staged, wrapped Cartesian position
    uint32_t cell;     // its cell word (filled by lookup_cell)
};

// float64 frac -> float64 Cartesian -> float32 -> wrap (the reference's staging).
// kDiag: orthorhombic cell given as a diagonal matrix (one multiply per axis).
template <bool kDiag>
__device__ __forceinline__ Staged stage_atom(const StatesParams &p, double f0, double f1, double f2) {
    double cx, cy, cz;
    if (kDiag) {
        cx = __dmul_rn(f0, p.m[0]);
        cy = __dmul_rn(f1, p.m[4]);
        cz = __dmul_rn(f2, p.m[8]);
    } else {
        cx = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[0]), __dmul_rn(f1, p.m[3])), __dmul_rn(f2, p.m[6]));
        cy = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[1]), __dmul_rn(f1, p.m[4])), __dmul_rn(f2, p.m[7]));
        cz = __dadd_rn(__dadd_rn(__dmul_rn(f0, p.m[2]), __dmul_rn(f1, p.m[5])), __dmul_rn(f2, p.m[8]));
    }
    Staged s;
    s.ax = __double2float_rn(cx);
    s.ay = __double2float_rn(cy);
    s.az = __double2float_rn(cz);
    s.cell = 0;
    // positions handed over by Trajectory.positions are in [0, 1): one test, wrap out of line
    // 0 <= x < L in ONE unsigned compare per axis: non-negative floats order like their bit
    // patterns, negative ones (and -0, NaN) have the sign bit set and land on the slow path,
    // which is exact for them
    const bool inside = (__float_as_uint(s.ax) < __float_as_uint(p.box[0])) &
                        (__float_as_uint(s.ay) < __float_as_uint(p.box[1])) &
                        (__float_as_uint(s.az) < __float_as_uint(p.box[2]));
    if (__builtin_expect(!inside, 0)) {
        s.ax = wrap_slow(s.ax, p.box[0], p.inv_box[0]);
        s.ay = wrap_slow(s.ay, p.box[1], p.inv_box[1]);
        s.az = wrap_slow(s.az, p.box[2], p.inv_box[2]);
    }
    return s;
}

__device__ __forceinline__ void lookup_cell(const StatesParams &p, Staged &s) {
    const unsigned ix = min((unsigned)cell_raw(s.ax, p.inv_cell[0]), (unsigned)(p.dims[0] - 1));
    const unsigned iy = min((unsigned)cell_raw(s.ay, p.inv_cell[1]), (unsigned)(p.dims[1] - 1));
    const unsigned iz = min((unsigned)cell_raw(s.az, p.inv_cell[2]), (unsigned)(p.dims[2] - 1));
    s.cell = __ldg(p.cells + ((ix * (unsigned)p.dims[1] + iy) * (unsigned)p.dims[2] + iz));
}

// The last site an atom was found in, kept in registers by the lane that owns the atom.
// While the atom stays inside the site's exclusive ball (squared radius excl2: no other
// site's sphere, nor any periodic image of one, reaches into it) the states follow from this
// one record and no table is read at all.
struct SiteCache {
    double x, y, z;  // float32-staged centre, widened
    double excl2;    // < 0: nothing cached
    int natgrp;      // natural | image axes << 20 | group << 24
};

template <bool kBand>
__device__ __forceinline__ void band_check(const GdGroupRec &g, double d2, bool count_band,
                                           BandCounters &bc) {
    if (kBand) {
        // boundary-band diagnostics: integer pre-filter on the high word of d2, exact test on
        // the rare path
        const int hi = __double2hiint(d2);
        if ((unsigned)(hi - g.band_lo_o) <= (unsigned)g.band_span_o ||
            (unsigned)(hi - g.band_lo_i) <= (unsigned)g.band_span_i) {
            const unsigned b = count_band ? band_diag(d2, &g) : 0u;
            bc.o12 += b & 1u;
            bc.of += (b >> 1) & 1u;
            bc.i12 += (b >> 2) & 1u;
            bc.fi += (b >> 3) & 1u;
        }
    }
}

// Squared distance site -> atom with the reference's arithmetic: float64 differences of the
// float32-staged coordinates, on the axes in `img` the atom's periodic image if it is closer,
// every operation individually rounded, summed x, y, z.
__device__ __forceinline__ double staged_d2(const StatesParams &p, double sx, double sy, double sz,
                                            float ax, float ay, float az, uint32_t img) {
    double dx = __dsub_rn(sx, (double)ax), dy = __dsub_rn(sy, (double)ay), dz = __dsub_rn(sz, (double)az);
    if (img) {  // a periodic image of the atom can be the closer one
#if GD_K1_IMAGE_FILTER
        // ... but only when site and atom sit on opposite sides of the box: the image differs
        // from the atom by +-L (float32 rounded), so with |d| <= L/4 it is at least 3L/4 - ulp
        // away and `fabs(e) < fabs(d)` is false whatever the rounding.  Most lanes whose site
        // lies near a face skip the image arithmetic (and most warps the divergent branch).
        if (!((img & 1u) && fabs(dx) > p.qbox[0])) img &= ~1u;
        if (!((img & 2u) && fabs(dy) > p.qbox[1])) img &= ~2u;
        if (!((img & 4u) && fabs(dz) > p.qbox[2])) img &= ~4u;
        if (img) {
#endif
        asm volatile("" : "+f"(ax), "+f"(ay), "+f"(az));  // keep the image arithmetic in here
        if (img & 1u) dx = axis_delta_image(dx, sx, ax, p.box[0], p.cutoff32);
        if (img & 2u) dy = axis_delta_image(dy, sy, ay, p.box[1], p.cutoff32);
        if (img & 4u) dz = axis_delta_image(dz, sz, az, p.box[2], p.cutoff32);
#if GD_K1_IMAGE_FILTER
        }
#endif
    }
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// Rare path (overlapping spheres): more than one site of the cell contains the atom, so sites
// other than the winner were "visited" too.  Re-walks the cell and flags every hit.
template <bool kOneGroup>
__device__ __noinline__ void mark_all_hits(const StatesParams &p, const GdGroupRec *s_groups,
                                           uint8_t *vis, Staged st) {
    const uint32_t count = st.cell >> 24, payload = st.cell & 0xFFFFFFu;
    for (uint32_t k = 0; k < count; ++k) {
        const uint32_t word = (count == 1) ? payload : __ldg(p.entries + payload + k);
        const float4 rec = __ldg(p.sites + (word & kRankMask));
        const int natgrp = __float_as_int(rec.w);
        const GdGroupRec &g = kOneGroup ? p.g0 : s_groups[natgrp >> 24];
        const double d2 = staged_d2(p, (double)rec.x, (double)rec.y, (double)rec.z, st.ax, st.ay, st.az,
                                    word >> 20);
        const int natural = natgrp & (int)kRankMask;
        if (d2 <= g.r2_outer) vis[natural] = 1;
        if (d2 <= g.r2_inner) vis[p.n_sites_total + natural] = 1;
    }
}

// Site states of one staged atom-frame from the tables: walk the candidate sites of its cell
// in overwrite order, exact float64 test on the float32-staged coordinates, last match wins.
// `multi` is set when a second site contains the atom (then mark_all_hits runs).  The last
// site hit is left in the lane's cache.
template <bool kOneGroup, bool kBand>
__device__ __forceinline__ void classify(const StatesParams &p, const GdGroupRec *s_groups,
                                         const Staged &st, bool count_band, int32_t &best_o,
                                         int32_t &best_i, bool &multi, SiteCache &cache,
                                         BandCounters &bc) {
    best_o = GD_NOSITE;
    best_i = GD_NOSITE;
    multi = false;
    uint32_t n = st.cell >> 24;
    if (n == 0) return;
    const uint32_t payload = st.cell & 0xFFFFFFu;
    const uint32_t *ep = p.entries + payload;
    uint32_t word = (n == 1) ? payload : __ldg(ep);
    for (;;) {
        const uint32_t rank = word & kRankMask;
        const float4 rec = __ldg(p.sites + rank);
        const int natgrp = __float_as_int(rec.w);
        const GdGroupRec &g = kOneGroup ? p.g0 : s_groups[natgrp >> 24];
        const double sx = (double)rec.x, sy = (double)rec.y, sz = (double)rec.z;
        const double d2 = staged_d2(p, sx, sy, sz, st.ax, st.ay, st.az, word >> 20);
        const int natural = natgrp & (int)kRankMask;
        const bool hit_o = d2 <= g.r2_outer, hit_i = d2 <= g.r2_inner;
        multi |= (hit_o & (best_o >= 0)) | (hit_i & (best_i >= 0));
        if (hit_o) best_o = natural;  // entries ascend in overwrite order: last match wins
        if (hit_i) best_i = natural;
        if (hit_o | hit_i) {
            cache.x = sx;
            cache.y = sy;
            cache.z = sz;
            cache.natgrp = natgrp;
            cache.excl2 = (double)__ldg(p.excl2 + rank);
        }
        band_check<kBand>(g, d2, count_band, bc);
        if (--n == 0) break;
        word = __ldg(++ep);
    }
}

template <bool kTma, bool kOneGroup, bool kDiag, bool kBand>
__global__ void __launch_bounds__(kThreads, GD_K1_MIN_CTAS)
    k_site_states(const __grid_constant__ StatesParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    double *ring = reinterpret_cast<double *>(smem);  // [kWarps][kStages][96]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)kWarps * kStages * kRowDoubles * 8);
    GdGroupRec *s_groups = reinterpret_cast<GdGroupRec *>(full + kWarps * kStages);
    uint8_t *s_vis = reinterpret_cast<uint8_t *>(s_groups + kMaxGroups);

    const int tid = threadIdx.x, lane = tid & 31;
    // warp index through a shuffle: tells the compiler it is warp-uniform, so the strip / chunk /
    // row-pointer arithmetic below stays on the uniform datapath
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);

    if (!kOneGroup)
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
        const int r = min(32, p.n_atoms - a0);  // atoms in this strip
        const bool lane_on = lane < r;
        // lanes past the end of a partial strip shadow its last atom (no divergence in the
        // walk); they never store and never count
        const int lsrc = min(lane, r - 1);
        const int32_t a = a0 + lsrc;
        const int64_t t0 = chunk * p.f;
        const int64_t t1 = min(p.n_frames, t0 + (int64_t)p.f);
        const int n_own = (int)(t1 - t0);
        const int64_t A3 = (int64_t)p.n_atoms * 3;
        const double *row0 = p.pos + t0 * A3 + (int64_t)a0 * 3;
        // halo row for the event at the chunk's last frame
        const double *halo_row = nullptr;
        if (p.work != nullptr) {
            if (t1 < p.n_frames)
                halo_row = p.pos + t1 * A3 + (int64_t)a0 * 3;
            else if (p.next_row != nullptr)
                halo_row = p.next_row + (int64_t)a0 * 3;
        }

        // -2 never equals a state: row 0 counts as a change of both (subtracted at the end),
        // which also flags the first state as visited
        int32_t prev_o = -2, prev_i = -2;
        int cnt = 0, n_do = 0, n_di = 0;
        int32_t *po = p.out_o + t0 * p.n_atoms + a;
        int32_t *pi = p.out_i + t0 * p.n_atoms + a;
        uint8_t *vis_g = p.visited;
        const int nst = p.n_sites_total;
        const int vis_mode = p.vis_mode;

        SiteCache cache;
        cache.x = cache.y = cache.z = 0.0;
        cache.excl2 = -1.0;
        cache.natgrp = 0;

        // event rows leave in the same pass (K2's fill then never re-reads the states): this
        // lane's staging region for the chunk, and the frame of the row being processed
        int2 *stage_row = nullptr;
        if (p.stage != nullptr && lane_on)
            stage_row = reinterpret_cast<int2 *>(p.stage) +
                        (chunk * p.n_atoms + a) * (int64_t)p.stage_cap * 3;
        const int stage_cap = p.stage_cap;
        const int32_t t_first = (int32_t)t0;

        // t_row: frame (inside this shard) of the row being processed -- warp-uniform
        auto process = [&](double f0, double f1, double f2, bool own, int32_t t_row) {
            Staged st = stage_atom<kDiag>(p, f0, f1, f2);
            int32_t so, si;
            bool multi = false;
            // fast path: still inside the exclusive ball of the cached site -> its record
            // alone decides both states (no table reads)
            const uint32_t c_img = ((uint32_t)cache.natgrp >> kSiteImgShift) & 7u;
            const double d2c = staged_d2(p, cache.x, cache.y, cache.z, st.ax, st.ay, st.az, c_img);
            if (GD_K1_LIKELY(d2c <= cache.excl2)) {
                const GdGroupRec &g = kOneGroup ? p.g0 : s_groups[cache.natgrp >> 24];
                const int natural = cache.natgrp & (int)kRankMask;
                so = d2c <= g.r2_outer ? natural : GD_NOSITE;
                si = d2c <= g.r2_inner ? natural : GD_NOSITE;
                band_check<kBand>(g, d2c, lane_on, bc);
            } else {
                lookup_cell(p, st);
                classify<kOneGroup, kBand>(p, s_groups, st, lane_on, so, si, multi, cache, bc);
            }
            const bool d_o = so != prev_o, d_i = si != prev_i;
            if (d_o | d_i) {  // everything a change triggers lives in this (divergent) block
                if (d_o) ++n_do;
                if (d_i) ++n_di;
                // cnt counts the chunk's first row as a pseudo event: real event k is cnt == k + 1
                if (stage_row != nullptr && cnt >= 1 && cnt <= stage_cap) {
                    int2 *e = stage_row + 3 * (cnt - 1);
                    e[0] = make_int2(a, prev_o);
                    e[1] = make_int2(so, prev_i);
                    e[2] = make_int2(si, t_row - 1);
                }
                ++cnt;
                // a site is visited when a state switches to it (or in mark_all_hits when
                // several sites contain the atom at once)
                uint8_t *vis = vis_mode == 1 ? s_vis : vis_g;  // generic pointer: shared or global
                if (vis_mode != 0) {
                    if (d_o & (so >= 0)) vis[so] = 1;
                    if (d_i & (si >= 0)) vis[nst + si] = 1;
                }
            }
            if (__builtin_expect(multi && vis_mode != 0, 0))
                mark_all_hits<kOneGroup>(p, s_groups, vis_mode == 1 ? s_vis : vis_g, st);
            prev_o = so;
            prev_i = si;
            if (own) {
                if (lane_on) {
                    *po = so;
                    *pi = si;
                }
                po += p.n_atoms;
                pi += p.n_atoms;
            }
        };

        if (kTma) {
            // rows of an even atom count are 16-byte aligned and r is even: the whole strip row
            // moves as one bulk copy
            const uint32_t row_bytes = (uint32_t)(r * 24);
            double *my_ring = ring + (size_t)warp * kStages * kRowDoubles;
            uint64_t *my_full = full + warp * kStages;
            if (lane == 0) {
                for (int s = 0; s < kStages && s < n_own; ++s) {
                    gd_mbar_arrive_expect_tx(&my_full[s], row_bytes);
                    gd_tma_load_1d(my_ring + s * kRowDoubles, row0 + (int64_t)s * A3, row_bytes, &my_full[s]);
                }
            }
            const double *next_src = row0 + (int64_t)kStages * A3;  // row k + kStages
            const double *my_row = my_ring + 3 * lsrc;
            // ONE copy of the row body in the loop (the kernel is ~2000 SASS instructions; with the
            // ring slots unrolled, four copies of it overflow the instruction cache: +11 % time,
            // `no_instruction` stalls): slot and phase are computed from the row number instead.
            // (Folding the halo row into this loop as row n_own was measured too: 6 % slower.)
            static_assert((kStages & (kStages - 1)) == 0, "kStages must be a power of two");
#pragma unroll 1
            for (int k = 0; k < n_own; ++k) {
                const int s = k & (kStages - 1);
                const uint32_t phase = (uint32_t)(k / kStages) & 1u;
                gd_mbar_wait(&my_full[s], phase);
                const double *sp = my_row + s * kRowDoubles;
                const double f0 = sp[0], f1 = sp[1], f2 = sp[2];
                // The slot is refilled by the async proxy (TMA); the reads above are generic-proxy
                // LDS.  Order them before the refill (observed on B200 without the fence: whole
                // warps read the data of a later row).
                gd_fence_proxy_async();
                __syncwarp();
                if (k + kStages < n_own && lane == 0) {
                    gd_mbar_arrive_expect_tx(&my_full[s], row_bytes);
                    gd_tma_load_1d(my_ring + s * kRowDoubles, next_src, row_bytes, &my_full[s]);
                }
                next_src += A3;
                process(f0, f1, f2, true, t_first + k);
            }
        } else {
            // odd n_atoms: rows are only 8-byte aligned, no bulk copies; register prefetch
            const double *q = row0 + 3 * lsrc;
            double n0 = __ldg(q), n1 = __ldg(q + 1), n2 = __ldg(q + 2);
            for (int k = 0; k < n_own; ++k) {
                const double f0 = n0, f1 = n1, f2 = n2;
                if (k + 1 < n_own) {
                    q += A3;
                    n0 = __ldg(q);
                    n1 = __ldg(q + 1);
                    n2 = __ldg(q + 2);
                }
                process(f0, f1, f2, true, t_first + k);
            }
        }
        if (halo_row != nullptr) {
            const double *q = halo_row + 3 * lsrc;
            process(__ldg(q), __ldg(q + 1), __ldg(q + 2), false, t_first + n_own);
        }
        if (p.work && lane_on) {
            p.work[(int64_t)a * p.n_chunks + chunk] = cnt - 1;
            if (n_do > 1) p.changed[a] = 1;
            if (n_di > 1) p.changed[p.n_atoms + a] = 1;
            if (stage_row != nullptr && cnt - 1 > stage_cap) *p.overflow = 1;
        }
    }

    if (kBand && p.band) {
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

// Squared "exclusive radius" of every ranked site: inside that ball around the (staged) site
// centre, measured with the minimum-image metric, no OTHER site's outer or inner sphere -- nor a
// periodic image of one, nor an image of the site itself -- can contain an atom.  There the
// site's own record decides both states, which is what the kernel's register cache relies on.
//   R(X) = min_Y ( dist(X, Y) - reach(Y) ) - margin,  capped at 4 * the largest reach,
// float32 rounded towards zero; -1 when there is no such ball (duplicate / touching sites).
// GD_K1_NOCACHE=1 disables the cache (every entry -1) for A/B measurements.
static std::vector<float> exclusive_radii2(const std::vector<float> &sx, int32_t n,
                                           const std::vector<int32_t> &order,
                                           const int32_t *h_site_group,
                                           const std::vector<GdGroupRec> &groups, const float *box,
                                           double margin) {
    std::vector<float> out((size_t)std::max(n, 1), -1.0f);
    if (const char *env = getenv("GD_K1_NOCACHE"))
        if (atoi(env) != 0) return out;
    std::vector<double> reach(n);
    double reach_max = 0.0;
    for (int32_t e = 0; e < n; ++e) {
        const GdGroupRec &g = groups[h_site_group[order[e]]];
        reach[e] = std::max(g.r_outer, g.r_inner);
        reach_max = std::max(reach_max, reach[e]);
    }
    const double cap = 4.0 * reach_max;
    const double width = cap + reach_max + margin;  // sites farther than this cannot matter
    int nb[3];
    for (int j = 0; j < 3; ++j) nb[j] = std::min(64, std::max(1, (int)floor((double)box[j] / width)));
    auto bin_of = [&](int32_t e, int j) {
        int b = (int)((double)sx[3 * e + j] / (double)box[j] * nb[j]);
        return std::min(std::max(b, 0), nb[j] - 1);
    };
    std::vector<std::vector<int32_t>> bins((size_t)nb[0] * nb[1] * nb[2]);
    for (int32_t e = 0; e < n; ++e)
        bins[((size_t)bin_of(e, 0) * nb[1] + bin_of(e, 1)) * nb[2] + bin_of(e, 2)].push_back(e);
    for (int32_t e = 0; e < n; ++e) {
        double r = cap;
        for (int j = 0; j < 3; ++j) r = std::min(r, (double)box[j] - reach[e] - margin);  // own images
        int lo[3], cnt[3];
        for (int j = 0; j < 3; ++j) {
            if (nb[j] <= 3) {
                lo[j] = 0;
                cnt[j] = nb[j];
            } else {
                lo[j] = bin_of(e, j) - 1;
                cnt[j] = 3;
            }
        }
        for (int ix = 0; ix < cnt[0]; ++ix)
            for (int iy = 0; iy < cnt[1]; ++iy)
                for (int iz = 0; iz < cnt[2]; ++iz) {
                    const int bx = ((lo[0] + ix) % nb[0] + nb[0]) % nb[0];
                    const int by = ((lo[1] + iy) % nb[1] + nb[1]) % nb[1];
                    const int bz = ((lo[2] + iz) % nb[2] + nb[2]) % nb[2];
                    for (int32_t y : bins[((size_t)bx * nb[1] + by) * nb[2] + bz]) {
                        if (y == e) continue;
                        double d2 = 0.0;
                        for (int j = 0; j < 3; ++j) {
                            const double L = (double)box[j];
                            double d = (double)sx[3 * y + j] - (double)sx[3 * e + j];
                            d -= L * nearbyint(d / L);
                            d2 += d * d;
                        }
                        r = std::min(r, sqrt(d2) - reach[y] - margin);
                    }
                }
        if (r > 0.0) {
            float f = (float)(r * r);
            if ((double)f > r * r) f = nextafterf(f, 0.0f);
            out[e] = f;
        }
    }
    return out;
}


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
    std::vector<uint32_t> site_img((size_t)std::max(n_ranked, 1), 0u);  // axes flagged in any entry
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
                    site_img[e] |= flags;
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
    const size_t off_excl = align256(off_sites + (size_t)n_ranked * 16);
    const size_t off_groups = align256(off_excl + (size_t)n_ranked * 4);
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
        const int32_t packed = order[e] | (int32_t)(site_img[e] << kSiteImgShift) | (h_site_group[order[e]] << 24);
        memcpy(&srec[4 * e + 3], &packed, 4);
    }
    const std::vector<float> excl2 = exclusive_radii2(sx, n_ranked, order, h_site_group, groups, box, margin);
    memcpy(base + off_excl, excl2.data(), (size_t)n_ranked * 4);
    h->off_excl = off_excl;
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
    return gd_site_states_staged(d_pos_frac, n_frames, n_atoms, d_next_pos_row, d_blob, h_blob,
                                 d_states_outer, d_states_inner, d_visited, d_band_counts,
                                 d_event_work, d_changed, nullptr, 0, nullptr, stream);
}

extern "C" int32_t gd_events_stage_cap(int64_t n_frames, int32_t n_atoms) {
    if (n_frames <= 0 || n_atoms <= 0) return 0;
    const int32_t f = gd_chunks(n_frames, n_atoms).f;
    return f / 4 > 8 ? f / 4 : 8;  // room for one event every 4th frame of every atom
}

extern "C" int64_t gd_events_stage_elems(int64_t n_frames, int32_t n_atoms, int32_t stage_cap) {
    if (n_frames <= 0 || n_atoms <= 0 || stage_cap <= 0) return 0;
    return gd_chunks(n_frames, n_atoms).n_chunks * (int64_t)n_atoms * stage_cap * 6;
}

extern "C" gd_status_t gd_site_states_staged(const double *d_pos_frac, int64_t n_frames,
                                             int32_t n_atoms, const double *d_next_pos_row,
                                             const void *d_blob, const void *h_blob,
                                             int32_t *d_states_outer, int32_t *d_states_inner,
                                             uint8_t *d_visited, unsigned long long *d_band_counts,
                                             int32_t *d_event_work, uint8_t *d_changed,
                                             int32_t *d_stage, int32_t stage_cap,
                                             int32_t *d_overflow, void *stream) {
    GD_REQUIRE(d_stage == nullptr || (d_event_work && d_overflow && stage_cap > 0), GD_ERR_INVALID,
               "gd_site_states_staged: staging needs d_event_work, d_overflow and a positive capacity");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_stage) & 7) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states_staged: d_stage must be 8-byte aligned");
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
    int spg_max = kWarps;  // strips per CTA (GD_K1_SPG overrides for tuning: 1, 2, 4, 8)
    if (const char *env = getenv("GD_K1_SPG")) {
        const int v = atoi(env);
        if (v == 1 || v == 2 || v == 4 || v == 8) spg_max = v;
    }
    while ((1 << spg_shift) < spg_max && (1 << spg_shift) < p.n_strips) ++spg_shift;
    p.spg_shift = spg_shift;
    const int cpg = kWarps >> spg_shift;
    p.n_cblocks = gd_ceil_div(geo.n_chunks, cpg);
    const int64_t n_groups_of_strips = gd_ceil_div(p.n_strips, 1 << spg_shift);
    const int64_t grid = n_groups_of_strips * p.n_cblocks;
    GD_REQUIRE(grid < (int64_t(1) << 31), GD_ERR_INVALID, "gd_site_states: shard too large");
    p.sites = reinterpret_cast<const float4 *>(base + h->off_sites);
    p.excl2 = reinterpret_cast<const float *>(base + h->off_excl);
    p.cells = reinterpret_cast<const uint32_t *>(base + h->off_cells);
    p.entries = reinterpret_cast<const uint32_t *>(base + h->off_entries);
    p.groups = reinterpret_cast<const GdGroupRec *>(base + h->off_groups);
    p.out_o = d_states_outer;
    p.out_i = d_states_inner;
    p.visited = d_visited;
    p.band = d_band_counts;
    p.work = d_event_work;
    p.changed = d_changed;
    p.stage = d_stage;
    p.stage_cap = stage_cap;
    p.overflow = d_overflow;
    p.n_sites_total = h->n_sites_total;
    p.n_groups = h->n_groups;
    p.cutoff32 = h->cutoff32;
    for (int j = 0; j < 3; ++j) {
        p.dims[j] = h->dims[j];
        p.box[j] = h->box[j];
        p.inv_cell[j] = h->inv_cell[j];
        p.inv_box[j] = h->inv_box[j];
        p.qbox[j] = 0.25 * (double)h->box[j];
    }
    for (int j = 0; j < 9; ++j) p.m[j] = h->matrix[j];

    const size_t vis_bytes = 2 * (size_t)h->n_sites_total;
    p.vis_mode = d_visited == nullptr ? 0 : (vis_bytes <= (size_t)kVisSmemMax ? 1 : 2);
#ifdef GD_K1_PROBE_NO_VIS  // tools/k1_variants.sh: upper bound of what the visited marks cost
    p.vis_mode = 0;
#endif
    p.g0 = *reinterpret_cast<const GdGroupRec *>(static_cast<const unsigned char *>(h_blob) + h->off_groups);
    const size_t smem = (size_t)kWarps * kStages * kRowDoubles * 8 + kWarps * kStages * sizeof(uint64_t) +
                        kMaxGroups * sizeof(GdGroupRec) +
                        (p.vis_mode == 1 ? ((vis_bytes + 15) & ~size_t(15)) : 0);
    // rows of an odd-sized atom axis start on 8-byte boundaries only: no bulk copies then
    const bool tma = (n_atoms % 2) == 0;
    const bool one = h->n_groups == 1;
    // variant: 0 = diagonal lattice + band counters, 1 = diagonal, no band counters,
    //          2 = general orthorhombic matrix (band code kept)
    const int variant = h->is_diag ? (d_band_counts ? 0 : 1) : 2;
    typedef void (*Kern)(StatesParams);
    static const Kern table[2][2][3] = {
        {{k_site_states<false, false, true, true>, k_site_states<false, false, true, false>,
          k_site_states<false, false, false, true>},
         {k_site_states<false, true, true, true>, k_site_states<false, true, true, false>,
          k_site_states<false, true, false, true>}},
        {{k_site_states<true, false, true, true>, k_site_states<true, false, true, false>,
          k_site_states<true, false, false, true>},
         {k_site_states<true, true, true, true>, k_site_states<true, true, true, false>,
          k_site_states<true, true, false, true>}}};
    const Kern kern = table[tma ? 1 : 0][one ? 1 : 0][variant];
    GD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
