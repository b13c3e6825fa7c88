// K1 -- site-state assignment for every (frame, floating atom).
//
// Replaces _calculate_atom_states (reference src/gemdat/transitions.py:736-809): the
// MDAnalysis PeriodicKDTree build + query (float32 wrap, periodic images,
// scipy cKDTree ball query) and the last-write-wins scatter, for the outer and the
// inner sphere in one pass over the positions.
//
// B200 design
//  * positions stream HBM -> shared memory through a 1-D TMA bulk-copy ring
//    (cp.async.bulk + mbarrier, SASS UBLKCP); every thread then owns one
//    atom-frame (3 doubles, bank-conflict free at a 24-byte stride);
//  * the site list (16 B / site) and the per-label radii live in shared memory;
//  * instead of a KD-tree, a uniform cell grid over the wrapped float32 box maps a
//    position to the few sites whose sphere can reach that cell (conservative
//    superset built on the host, L1/L2 resident); the exact test below decides;
//  * the exact test reproduces the reference's arithmetic bit for bit:
//    float64 frac*L -> float32, float32 wrap, float32 periodic image (+L if
//    x <= cutoff else -L if L-x <= cutoff), float64 dx*dx+dy*dy+dz*dz <= r*r with
//    every operation individually rounded (no FMA);
//  * persistent CTAs, each walking a contiguous range of 512-atom-frame chunks so
//    the cell records of "its" frames stay in L1.
//
// Algorithmic traffic: 24 B read + 8 B written per atom-frame.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "gd_common.cuh"

namespace {

constexpr uint32_t kMagic = 0x47534447u;  // "GDSG"
constexpr int kMaxGroups = 64;
constexpr int kThreads = 512;  // atom-frames per pipeline stage
constexpr int kStages = 4;
constexpr float kNoImage = 1.0e30f;

struct GdGridHeader {  // 256 bytes, little endian, position independent
    uint32_t magic;
    uint32_t version;
    int32_t n_sites;   // sites that belong to a group (= rows of the site table)
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

struct GdGroupRec {  // 32 bytes
    double r2_outer;
    double r2_inner;
    double r_outer;
    double r_inner;
};

inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// MDAnalysis _ortho_pbc, float32 (lib/src/calc_distances.h) [recalled]
__host__ __device__ inline float wrap_ortho(float c, float L, double invL) {
#ifdef __CUDA_ARCH__
#define GD_FADD(a, b) __fadd_rn((a), (b))
#define GD_FSUB(a, b) __fsub_rn((a), (b))
#define GD_FMUL(a, b) __fmul_rn((a), (b))
#else
#define GD_FADD(a, b) ((a) + (b))
#define GD_FSUB(a, b) ((a) - (b))
#define GD_FMUL(a, b) ((a) * (b))
#endif
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
#undef GD_FADD
#undef GD_FSUB
#undef GD_FMUL
}

struct StatesParams {
    const double *pos;
    int64_t n_af;
    int64_t n_chunks;
    const float4 *sites;     // [n_sites] x, y, z, bits(natural | group << 24)
    const uint2 *cells;      // [ncell] {count, payload}
    const int32_t *entries;  // ranks, ascending inside a cell
    const GdGroupRec *groups;
    int32_t *out_o;
    int32_t *out_i;
    uint8_t *visited;  // [2*n_sites_total] or null
    unsigned long long *band;
    int32_t n_sites;
    int32_t n_sites_total;
    int32_t n_groups;
    int32_t dims[3];
    int32_t is_diag;
    float box[3];
    float cutoff32;
    float inv_cell[3];
    double inv_box[3];
    double m[9];
};

template <bool kSmemSites>
__global__ void __launch_bounds__(kThreads, 2)
    k_site_states(const __grid_constant__ StatesParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    double *ring = reinterpret_cast<double *>(smem);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)kStages * kThreads * 24);
    uint64_t *empty = full + kStages;
    GdGroupRec *s_groups = reinterpret_cast<GdGroupRec *>(empty + kStages);
    float4 *s_sites = reinterpret_cast<float4 *>(s_groups + kMaxGroups);

    const int tid = threadIdx.x;

    // contiguous chunk range of this CTA
    const int64_t per = p.n_chunks / gridDim.x;
    const int64_t rem = p.n_chunks % gridDim.x;
    const int64_t bid = blockIdx.x;
    const int64_t chunk0 = bid * per + (bid < rem ? bid : rem);
    const int64_t n_my = per + (bid < rem ? 1 : 0);

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            gd_mbar_init(&full[s], 1);
            gd_mbar_init(&empty[s], kThreads);
        }
        gd_fence_barrier_init();
    }
    for (int g = tid; g < p.n_groups; g += kThreads) s_groups[g] = p.groups[g];
    if (kSmemSites) {
        for (int e = tid; e < p.n_sites; e += kThreads) s_sites[e] = __ldg(p.sites + e);
    }
    __syncthreads();

    auto issue = [&](int64_t local_chunk, int slot) {
        const int64_t chunk = chunk0 + local_chunk;
        const int64_t first = chunk * kThreads;
        int64_t valid = p.n_af - first;
        if (valid > kThreads) valid = kThreads;
        const uint32_t bytes = (uint32_t)(valid * 24) & ~15u;
        gd_mbar_arrive_expect_tx(&full[slot], bytes);
        gd_tma_load_1d(ring + (size_t)slot * kThreads * 3, p.pos + first * 3, bytes, &full[slot]);
    };

    if (tid == 0) {
        for (int s = 0; s < kStages && s < n_my; ++s) issue(s, s);
    }

    const float Lx = p.box[0], Ly = p.box[1], Lz = p.box[2];
    const float r32 = p.cutoff32;
    const int dimx = p.dims[0], dimy = p.dims[1], dimz = p.dims[2];
    const float4 *sites = kSmemSites ? s_sites : p.sites;

    unsigned int band_o12 = 0, band_i12 = 0, band_of = 0, band_if = 0;

    for (int64_t i = 0; i < n_my; ++i) {
        const int slot = (int)(i % kStages);
        const uint32_t phase = (uint32_t)((i / kStages) & 1);

        // producer duty: refill the slot everybody consumed in the previous iteration
        if (tid == 0 && i >= 1 && (i - 1 + kStages) < n_my) {
            const int ps = (int)((i - 1) % kStages);
            const uint32_t pp = (uint32_t)(((i - 1) / kStages) & 1);
            gd_mbar_wait(&empty[ps], pp);
            issue(i - 1 + kStages, ps);
        }

        gd_mbar_wait(&full[slot], phase);

        const int64_t n = (chunk0 + i) * kThreads + tid;
        const bool active = n < p.n_af;
        int64_t valid = p.n_af - (chunk0 + i) * kThreads;
        if (valid > kThreads) valid = kThreads;
        const int covered = (int)(((uint32_t)(valid * 24) & ~15u) / 24);  // whole triples in smem

        double f0 = 0.0, f1 = 0.0, f2 = 0.0;
        if (active) {
            if (tid < covered) {
                const double *sp = ring + (size_t)slot * kThreads * 3 + 3 * tid;
                f0 = sp[0];
                f1 = sp[1];
                f2 = sp[2];
            } else {  // odd tail element: its last 8 bytes were not part of the bulk copy
                const double *gp = p.pos + n * 3;
                f0 = __ldg(gp);
                f1 = __ldg(gp + 1);
                f2 = __ldg(gp + 2);
            }
        }
        // The slot is refilled by the async proxy (TMA) once all 512 arrivals are in.  The
        // reads above are generic-proxy LDS: without this cross-proxy fence the arrive can
        // overtake a still-queued LDS and the refill lands first (observed on B200: whole
        // warps saw the data of chunk i+4).
        gd_fence_proxy_async();
        gd_mbar_arrive(&empty[slot]);
        if (!active) continue;

        // ---- reference staging: float64 Cartesian -> float32 -> wrap -> images
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
        const float ax = wrap_ortho(__double2float_rn(cx), Lx, p.inv_box[0]);
        const float ay = wrap_ortho(__double2float_rn(cy), Ly, p.inv_box[1]);
        const float az = wrap_ortho(__double2float_rn(cz), Lz, p.inv_box[2]);

        const bool lox = ax <= r32, loy = ay <= r32, loz = az <= r32;
        const float bx = lox ? __fadd_rn(ax, Lx) : ((__fsub_rn(Lx, ax) <= r32) ? __fsub_rn(ax, Lx) : kNoImage);
        const float by = loy ? __fadd_rn(ay, Ly) : ((__fsub_rn(Ly, ay) <= r32) ? __fsub_rn(ay, Ly) : kNoImage);
        const float bz = loz ? __fadd_rn(az, Lz) : ((__fsub_rn(Lz, az) <= r32) ? __fsub_rn(az, Lz) : kNoImage);

        const double a0x = (double)ax, a0y = (double)ay, a0z = (double)az;
        const double a1x = (double)bx, a1y = (double)by, a1z = (double)bz;

        // ---- candidate sites of this cell
        int ix = (int)(ax * p.inv_cell[0]);
        int iy = (int)(ay * p.inv_cell[1]);
        int iz = (int)(az * p.inv_cell[2]);
        ix = min(max(ix, 0), dimx - 1);
        iy = min(max(iy, 0), dimy - 1);
        iz = min(max(iz, 0), dimz - 1);
        const uint2 cell = __ldg(p.cells + ((int64_t)ix * dimy + iy) * dimz + iz);

        int32_t best_o = GD_NOSITE, best_i = GD_NOSITE;
        for (uint32_t k = 0; k < cell.x; ++k) {
            const int e = (cell.x == 1) ? (int)cell.y : __ldg(p.entries + cell.y + k);
            const float4 rec = sites[e];
            const int natgrp = __float_as_int(rec.w);
            const GdGroupRec g = s_groups[natgrp >> 24];
            const double sx = (double)rec.x, sy = (double)rec.y, sz = (double)rec.z;
            double dx = __dsub_rn(sx, a0x), ex = __dsub_rn(sx, a1x);
            double dy = __dsub_rn(sy, a0y), ey = __dsub_rn(sy, a1y);
            double dz = __dsub_rn(sz, a0z), ez = __dsub_rn(sz, a1z);
            const double qx = fmin(__dmul_rn(dx, dx), __dmul_rn(ex, ex));
            const double qy = fmin(__dmul_rn(dy, dy), __dmul_rn(ey, ey));
            const double qz = fmin(__dmul_rn(dz, dz), __dmul_rn(ez, ez));
            const double d2 = __dadd_rn(__dadd_rn(qx, qy), qz);
            const int natural = natgrp & 0x00ffffff;
            if (d2 <= g.r2_outer) {
                best_o = natural;  // entries ascend in overwrite order: last match wins
                if (p.visited && !p.visited[natural]) p.visited[natural] = 1;
            }
            if (d2 <= g.r2_inner) {
                best_i = natural;
                if (p.visited && !p.visited[p.n_sites_total + natural])
                    p.visited[p.n_sites_total + natural] = 1;
            }
            // boundary-band diagnostics (rare path)
            if (fabs(d2 - g.r2_outer) < g.r2_outer * 1e-3) {
                const double dd = fabs(sqrt(d2) - g.r_outer);
                band_o12 += dd < 1e-12;
                band_of += dd < 4e-6;
            }
            if (fabs(d2 - g.r2_inner) < g.r2_inner * 1e-3) {
                const double dd = fabs(sqrt(d2) - g.r_inner);
                band_i12 += dd < 1e-12;
                band_if += dd < 4e-6;
            }
        }
        p.out_o[n] = best_o;
        p.out_i[n] = best_i;
    }

    if (p.band) {
        const int o12 = gd_warp_sum((int)band_o12), i12 = gd_warp_sum((int)band_i12);
        const int of = gd_warp_sum((int)band_of), fi = gd_warp_sum((int)band_if);
        if ((tid & 31) == 0) {
            if (o12) atomicAdd(p.band + 0, (unsigned long long)o12);
            if (i12) atomicAdd(p.band + 1, (unsigned long long)i12);
            if (of) atomicAdd(p.band + 2, (unsigned long long)of);
            if (fi) atomicAdd(p.band + 3, (unsigned long long)fi);
        }
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
    GD_REQUIRE(n_sites > 0 && n_sites < (1 << 24), GD_ERR_INVALID,
               "gd_sitegrid_build: n_sites must be in [1, 2^24)");
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
    double reach_max = 0.0;
    for (int g = 0; g < n_groups; ++g) {
        const double ro = h_group_radius[g] * 1.0;
        const double ri = h_group_radius[g] * h_group_inner[g];
        GD_REQUIRE(!(cutoff < ri), GD_ERR_RADIUS_GT_CUTOFF,
                   "Set cutoff greater or equal to the radius.");
        groups[g] = GdGroupRec{ro * ro, ri * ri, ro, ri};
        reach_max = std::max(reach_max, std::max(ro, ri));
    }

    float box[3];
    double inv_box[3];
    for (int i = 0; i < 3; ++i) {
        box[i] = (float)len[i];
        inv_box[i] = 1.0 / (double)box[i];
    }

    // rank order = overwrite order of the reference: groups in dict order, ascending
    // site index inside a group
    std::vector<int32_t> order;
    order.reserve(n_sites);
    for (int g = 0; g < n_groups; ++g)
        for (int32_t s = 0; s < n_sites; ++s)
            if (h_site_group[s] == g) order.push_back(s);
    for (int32_t s = 0; s < n_sites; ++s)
        GD_REQUIRE(h_site_group[s] >= -1 && h_site_group[s] < n_groups, GD_ERR_INVALID,
                   "gd_sitegrid_build: site %d has group %d", s, h_site_group[s]);
    const int32_t n_ranked = (int32_t)order.size();

    // float32 staging of the site centres (float64 frac @ M -> float32 -> wrap)
    std::vector<float> sx(n_ranked * 3);
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

    // grid resolution: ~0.4 * cutoff per cell, at most 256 cells per axis / 2^22 cells
    int dims[3];
    double target = std::max(0.4 * cutoff, 1e-3);
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

    // cells reachable from each site (periodic), margin covers every float32 rounding
    const double margin = 2e-3;
    auto axis_cells = [&](double s, double reach, int j, std::vector<int> &out) {
        out.clear();
        const double w = (double)box[j] / dims[j];
        long lo = (long)floor((s - reach) / w) - 1;  // one extra cell each side: the device
        long hi = (long)floor((s + reach) / w) + 1;  // computes the cell index in float32
        if (hi - lo + 1 >= dims[j]) {
            for (int i = 0; i < dims[j]; ++i) out.push_back(i);
            return;
        }
        for (long i = lo; i <= hi; ++i) {
            long c = i % dims[j];
            if (c < 0) c += dims[j];
            out.push_back((int)c);
        }
    };

    // pass 1: count candidates per cell
    std::vector<uint32_t> count(ncell, 0);
    std::vector<int> cxs, cys, czs;
    auto for_each_cell = [&](int32_t e, auto &&fn) {
        const GdGroupRec &g = groups[h_site_group[order[e]]];
        const double reach = std::max(g.r_outer, g.r_inner) + margin;
        axis_cells(sx[3 * e + 0], reach, 0, cxs);
        axis_cells(sx[3 * e + 1], reach, 1, cys);
        axis_cells(sx[3 * e + 2], reach, 2, czs);
        for (int a : cxs)
            for (int b : cys)
                for (int c : czs) fn(((int64_t)a * dims[1] + b) * dims[2] + c);
    };
    for (int32_t e = 0; e < n_ranked; ++e) for_each_cell(e, [&](int64_t cell) { ++count[cell]; });

    std::vector<uint32_t> start(ncell + 1, 0);
    for (int64_t c = 0; c < ncell; ++c) start[c + 1] = start[c] + count[c];
    const size_t n_entries = start[ncell];

    const size_t off_sites = sizeof(GdGridHeader);
    const size_t off_groups = align256(off_sites + (size_t)n_ranked * 16);
    const size_t off_cells = align256(off_groups + (size_t)n_groups * sizeof(GdGroupRec));
    const size_t off_entries = align256(off_cells + (size_t)ncell * 8);
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
    h->version = 1;
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
    h->cutoff32 = (float)cutoff;
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

    // pass 2: fill; sites are visited in ascending rank so every cell list ascends too
    int32_t *entries = reinterpret_cast<int32_t *>(base + off_entries);
    std::vector<uint32_t> cursor(start.begin(), start.end() - 1);
    for (int32_t e = 0; e < n_ranked; ++e)
        for_each_cell(e, [&](int64_t cell) { entries[cursor[cell]++] = e; });

    uint32_t *cells = reinterpret_cast<uint32_t *>(base + off_cells);
    uint32_t maxc = 0;
    for (int64_t c = 0; c < ncell; ++c) {
        cells[2 * c] = count[c];
        cells[2 * c + 1] = count[c] == 1 ? (uint32_t)entries[start[c]] : start[c];
        maxc = std::max(maxc, count[c]);
    }
    h->max_count = (int32_t)maxc;
    return GD_OK;
}

// ------------------------------------------------------------------------------------ launch

extern "C" gd_status_t gd_site_states(const double *d_pos_frac, int64_t n_atom_frames,
                                      const void *d_blob, const void *h_blob,
                                      int32_t *d_states_outer, int32_t *d_states_inner,
                                      uint8_t *d_visited, unsigned long long *d_band_counts,
                                      void *stream) {
    GD_REQUIRE(d_pos_frac && d_blob && h_blob && d_states_outer && d_states_inner, GD_ERR_INVALID,
               "gd_site_states: null argument");
    GD_REQUIRE(n_atom_frames >= 0, GD_ERR_INVALID, "gd_site_states: negative size");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_pos_frac) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states: d_pos_frac must be 16-byte aligned");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_blob) & 255) == 0, GD_ERR_ALIGNMENT,
               "gd_site_states: d_blob must be 256-byte aligned");
    const GdGridHeader *h = static_cast<const GdGridHeader *>(h_blob);
    GD_REQUIRE(h->magic == kMagic && h->version == 1, GD_ERR_INVALID,
               "gd_site_states: h_blob is not a site grid");
    if (n_atom_frames == 0) return GD_OK;
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");

    const unsigned char *base = static_cast<const unsigned char *>(d_blob);
    StatesParams p;
    memset(&p, 0, sizeof(p));
    p.pos = d_pos_frac;
    p.n_af = n_atom_frames;
    p.n_chunks = gd_ceil_div(n_atom_frames, kThreads);
    p.sites = reinterpret_cast<const float4 *>(base + h->off_sites);
    p.cells = reinterpret_cast<const uint2 *>(base + h->off_cells);
    p.entries = reinterpret_cast<const int32_t *>(base + h->off_entries);
    p.groups = reinterpret_cast<const GdGroupRec *>(base + h->off_groups);
    p.out_o = d_states_outer;
    p.out_i = d_states_inner;
    p.visited = d_visited;
    p.band = d_band_counts;
    p.n_sites = h->n_sites;
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

    const size_t fixed = (size_t)kStages * kThreads * 24 + 2 * kStages * sizeof(uint64_t) +
                         kMaxGroups * sizeof(GdGroupRec);
    const size_t with_sites = fixed + (size_t)h->n_sites * 16;
    // keep two CTAs per SM when the site table fits twice, else one; fall back to global
    const bool smem_sites = with_sites <= (size_t)dev.max_smem_optin;
    const size_t smem = smem_sites ? with_sites : fixed;
    auto kern = smem_sites ? k_site_states<true> : k_site_states<false>;
    GD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    GD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem));
    GD_REQUIRE(per_sm >= 1, GD_ERR_CUDA, "gd_site_states: kernel does not fit on an SM");
    int64_t grid = (int64_t)dev.sm_count * per_sm;
    if (grid > p.n_chunks) grid = p.n_chunks;
    kern<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
