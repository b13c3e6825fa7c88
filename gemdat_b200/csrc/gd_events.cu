// K2 -- transition events and forward/backward fill over the [n_frames, n_atoms]
// int32 state arrays produced by K1.
//
// Replaces _calculate_transition_events (reference src/gemdat/transitions.py:413-482:
// a Python loop over atoms with np.roll/np.nonzero/np.unique inside) and
// ffill/bfill (src/gemdat/utils.py:59-98).
//
// The frame axis is cut into chunks of f frames (gd_chunks, shared with K1); one thread
// owns one (chunk, atom) column segment, so a warp reads 32 neighbouring atoms of the
// same frame (one coalesced 128-byte piece per state array per step) and keeps the
// previous row in registers: every state is read exactly once per pass.  Rows come
// out sorted by (atom, time) through a count -> per-atom block scan -> fill sequence;
// the row offsets live in the caller's scratch (atom-major so the scan is contiguous).
// On the fused path the count pass is free: K1 (gd_site_states) produces the per-segment
// counts while the states are still in registers, and only the fill pass re-reads them.
//
// Algorithmic traffic: 8 B read per atom-frame per pass; events are O(1e-2) of the
// atom-frames (24 B each).
#include <stdlib.h>

#include <mutex>

#include "gd_common.cuh"

namespace {

typedef GdChunks Geometry;
inline Geometry geometry(int64_t n_frames, int32_t n_atoms) { return gd_chunks(n_frames, n_atoms); }

// ---------------------------------------------------------------- events: count / fill

__device__ __forceinline__ void put_event(int32_t *__restrict__ events, int64_t row, int32_t a,
                                          int32_t po, int32_t no, int32_t pi, int32_t ni, int32_t t) {
    int2 *e = reinterpret_cast<int2 *>(events + row * 6);  // rows are 24 B: 8-byte aligned
    e[0] = make_int2(a, po);
    e[1] = make_int2(no, pi);
    e[2] = make_int2(ni, t);
}

// One thread owns one (chunk, atom) column segment; a warp reads 32 neighbouring atoms of
// the same frame (one coalesced 128-byte row piece per state array per step) and keeps the
// previous row in registers: every state is read once per pass.
template <bool kFill>
__global__ void __launch_bounds__(256)
    k_events(const int32_t *__restrict__ so, const int32_t *__restrict__ si, Geometry g,
             const int32_t *__restrict__ next_o, const int32_t *__restrict__ next_i,
             int32_t *__restrict__ work, uint8_t *__restrict__ changed,
             const uint8_t *__restrict__ changed_global, const int64_t *__restrict__ atom_offsets,
             int64_t frame0, int32_t *__restrict__ events) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = item < g.n_chunks * g.n_atoms;
    const int64_t c = on ? item / g.n_atoms : 0;
    const int32_t a = on ? (int32_t)(item - c * g.n_atoms) : 0;
    const int64_t t0 = c * g.f;
    const int64_t t1 = min(g.n_frames, t0 + g.f);
    const int64_t A = g.n_atoms;

    int64_t row = 0;
    if (kFill) {
        // rows of this segment: [first, last) in the atom's block of the table
        int32_t first = 0, last = 0;
        if (on && changed_global[a]) {  // outer state never changes: atom is skipped
            const int32_t *w = work + (int64_t)a * g.n_chunks + c;
            first = w[0];
            last = (c + 1 < g.n_chunks) ? w[1] : (int32_t)(atom_offsets[a + 1] - atom_offsets[a]);
            row = atom_offsets[a] + first;
        }
        const bool mine = last > first;
        if (!__any_sync(0xffffffffu, mine)) return;  // nothing happens to these 32 atoms here
        if (!mine) return;
    } else if (!on) {
        return;
    }

    const int32_t *qo = so + t0 * A + a;
    const int32_t *qi = si + t0 * A + a;
    int32_t po = *qo, pi = *qi;
    int cnt = 0;
    bool ch_o = false, ch_i = false;
    // transitions t -> t+1 with both frames inside this shard.  The walk is latency bound (two
    // 128-byte requests per warp and step), so kAhead rows are fetched before they are compared.
    const int64_t t_in = min(t1, g.n_frames - 1);
    constexpr int kAhead = 8;
    auto step = [&](int32_t no, int32_t ni, int64_t t) {
        const bool d_o = no != po, d_i = ni != pi;
        if (d_o | d_i) {
            if (kFill) put_event(events, row++, a, po, no, pi, ni, (int32_t)(frame0 + t));
            ++cnt;
        }
        ch_o |= d_o;
        ch_i |= d_i;
        po = no;
        pi = ni;
    };
    int64_t t = t0;
    for (; t + kAhead <= t_in; t += kAhead) {
        int32_t vo[kAhead], vi[kAhead];
#pragma unroll
        for (int j = 0; j < kAhead; ++j) {
            vo[j] = __ldcs(qo + (int64_t)(j + 1) * A);
            vi[j] = __ldcs(qi + (int64_t)(j + 1) * A);
        }
        qo += kAhead * A;
        qi += kAhead * A;
#pragma unroll
        for (int j = 0; j < kAhead; ++j) step(vo[j], vi[j], t + j);
    }
    for (; t < t_in; ++t) {
        qo += A;
        qi += A;
        step(__ldcs(qo), __ldcs(qi), t);
    }
    // last frame of the shard: needs the first row of the following shard; without one the
    // np.roll wrap-around pseudo event of the reference is dropped (transitions.py:439-452)
    if (t1 == g.n_frames && next_o != nullptr) {
        const int32_t no = next_o[a], ni = next_i[a];
        const bool d_o = no != po, d_i = ni != pi;
        if (d_o | d_i) {
            if (kFill) put_event(events, row++, a, po, no, pi, ni, (int32_t)(frame0 + g.n_frames - 1));
            ++cnt;
        }
        ch_o |= d_o;
        ch_i |= d_i;
    }
    if (!kFill) {
        work[(int64_t)a * g.n_chunks + c] = cnt;
        if (ch_o) changed[a] = 1;
        if (ch_i) changed[A + a] = 1;
    }
}

// Staged event rows (written by K1 as it found them, [chunk, atom, cap] rows) -> the table.
// kLanes lanes copy one (chunk, atom) segment: its rows are contiguous on both sides, so the copy
// moves 8-byte words with coalesced pieces of kLanes * 8 bytes; the shard's first frame is added
// to the time column on the way.
#ifndef GD_COMPACT_LANES
#define GD_COMPACT_LANES 16
#endif
#ifndef GD_COMPACT_ATOM_MAJOR
#define GD_COMPACT_ATOM_MAJOR 0
#endif
constexpr int kCompactLanes = GD_COMPACT_LANES;
#ifndef GD_COMPACT_WARP_DEFAULT
#define GD_COMPACT_WARP_DEFAULT 1
#endif

// where a (chunk, atom) segment's rows go: n_words 8-byte words from `src` to `dst` (0 = nothing)
struct CompactJob {
    const int2 *src;
    int2 *dst;
    int n_words;
};

__device__ __forceinline__ CompactJob compact_job(int64_t item, const int2 *__restrict__ stage,
                                                  int32_t cap, const Geometry &g,
                                                  const int32_t *__restrict__ work,
                                                  const uint8_t *__restrict__ changed_global,
                                                  const int64_t *__restrict__ atom_offsets,
                                                  int2 *__restrict__ events, int64_t row_capacity) {
    CompactJob j;
    j.src = nullptr;
    j.dst = nullptr;
    j.n_words = 0;
    if (item >= g.n_chunks * g.n_atoms) return j;
#if GD_COMPACT_ATOM_MAJOR
    // neighbouring groups copy neighbouring chunks of one atom: contiguous destination
    const int32_t a = (int32_t)(item / g.n_chunks);
    const int64_t c = item - (int64_t)a * g.n_chunks;
#else
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
#endif
    if (!changed_global[a]) return j;  // outer state never changes: atom is skipped
    const int32_t *w = work + (int64_t)a * g.n_chunks + c;
    const int64_t off = atom_offsets[a];
    const int32_t first = w[0];
    const int32_t last = (c + 1 < g.n_chunks) ? w[1] : (int32_t)(atom_offsets[a + 1] - off);
    // a table smaller than the row count (capacity guessed from an earlier run): write nothing
    // out of bounds, the caller sees the count and repeats with the right size
    if (last <= first || off + last > row_capacity) return j;
    j.src = stage + (c * g.n_atoms + a) * (int64_t)cap * 3;
    j.dst = events + (off + first) * 3;
    j.n_words = 3 * (last - first);
    return j;
}

// kLanes lanes own one segment at a time (grid-stride when the grid is capped; then the NEXT
// segment's look-ups -- skip flag -> chunk offsets -> atom offset, a chain of three dependent
// loads -- are issued before the current segment is copied).
__global__ void __launch_bounds__(256)
    k_events_compact(const int2 *__restrict__ stage, int32_t cap, Geometry g,
                     const int32_t *__restrict__ work, const uint8_t *__restrict__ changed_global,
                     const int64_t *__restrict__ atom_offsets, int32_t frame0,
                     int2 *__restrict__ events, int64_t row_capacity) {
    const int sub = threadIdx.x % kCompactLanes;
    const int64_t n_items = g.n_chunks * g.n_atoms;
    const int64_t stride = (int64_t)gridDim.x * (blockDim.x / kCompactLanes);
    int64_t item = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / kCompactLanes;
    CompactJob cur = compact_job(item, stage, cap, g, work, changed_global, atom_offsets, events, row_capacity);
    for (; item < n_items; item += stride) {
        const CompactJob nxt = compact_job(item + stride, stage, cap, g, work, changed_global,
                                           atom_offsets, events, row_capacity);
        const int2 *src = cur.src;
        int2 *dst = cur.dst;
        const int n_words = cur.n_words;
        // (a segment is ~54 words, 3-4 trips of 16 lanes: all of a lane's loads are issued before
        // its first store)
        constexpr int kTrips = 4;
        int k = sub;
        for (; k + (kTrips - 1) * kCompactLanes < n_words; k += kTrips * kCompactLanes) {
            int2 v[kTrips];
#pragma unroll
            for (int j = 0; j < kTrips; ++j) v[j] = __ldcs(src + k + j * kCompactLanes);
#pragma unroll
            for (int j = 0; j < kTrips; ++j) {
                const int kk = k + j * kCompactLanes;
                if (kk % 3 == 2) v[j].y += frame0;  // (destination inner site, time)
                dst[kk] = v[j];
            }
        }
        for (; k < n_words; k += kCompactLanes) {
            int2 v = __ldcs(src + k);
            if (k % 3 == 2) v.y += frame0;
            dst[k] = v;
        }
        cur = nxt;
    }
}

// Warp-cooperative variant (GD_COMPACT_WARP, default): a warp owns 32 consecutive (atom, chunk)
// segments in ATOM-major order.  The look-ups of the 32 segments (skip flag, chunk offsets, atom
// offset -- independent loads, coalesced over the chunks) are issued at once, one per lane; the
// rows of consecutive chunks of an atom are consecutive in the table, so the warp then writes one
// contiguous run of words, lane by lane (32 x 8 B per store instruction), each lane finding the
// segment its word comes from in a small shared table.  Four loads per lane are in flight before
// the first store.  Against the 16-lanes-per-segment kernel: one dependent load latency instead
// of three before the copy starts, full-width stores, no lanes idle in the last trip of a segment.
constexpr int kCwWarps = 8;

__global__ void __launch_bounds__(kCwWarps * 32)
    k_events_compact_warp(const int2 *__restrict__ stage, int32_t cap, Geometry g,
                          const int32_t *__restrict__ work, const uint8_t *__restrict__ changed_global,
                          const int64_t *__restrict__ atom_offsets, int32_t frame0,
                          int2 *__restrict__ events, int64_t row_capacity) {
    __shared__ int s_prefix[kCwWarps][33];
    __shared__ const int2 *s_src[kCwWarps][32];
    __shared__ int2 *s_dst[kCwWarps][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t n_items = g.n_chunks * g.n_atoms;
    const int64_t item = ((int64_t)blockIdx.x * kCwWarps + wid) * 32 + lane;
    int n_words = 0;
    const int2 *src = nullptr;
    int2 *dst = nullptr;
    if (item < n_items) {
        const int32_t a = (int32_t)(item / g.n_chunks);
        const int64_t c = item - (int64_t)a * g.n_chunks;
        const int32_t *w = work + item;  // == work[a * n_chunks + c]
        const bool last_chunk = c + 1 == g.n_chunks;
        // independent loads: all issued before the first use
        const uint8_t keep = changed_global[a];
        const int32_t first = w[0];
        const int32_t nxt = last_chunk ? 0 : w[1];
        const int64_t off = atom_offsets[a];
        const int64_t off1 = atom_offsets[a + 1];
        const int32_t last = last_chunk ? (int32_t)(off1 - off) : nxt;
        if (keep && last > first && off + last <= row_capacity) {
            src = stage + (c * g.n_atoms + a) * (int64_t)cap * 3;
            dst = events + (off + first) * 3;
            n_words = 3 * (last - first);
        }
    }
    int incl = n_words;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    s_prefix[wid][lane + 1] = incl;
    if (lane == 0) s_prefix[wid][0] = 0;
    s_src[wid][lane] = src;
    s_dst[wid][lane] = dst;
    __syncwarp();
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    const int *prefix = s_prefix[wid];
    int j = 0, lo = 0, hi = prefix[1];  // segment of the lane's current word: words [lo, hi)
    constexpr int kAhead = 4;
    for (int w0 = lane; w0 < total; w0 += 32 * kAhead) {
        int2 v[kAhead];
        int2 *to[kAhead];
        int kind[kAhead];
#pragma unroll
        for (int u = 0; u < kAhead; ++u) {
            const int wd = w0 + 32 * u;
            to[u] = nullptr;
            if (wd < total) {
                while (wd >= hi) {
                    ++j;
                    lo = hi;
                    hi = prefix[j + 1];
                }
                const int local = wd - lo;
                v[u] = __ldcs(s_src[wid][j] + local);
                to[u] = s_dst[wid][j] + local;
                kind[u] = local % 3;
            }
        }
#pragma unroll
        for (int u = 0; u < kAhead; ++u) {
            if (to[u] != nullptr) {
                if (kind[u] == 2) v[u].y += frame0;  // (destination inner site, time)
                *to[u] = v[u];
            }
        }
    }
}

// one block per atom: exclusive scan of the chunk counts (zeroed when the atom is
// skipped), total -> totals[a]
__global__ void __launch_bounds__(256)
    k_scan_chunks(int32_t *__restrict__ work, int64_t n_chunks, const uint8_t *__restrict__ changed,
                  int32_t n_atoms, int64_t *__restrict__ totals, int32_t *__restrict__ err_flag) {
    const int a = blockIdx.x;
    int32_t *col = work + (int64_t)a * n_chunks;
    const bool keep = changed[a] != 0;
    if (threadIdx.x == 0 && keep && !changed[n_atoms + a]) atomicExch(err_flag, 1);
    __shared__ int64_t warp_tot[8];
    __shared__ int64_t running_s;
    if (threadIdx.x == 0) running_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n_chunks; base += 256) {
        const int64_t idx = base + threadIdx.x;
        int64_t v = (idx < n_chunks && keep) ? col[idx] : 0;
        int64_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int64_t warp_base = 0;
        for (int w = 0; w < warp; ++w) warp_base += warp_tot[w];
        const int64_t running = running_s;
        if (idx < n_chunks) col[idx] = (int32_t)(running + warp_base + incl - v);
        __syncthreads();
        if (threadIdx.x == 255) running_s = running + warp_base + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[a + 1] = running_s;
}

// single block: in-place exclusive scan over atoms of totals[1..n] -> offsets[0..n]
__global__ void __launch_bounds__(1024) k_scan_atoms(int64_t *__restrict__ off, int32_t n_atoms) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t running_s;
    if (threadIdx.x == 0) {
        running_s = 0;
        off[0] = 0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n_atoms; base += 1024) {
        const int idx = base + threadIdx.x;
        int64_t v = idx < n_atoms ? off[idx + 1] : 0;
        int64_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int64_t warp_base = 0;
        for (int w = 0; w < warp; ++w) warp_base += warp_tot[w];
        const int64_t running = running_s;
        __syncthreads();
        if (idx < n_atoms) off[idx + 1] = running + warp_base + incl;  // inclusive = offset of a+1
        if (threadIdx.x == 1023) running_s = running + warp_base + incl;
        __syncthreads();
    }
}

// ---------------------------------------------------------------- ffill / bfill
constexpr int kFillAhead = 8;

// phase 1: last valid state of every (chunk, atom) segment, in scan direction
__global__ void __launch_bounds__(256)
    k_fill_summary(const int32_t *__restrict__ s, Geometry g, int direction,
                   int32_t *__restrict__ work) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= g.n_chunks * g.n_atoms) return;
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
    const int64_t t0 = c * g.f, t1 = min(g.n_frames, t0 + g.f);
    int32_t last = GD_NOSITE;
    // kFillAhead rows are in flight per thread (the loads do not depend on `last`)
    const int64_t A = g.n_atoms;
    const int64_t step = direction > 0 ? A : -A;
    const int32_t *q = s + (direction > 0 ? t0 : t1 - 1) * A + a;
    int64_t left = t1 - t0;
    for (; left >= kFillAhead; left -= kFillAhead) {
        int32_t v[kFillAhead];
#pragma unroll
        for (int j = 0; j < kFillAhead; ++j) v[j] = __ldcs(q + j * step);
        q += kFillAhead * step;
#pragma unroll
        for (int j = 0; j < kFillAhead; ++j) last = v[j] != GD_NOSITE ? v[j] : last;
    }
    for (; left > 0; --left, q += step) {
        const int32_t v = __ldcs(q);
        last = v != GD_NOSITE ? v : last;
    }
    work[c * g.n_atoms + a] = last;
}

// phase 2: one WARP per atom turns the chunk summaries into carry-ins: an exclusive scan over
// the chunks in scan direction with the operator "the later valid state wins" (32 chunks per
// step through shuffles; a single thread walking ~10^3-10^4 chunks took as long as a pass over
// the states)
__global__ void __launch_bounds__(128)
    k_fill_carry(Geometry g, int direction, const int32_t *__restrict__ carry_in,
                 int32_t *__restrict__ carry_out, int32_t *__restrict__ work) {
    const int32_t a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (a >= g.n_atoms) return;
    int32_t carry = carry_in ? carry_in[a] : GD_NOSITE;
    for (int64_t p0 = 0; p0 < g.n_chunks; p0 += 32) {
        const int64_t p = p0 + lane;  // position in scan order
        const bool on = p < g.n_chunks;
        const int64_t c = direction > 0 ? p : g.n_chunks - 1 - p;
        int32_t *slot = work + c * g.n_atoms + a;
        int32_t incl = on ? *slot : GD_NOSITE;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o && incl == GD_NOSITE) incl = up;
        }
        int32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) excl = GD_NOSITE;
        if (on) *slot = excl != GD_NOSITE ? excl : carry;
        const int32_t last = __shfl_sync(0xffffffffu, incl, 31);
        carry = last != GD_NOSITE ? last : carry;
    }
    if (carry_out && lane == 0) carry_out[a] = carry;
}

// phase 3: apply
__global__ void __launch_bounds__(256)
    k_fill_apply(const int32_t *__restrict__ s, Geometry g, int direction,
                 const int32_t *__restrict__ work, const int32_t *__restrict__ shard_carry,
                 int32_t *__restrict__ out) {
    const int64_t item = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= g.n_chunks * g.n_atoms) return;
    const int64_t c = item / g.n_atoms;
    const int32_t a = (int32_t)(item - c * g.n_atoms);
    const int64_t t0 = c * g.f, t1 = min(g.n_frames, t0 + g.f);
    int32_t last = work[c * g.n_atoms + a];
    // chunk carries computed without knowing the neighbouring shard: where no valid state was
    // seen yet inside this shard, the state carried over the shard boundary applies
    if (shard_carry && last == GD_NOSITE) last = shard_carry[a];
    const int64_t A = g.n_atoms;
    const int64_t step = direction > 0 ? A : -A;
    const int64_t first = (direction > 0 ? t0 : t1 - 1) * A + a;
    const int32_t *q = s + first;
    int32_t *o = out + first;
    int64_t left = t1 - t0;
    for (; left >= kFillAhead; left -= kFillAhead) {
        int32_t v[kFillAhead];
#pragma unroll
        for (int j = 0; j < kFillAhead; ++j) v[j] = __ldcs(q + j * step);
        q += kFillAhead * step;
#pragma unroll
        for (int j = 0; j < kFillAhead; ++j) {
            last = v[j] != GD_NOSITE ? v[j] : last;
            __stcs(o + j * step, last);
        }
        o += kFillAhead * step;
    }
    for (; left > 0; --left, q += step, o += step) {
        const int32_t v = __ldcs(q);
        last = v != GD_NOSITE ? v : last;
        __stcs(o, last);
    }
}

// one block per atom: copy this shard's rows of the atom to their place in the merged table
__global__ void __launch_bounds__(128)
    k_merge_rows(const int32_t *__restrict__ src, const int64_t *__restrict__ src_off,
                 const int64_t *__restrict__ dst_start, int32_t *__restrict__ dst, int64_t src_cap,
                 int64_t dst_cap) {
    const int a = blockIdx.x;
    const int64_t r0 = src_off[a], r1 = src_off[a + 1], d0 = dst_start[a];
    // offsets that were not validated by the host yet (optimistic pass): never leave the tables
    if (r0 < 0 || r1 < r0 || r1 > src_cap || d0 < 0 || d0 + (r1 - r0) > dst_cap) return;
    const int64_t n = (r1 - r0) * 6;
    const int32_t *s = src + r0 * 6;
    int32_t *d = dst + dst_start[a] * 6;
    for (int64_t k = threadIdx.x; k < n; k += blockDim.x) d[k] = s[k];
}

// block (atom, shard): place shard r's jump rows of the atom in the merged (atom, time) table.
// `gathered` is the all-gather of every shard's buffer [head int64 words: per-atom row offsets
// [n_atoms + 1] ... | rows]; offs_r[a] = rows of shard r before atom a, so the merged table has
//   first row of atom a            = sum_r offs_r[a]
//   shard r's rows of a start at   that + sum_{r' < r} (offs_r'[a + 1] - offs_r'[a])
// -- the whole placement plan is a handful of loads per block, no scan and no host round trip.
__global__ void __launch_bounds__(128)
    k_merge_shards(const int32_t *__restrict__ gathered, int64_t stride_words, int32_t head_words,
                   int32_t world, int32_t n_atoms, int64_t src_cap, int32_t *__restrict__ dst,
                   int64_t dst_cap) {
    const int a = blockIdx.x, r = blockIdx.y;
    int64_t start = 0;
    for (int q = 0; q < world; ++q) {
        const int64_t *off = reinterpret_cast<const int64_t *>(gathered + q * stride_words);
        const int64_t o0 = off[a];
        start += o0;
        if (q < r) start += off[a + 1] - o0;
    }
    const int64_t *mine = reinterpret_cast<const int64_t *>(gathered + r * stride_words);
    const int64_t r0 = mine[a], r1 = mine[a + 1];
    // offsets the host has not validated yet (optimistic pass): never leave the tables
    if (r0 < 0 || r1 < r0 || r1 > src_cap || start < 0 || start + (r1 - r0) > dst_cap) return;
    const int32_t *src = gathered + r * stride_words + head_words + r0 * 6;
    int32_t *d = dst + start * 6;
    const int64_t n = (r1 - r0) * 6;
    for (int64_t k = threadIdx.x; k < n; k += blockDim.x) d[k] = src[k];
}

inline unsigned blocks_for(int64_t items, int threads) { return (unsigned)gd_ceil_div(items, threads); }

}  // namespace

extern "C" int64_t gd_events_work_elems(int64_t n_frames, int32_t n_atoms) {
    if (n_frames <= 0 || n_atoms <= 0) return 16;
    Geometry g = geometry(n_frames, n_atoms);
    return g.n_chunks * g.n_atoms + 16;  // + error flag / padding
}

extern "C" gd_status_t gd_events_count(const int32_t *d_states_outer, const int32_t *d_states_inner,
                                       int64_t n_frames, int32_t n_atoms,
                                       const int32_t *d_next_outer, const int32_t *d_next_inner,
                                       int32_t *d_work, uint8_t *d_changed, void *stream) {
    GD_REQUIRE(d_states_outer && d_states_inner && d_work && d_changed, GD_ERR_INVALID,
               "gd_events_count: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_events_count: empty input");
    GD_REQUIRE((d_next_outer == nullptr) == (d_next_inner == nullptr), GD_ERR_INVALID,
               "gd_events_count: halo rows must both be given or both be NULL");
    Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_events<false><<<blocks_for(items, 256), 256, 0, (cudaStream_t)stream>>>(
        d_states_outer, d_states_inner, g, d_next_outer, d_next_inner, d_work, d_changed, nullptr,
        nullptr, 0, nullptr);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_scan(int32_t *d_work, int64_t n_frames, int32_t n_atoms,
                                      const uint8_t *d_changed_global, int64_t *d_atom_offsets,
                                      int64_t *h_n_events, void *stream) {
    GD_REQUIRE(d_work && d_changed_global && d_atom_offsets, GD_ERR_INVALID,
               "gd_events_scan: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_events_scan: empty input");
    cudaStream_t st = (cudaStream_t)stream;
    Geometry g = geometry(n_frames, n_atoms);
    int32_t *err_flag = d_work + g.n_chunks * g.n_atoms;
    GD_CUDA_CHECK(cudaMemsetAsync(err_flag, 0, sizeof(int32_t), st));
    k_scan_chunks<<<(unsigned)n_atoms, 256, 0, st>>>(d_work, g.n_chunks, d_changed_global, n_atoms,
                                                    d_atom_offsets, err_flag);
    GD_LAUNCH_CHECK();
    k_scan_atoms<<<1, 1024, 0, st>>>(d_atom_offsets, n_atoms);
    GD_LAUNCH_CHECK();
    // asynchronous use: the total stays in d_atom_offsets[n_atoms], the INNER_CONSTANT flag in
    // d_work[gd_events_work_elems() - 16]; the caller reads both later
    if (h_n_events == nullptr) return GD_OK;
    int32_t h_err = 0;
    GD_CUDA_CHECK(cudaMemcpyAsync(h_n_events, d_atom_offsets + n_atoms, sizeof(int64_t),
                                  cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaMemcpyAsync(&h_err, err_flag, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    GD_CUDA_CHECK(cudaStreamSynchronize(st));
    GD_REQUIRE(h_err == 0, GD_ERR_INNER_CONSTANT,
               "index -1 is out of bounds for axis 0 with size 0");
    return GD_OK;
}

extern "C" gd_status_t gd_events_fill(const int32_t *d_states_outer, const int32_t *d_states_inner,
                                      int64_t n_frames, int32_t n_atoms,
                                      const int32_t *d_next_outer, const int32_t *d_next_inner,
                                      const int32_t *d_work, const uint8_t *d_changed_global,
                                      const int64_t *d_atom_offsets, int64_t frame0,
                                      int32_t *d_events, void *stream) {
    GD_REQUIRE(d_states_outer && d_states_inner && d_work && d_changed_global && d_atom_offsets &&
                   d_events,
               GD_ERR_INVALID, "gd_events_fill: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_events_fill: empty input");
    Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_events<true><<<blocks_for(items, 256), 256, 0, (cudaStream_t)stream>>>(
        d_states_outer, d_states_inner, g, d_next_outer, d_next_inner,
        const_cast<int32_t *>(d_work), nullptr, d_changed_global, d_atom_offsets, frame0, d_events);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_compact(const int32_t *d_stage, int32_t stage_cap, int64_t n_frames,
                                         int32_t n_atoms, const int32_t *d_work,
                                         const uint8_t *d_changed_global,
                                         const int64_t *d_atom_offsets, int64_t frame0,
                                         int32_t *d_events, int64_t row_capacity, void *stream) {
    GD_REQUIRE(d_stage && d_work && d_changed_global && d_atom_offsets && d_events, GD_ERR_INVALID,
               "gd_events_compact: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0 && stage_cap > 0, GD_ERR_INVALID,
               "gd_events_compact: empty input");
    GD_REQUIRE(((reinterpret_cast<uintptr_t>(d_stage) | reinterpret_cast<uintptr_t>(d_events)) & 7) == 0,
               GD_ERR_ALIGNMENT, "gd_events_compact: tables must be 8-byte aligned");
    GD_REQUIRE(frame0 >= 0 && frame0 + n_frames < (int64_t(1) << 31), GD_ERR_INVALID,
               "gd_events_compact: frame numbers must fit int32");
    Geometry g = geometry(n_frames, n_atoms);
    const GdDeviceInfo &dev = gd_device_info();
    GD_REQUIRE(dev.sm_count > 0, GD_ERR_CUDA, "no CUDA device");
    // One segment per lane group by default: measured on C3 (profiles/r2e_step_ab.txt) a capped,
    // persistent grid with the next segment's look-ups prefetched is SLOWER (8 CTAs/SM: 1.37 ms,
    // 16: 1.13, uncapped: 1.10) -- the copy wants all the parallelism it can get.
    // GD_COMPACT_CTAS_PER_SM caps the grid for A/B runs.
    static const int use_warp = [] {
        const char *e = getenv("GD_COMPACT_WARP");
        return e ? atoi(e) : GD_COMPACT_WARP_DEFAULT;
    }();
    if (use_warp) {
        const int64_t wblocks = gd_ceil_div(g.n_chunks * g.n_atoms, (int64_t)kCwWarps * 32);
        k_events_compact_warp<<<(unsigned)wblocks, kCwWarps * 32, 0, (cudaStream_t)stream>>>(
            reinterpret_cast<const int2 *>(d_stage), stage_cap, g, d_work, d_changed_global,
            d_atom_offsets, (int32_t)frame0, reinterpret_cast<int2 *>(d_events),
            row_capacity > 0 ? row_capacity : INT64_MAX);
        GD_LAUNCH_CHECK();
        return GD_OK;
    }
    int64_t blocks = gd_ceil_div(g.n_chunks * g.n_atoms * kCompactLanes, 256);
    if (const char *e = getenv("GD_COMPACT_CTAS_PER_SM"))
        if (atoi(e) > 0 && blocks > (int64_t)dev.sm_count * atoi(e)) blocks = (int64_t)dev.sm_count * atoi(e);
    k_events_compact<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const int2 *>(d_stage), stage_cap, g, d_work, d_changed_global,
        d_atom_offsets, (int32_t)frame0, reinterpret_cast<int2 *>(d_events),
        row_capacity > 0 ? row_capacity : INT64_MAX);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_fill_states(const int32_t *d_states, int64_t n_frames, int32_t n_atoms,
                                      int direction, const int32_t *d_carry_in,
                                      int32_t *d_carry_out, int32_t *d_work, int32_t *d_out,
                                      void *stream) {
    GD_REQUIRE(d_states && d_work && d_out, GD_ERR_INVALID, "gd_fill_states: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_fill_states: empty input");
    GD_REQUIRE(direction == 1 || direction == -1, GD_ERR_INVALID,
               "gd_fill_states: direction must be +1 (ffill) or -1 (bfill)");
    cudaStream_t st = (cudaStream_t)stream;
    Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_fill_summary<<<blocks_for(items, 256), 256, 0, st>>>(d_states, g, direction, d_work);
    GD_LAUNCH_CHECK();
    k_fill_carry<<<blocks_for((int64_t)n_atoms * 32, 128), 128, 0, st>>>(g, direction, d_carry_in, d_carry_out,
                                                          d_work);
    GD_LAUNCH_CHECK();
    k_fill_apply<<<blocks_for(items, 256), 256, 0, st>>>(d_states, g, direction, d_work, nullptr,
                                                         d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_fill_last_valid(const int32_t *d_states, int64_t n_frames,
                                          int32_t n_atoms, int direction, int32_t *d_work,
                                          int32_t *d_last_valid, void *stream) {
    GD_REQUIRE(d_states && d_work && d_last_valid, GD_ERR_INVALID,
               "gd_fill_last_valid: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_fill_last_valid: empty input");
    GD_REQUIRE(direction == 1 || direction == -1, GD_ERR_INVALID,
               "gd_fill_last_valid: direction must be +1 (ffill) or -1 (bfill)");
    cudaStream_t st = (cudaStream_t)stream;
    Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_fill_summary<<<blocks_for(items, 256), 256, 0, st>>>(d_states, g, direction, d_work);
    GD_LAUNCH_CHECK();
    k_fill_carry<<<blocks_for((int64_t)n_atoms * 32, 128), 128, 0, st>>>(g, direction, nullptr, d_last_valid,
                                                          d_work);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_fill_apply(const int32_t *d_states, int64_t n_frames, int32_t n_atoms,
                                     int direction, const int32_t *d_carry_in,
                                     const int32_t *d_work, int32_t *d_out, void *stream) {
    GD_REQUIRE(d_states && d_work && d_out, GD_ERR_INVALID, "gd_fill_apply: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_fill_apply: empty input");
    GD_REQUIRE(direction == 1 || direction == -1, GD_ERR_INVALID,
               "gd_fill_apply: direction must be +1 (ffill) or -1 (bfill)");
    Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    k_fill_apply<<<blocks_for(items, 256), 256, 0, (cudaStream_t)stream>>>(
        d_states, g, direction, d_work, d_carry_in, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_merge_capped(const int32_t *d_src_rows, const int64_t *d_src_atom_offsets,
                                              const int64_t *d_dst_start, int32_t n_atoms,
                                              int32_t *d_dst_rows, int64_t src_capacity,
                                              int64_t dst_capacity, void *stream) {
    GD_REQUIRE(d_src_atom_offsets && d_dst_start && d_dst_rows && n_atoms > 0, GD_ERR_INVALID,
               "gd_events_merge: bad argument");
    k_merge_rows<<<(unsigned)n_atoms, 128, 0, (cudaStream_t)stream>>>(
        d_src_rows, d_src_atom_offsets, d_dst_start, d_dst_rows, src_capacity, dst_capacity);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_events_merge(const int32_t *d_src_rows, const int64_t *d_src_atom_offsets,
                                       const int64_t *d_dst_start, int32_t n_atoms,
                                       int32_t *d_dst_rows, void *stream) {
    return gd_events_merge_capped(d_src_rows, d_src_atom_offsets, d_dst_start, n_atoms, d_dst_rows,
                                  INT64_MAX, INT64_MAX, stream);
}

extern "C" gd_status_t gd_jumps_merge_shards(const int32_t *d_gathered, int64_t stride_words,
                                             int32_t head_words, int32_t world, int32_t n_atoms,
                                             int64_t src_capacity, int32_t *d_dst_rows,
                                             int64_t dst_capacity, void *stream) {
    GD_REQUIRE(d_gathered && d_dst_rows && world > 0 && world <= 65535 && n_atoms > 0 &&
                   head_words >= 2 * (n_atoms + 1) && stride_words >= head_words,
               GD_ERR_INVALID, "gd_jumps_merge_shards: bad argument");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_gathered) & 7) == 0 && (stride_words & 1) == 0,
               GD_ERR_ALIGNMENT, "gd_jumps_merge_shards: buffers must be 8-byte aligned");
    dim3 grid((unsigned)n_atoms, (unsigned)world);
    k_merge_shards<<<grid, 128, 0, (cudaStream_t)stream>>>(d_gathered, stride_words, head_words, world,
                                                          n_atoms, src_capacity, d_dst_rows,
                                                          dst_capacity);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

// ---------------------------------------------------------------- allocation-free ffill / bfill
//
// gd_fill_states for callers that cannot allocate (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc): the per-(frame chunk, atom) carries live in a
// file-scope device array instead of a caller-provided d_work.  Every GPU that loads the library
// holds the array (16 MiB) until exit.  All calls share it: calls that overlap on the GPU race.
// Nothing in it has to start at zero -- k_fill_summary overwrites every slot the later two
// kernels read -- so it is not cleared; it keeps the last call's carries.
constexpr int64_t kFillStaticElems = GD_FILL_NOALLOC_MAX_WORK_ELEMS;
static __device__ __align__(256) int32_t g_fill_scratch[kFillStaticElems];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_fill_scratch_mutex;
int32_t *g_fill_scratch_ptr[kMaxDevices];  // device address of g_fill_scratch, per device

// In host code `g_fill_scratch` is a host-side stand-in, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
gd_status_t fill_scratch_address(int32_t **out) {
    int dev = -1;
    GD_CUDA_CHECK(cudaGetDevice(&dev));
    GD_REQUIRE(dev >= 0 && dev < kMaxDevices, GD_ERR_INVALID,
               "gd_fill_states_noalloc: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_fill_scratch_mutex);
    if (g_fill_scratch_ptr[dev] == nullptr) {
        void *p = nullptr;
        GD_CUDA_CHECK(cudaGetSymbolAddress(&p, g_fill_scratch));
        GD_REQUIRE(p != nullptr, GD_ERR_CUDA, "gd_fill_states_noalloc: no address for the scratch array");
        g_fill_scratch_ptr[dev] = static_cast<int32_t *>(p);
    }
    *out = g_fill_scratch_ptr[dev];
    return GD_OK;
}

}  // namespace

extern "C" gd_status_t gd_fill_states_noalloc(const int32_t *d_states, int64_t n_frames,
                                              int32_t n_atoms, int direction,
                                              const int32_t *d_carry_in, int32_t *d_carry_out,
                                              int32_t *d_out, void *stream) {
    GD_REQUIRE(d_states && d_out, GD_ERR_INVALID, "gd_fill_states_noalloc: null argument");
    GD_REQUIRE(n_frames > 0 && n_atoms > 0, GD_ERR_INVALID, "gd_fill_states_noalloc: empty input");
    GD_REQUIRE(direction == 1 || direction == -1, GD_ERR_INVALID,
               "gd_fill_states_noalloc: direction must be +1 (ffill) or -1 (bfill)");
    const Geometry g = geometry(n_frames, n_atoms);
    const int64_t items = g.n_chunks * g.n_atoms;
    GD_REQUIRE(items <= kFillStaticElems, GD_ERR_CAPACITY,
               "gd_fill_states_noalloc: %lld frames x %d atoms need %lld scratch slots, the static "
               "array holds %lld (use gd_fill_states with a caller-provided d_work)",
               (long long)n_frames, (int)n_atoms, (long long)items, (long long)kFillStaticElems);
    int32_t *work = nullptr;
    const gd_status_t st_addr = fill_scratch_address(&work);
    if (st_addr != GD_OK) return st_addr;
    cudaStream_t st = (cudaStream_t)stream;
    k_fill_summary<<<blocks_for(items, 256), 256, 0, st>>>(d_states, g, direction, work);
    GD_LAUNCH_CHECK();
    k_fill_carry<<<blocks_for((int64_t)n_atoms * 32, 128), 128, 0, st>>>(g, direction, d_carry_in,
                                                                         d_carry_out, work);
    GD_LAUNCH_CHECK();
    k_fill_apply<<<blocks_for(items, 256), 256, 0, st>>>(d_states, g, direction, work, nullptr, d_out);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
