// K3 -- transition events -> jumps, plus the (start, destination) count matrix.
//
// Replaces _generic_transitions_to_jumps (reference src/gemdat/jumps.py:24-119: pandas
// groupby + DataFrame.iterrows, >= 25 us per event) and _calculate_transitions_matrix
// (src/gemdat/transitions.py:812-834: np.unique(axis=0) + fancy assignment).
//
// The conversion is a per-atom sequential state machine over that atom's events
// (already contiguous and time-ordered in the K2 table).  Atoms run in parallel, and so
// do segments of 128 events of the same atom (speculate / chain / fill, see below).
// Two phases (count, fill) give the exact output size without any allocation in the
// library.
#include <stdlib.h>

#include "gd_common.cuh"

namespace {

struct Jump {
    int32_t atom, start_site, dest_site, start_time, stop_time, residence;
};

// The state machine of jumps.py:45-105, on int32 rows
// (atom, start site, destination site, start inner, destination inner, time).
struct JumpMachine {
    bool have_from = false;
    int32_t from_start_site = 0, from_dest_site = 0, from_start_time = 0;
    bool have_cand = false;
    int32_t c_atom = 0, c_start_site = 0, c_dest_site = 0, c_dest_inner = 0, c_start_time = 0,
            c_stop_time = 0;
    int64_t minimal_residence;

    __device__ explicit JumpMachine(int64_t mr) : minimal_residence(mr) {}

    template <typename Emit>
    __device__ __forceinline__ void close_candidate(int64_t until, Emit &&emit) {
        const int64_t residence = until - ((int64_t)c_stop_time - 1);
        if (residence >= minimal_residence && c_dest_inner != GD_NOSITE &&
            c_start_site != c_dest_site) {  // jumps.py:113 start != destination
            emit(Jump{c_atom, c_start_site, c_dest_site, c_start_time, c_stop_time,
                      (int32_t)residence});
        }
        have_cand = false;
    }

    template <typename Emit>
    __device__ __forceinline__ void step(int32_t atom, int32_t s0, int32_t s1, int32_t i1,
                                         int32_t t, Emit &&emit) {
        if (have_cand) {
            if (s1 == c_dest_site) {
                if (i1 != GD_NOSITE) c_dest_inner = i1;
            } else {
                close_candidate(t, emit);
            }
        }
        if (s0 != GD_NOSITE && s0 != s1) {
            have_from = true;
            from_start_site = s0;
            from_dest_site = s1;
            from_start_time = t;
        }
        if (have_from) {
            if (s1 == from_start_site) {
                have_from = false;
                have_cand = false;
            } else if (i1 != GD_NOSITE || s1 != from_dest_site) {
                have_cand = true;
                c_atom = atom;
                c_start_site = from_start_site;
                c_dest_site = s1;
                c_dest_inner = i1;
                c_start_time = from_start_time;
                c_stop_time = t + 1;
                have_from = false;
            }
        }
    }
};

// ---------------------------------------------------------------- time-parallel conversion
//
// The machine is sequential per atom, but its state forgets the past quickly: after any
// "departure" event (start site != NOSITE and != destination site) the pending `from`
// is overwritten by that event and a pending candidate has been closed by it, so the state
// is a function of that event alone.  An atom's events are therefore cut into segments of
// kSeg rows that run in parallel:
//   1. k_jumps_spec   (thread per segment) runs the machine from the EMPTY state, records
//      the row after which its state can no longer depend on the true entry state (the
//      first departure), the jumps it emits after that row, and its exit state;
//   2. k_jumps_chain  (thread per atom) walks the atom's segments in order with the TRUE
//      state, replaying only each segment's short prefix (in lockstep with the speculative
//      run until both states are equal -- so odd inputs degrade to sequential, never to
//      wrong), and leaves the true entry state and jump count of every segment;
//   3. an exclusive scan of the counts, then k_jumps_fill (thread per segment) re-runs
//      each segment from its true entry state and writes its jumps in place.
// Output order = segment order = (atom, time), exactly the reference's.

#ifndef GD_K3_SEG
#define GD_K3_SEG 128
#endif
#ifndef GD_K3_SPEC_MIN_CTAS
#define GD_K3_SPEC_MIN_CTAS 1
#endif
constexpr int kSeg = GD_K3_SEG;      // event rows per segment

struct PackedState {  // 40 bytes == GD_JUMP_STATE_BYTES
    int32_t have;  // bit 0: from pending, bit 1: candidate pending
    int32_t f_start, f_dest, f_time;
    int32_t c_start, c_dest, c_inner, c_start_time, c_stop_time;
    int32_t flags;  // per-atom shard records only (kHasEvents | kSynced | kKnown), else 0
};
static_assert(sizeof(PackedState) == GD_JUMP_STATE_BYTES, "jump state record size is part of the ABI");

// flags of a per-atom record of a frame shard
constexpr int32_t kHasEvents = 1;  // the shard holds events of this atom
constexpr int32_t kSynced = 2;     // ... one of them a departure: the exit state does not depend on the entry
constexpr int32_t kKnown = 4;      // the state in the record is the true one
constexpr int32_t kUnproven = 8;   // (work only) the parallel chain could not prove one of the atom's segments

__device__ __forceinline__ PackedState pack(const JumpMachine &m) {
    PackedState s;
    s.have = (m.have_from ? 1 : 0) | (m.have_cand ? 2 : 0);
    s.f_start = m.from_start_site;
    s.f_dest = m.from_dest_site;
    s.f_time = m.from_start_time;
    s.c_start = m.c_start_site;
    s.c_dest = m.c_dest_site;
    s.c_inner = m.c_dest_inner;
    s.c_start_time = m.c_start_time;
    s.c_stop_time = m.c_stop_time;
    s.flags = 0;
    return s;
}

__device__ __forceinline__ void unpack(JumpMachine &m, const PackedState &s, int32_t atom) {
    m.have_from = (s.have & 1) != 0;
    m.have_cand = (s.have & 2) != 0;
    m.from_start_site = s.f_start;
    m.from_dest_site = s.f_dest;
    m.from_start_time = s.f_time;
    m.c_atom = atom;
    m.c_start_site = s.c_start;
    m.c_dest_site = s.c_dest;
    m.c_dest_inner = s.c_inner;
    m.c_start_time = s.c_start_time;
    m.c_stop_time = s.c_stop_time;
}

__device__ __forceinline__ bool same_state(const JumpMachine &x, const JumpMachine &y) {
    if (x.have_from != y.have_from || x.have_cand != y.have_cand) return false;
    if (x.have_from && (x.from_start_site != y.from_start_site || x.from_dest_site != y.from_dest_site ||
                        x.from_start_time != y.from_start_time))
        return false;
    if (x.have_cand && (x.c_start_site != y.c_start_site || x.c_dest_site != y.c_dest_site ||
                        x.c_dest_inner != y.c_dest_inner || x.c_start_time != y.c_start_time ||
                        x.c_stop_time != y.c_stop_time))
        return false;
    return true;
}

struct Row {
    int32_t atom, s0, s1, i0, i1, t;
};
__device__ __forceinline__ Row load_row(const int32_t *__restrict__ events, int64_t r) {
    const int2 *e = reinterpret_cast<const int2 *>(events + r * 6);
    const int2 a = __ldg(e), b = __ldg(e + 1), c = __ldg(e + 2);
    return Row{a.x, a.y, b.x, b.y, c.x, c.y};
}

// Walks rows [r0, r0 + n) in order, calling f(k, row) (return false to stop).  One thread owns
// the run, so every load instruction of a warp touches 32 different lines and the L1 tag stage
// (one line per cycle) is the limit: rows are therefore fetched with 16-byte loads, 8 rows =
// 12 loads ahead of the machine.  Rows are 24 B, so a row pair starting at an even row is
// 16-byte aligned when the table is (`wide`); an odd first row and the tail go one by one.
template <typename F>
__device__ __forceinline__ void walk_rows(const int32_t *__restrict__ events, int64_t r0, int n,
                                          bool wide, F &&f) {
    int k = 0;
    if (wide) {
        if ((r0 & 1) && n > 0) {
            if (!f(0, load_row(events, r0))) return;
            k = 1;
        }
        for (; k + 8 <= n; k += 8) {
            const int4 *q = reinterpret_cast<const int4 *>(events + (r0 + k) * 6);
            int4 v[12];
#pragma unroll
            for (int j = 0; j < 12; ++j) v[j] = __ldg(q + j);
#pragma unroll
            for (int j = 0; j < 4; ++j) {  // 3 int4 = 2 rows
                const int4 a = v[3 * j], b = v[3 * j + 1], c = v[3 * j + 2];
                if (!f(k + 2 * j, Row{a.x, a.y, a.z, a.w, b.x, b.y})) return;
                if (!f(k + 2 * j + 1, Row{b.z, b.w, c.x, c.y, c.z, c.w})) return;
            }
        }
    }
    for (; k < n; ++k)
        if (!f(k, load_row(events, r0 + k))) return;
}

struct SegSpec {  // 48 bytes: result of the speculative run of one segment
    PackedState exit;
    int32_t sync;     // rows [0, sync) must be replayed from the true entry state
    int32_t n_after;  // jumps the speculative run emits in rows [sync, n)
};

// segment bookkeeping: seg_off[a] = first segment of atom a (exclusive scan of
// ceil(n_events_a / kSeg)); single block
__global__ void __launch_bounds__(1024)
    k_seg_offsets(const int64_t *__restrict__ atom_offsets, int32_t n_atoms, int64_t *__restrict__ seg_off,
                  int64_t row_cap, int32_t *__restrict__ flag) {
    __shared__ int64_t warp_tot[32];
    __shared__ int64_t running_s;
    // the table may hold fewer rows than the offsets say (capacity guessed from an earlier run,
    // the count not read back yet): then there are NO segments, nothing downstream touches a
    // row, and flag[4] tells the caller to repeat with the right size
    const bool over = atom_offsets[n_atoms] > row_cap;
    if (threadIdx.x == 0) {
        running_s = 0;
        flag[4] = over ? 1 : 0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n_atoms; base += 1024) {
        const int idx = base + threadIdx.x;
        int64_t v = 0;  // every atom owns at least one (possibly empty) segment: its last segment
        if (idx < n_atoms && !over) {  // carries the atom's exit state and the end-of-trajectory close
            v = (atom_offsets[idx + 1] - atom_offsets[idx] + kSeg - 1) / kSeg;
            if (v < 1) v = 1;
        }
        int64_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        int64_t warp_base = 0;
        for (int w = 0; w < warp; ++w) warp_base += warp_tot[w];
        const int64_t running = running_s;
        __syncthreads();
        if (idx < n_atoms) seg_off[idx] = running + warp_base + incl - v;
        if (threadIdx.x == 1023) running_s = running + warp_base + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) seg_off[n_atoms] = running_s;
}

// (atom, first row, row count) of global segment `seg`
__device__ __forceinline__ bool locate_segment(int64_t seg, const int64_t *__restrict__ seg_off,
                                               const int64_t *__restrict__ atom_offsets, int32_t n_atoms,
                                               int32_t &atom, int64_t &r0, int &n) {
    if (seg >= seg_off[n_atoms]) return false;
    int lo = 0, hi = n_atoms;  // last atom with seg_off[atom] <= seg
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (seg_off[mid] <= seg) lo = mid; else hi = mid;
    }
    atom = lo;
    const int64_t j = seg - seg_off[lo];
    r0 = atom_offsets[lo] + j * kSeg;
    const int64_t left = atom_offsets[lo + 1] - r0;
    n = (int)(left < kSeg ? left : kSeg);
    return true;
}

__global__ void __launch_bounds__(128, GD_K3_SPEC_MIN_CTAS)
    k_jumps_spec(const int32_t *__restrict__ events, const int64_t *__restrict__ atom_offsets,
                 const int64_t *__restrict__ seg_off, int32_t n_atoms, int64_t minimal_residence,
                 SegSpec *__restrict__ spec, int32_t *__restrict__ atom_flags) {
    const int64_t seg = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int32_t atom;
    int64_t r0;
    int n;
    if (!locate_segment(seg, seg_off, atom_offsets, n_atoms, atom, r0, n)) return;
    JumpMachine m(minimal_residence);
    int sync = n, n_after = 0;
    bool synced = false;
    auto emit = [&](const Jump &) { n_after += synced ? 1 : 0; };
    walk_rows(events, r0, n, (reinterpret_cast<uintptr_t>(events) & 15) == 0, [&](int k, const Row &e) {
        m.step(e.atom, e.s0, e.s1, e.i1, e.t, emit);
        if (!synced && e.s0 != GD_NOSITE && e.s0 != e.s1) {
            synced = true;
            sync = k + 1;
        }
        return true;
    });
    SegSpec out;
    out.exit = pack(m);
    out.sync = sync;
    out.n_after = n_after;
    spec[seg] = out;
    const int32_t fl = (n > 0 ? kHasEvents : 0) | (synced ? kSynced : 0);
    if (fl) atomicOr(atom_flags + atom, fl);
}

// first-segment entry state of an atom: the caller's record (frame shards) or the empty state
__device__ __forceinline__ bool atom_entry_state(JumpMachine &m, const PackedState *__restrict__ atom_entry,
                                                 int32_t a) {
    m.c_atom = a;
    if (!atom_entry) return true;
    const PackedState e = atom_entry[a];
    unpack(m, e, a);
    return (e.flags & kKnown) != 0;
}

// record of the atom's state when it leaves this shard (before any end-of-trajectory close)
__device__ __forceinline__ void write_atom_exit(PackedState *__restrict__ atom_exit, int32_t a,
                                                const JumpMachine &m, int32_t shard_flags,
                                                bool entry_known) {
    if (!atom_exit) return;
    PackedState e = pack(m);
    e.flags = shard_flags | ((entry_known || (shard_flags & kSynced)) ? kKnown : 0);
    atom_exit[a] = e;
}

__global__ void __launch_bounds__(64)
    k_jumps_chain(const int32_t *__restrict__ events, const int64_t *__restrict__ atom_offsets,
                  const int64_t *__restrict__ seg_off, int32_t n_atoms, int64_t n_frames_total,
                  int64_t minimal_residence, const SegSpec *__restrict__ spec,
                  PackedState *__restrict__ entry, int64_t *__restrict__ counts,
                  const PackedState *__restrict__ atom_entry, PackedState *__restrict__ atom_exit,
                  const int32_t *__restrict__ atom_flags, int is_last_shard) {
    const int32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_atoms) return;
    const int64_t s0 = seg_off[a], s1 = seg_off[a + 1];
    const int64_t row_end = atom_offsets[a + 1];
    JumpMachine truth(minimal_residence);
    const bool entry_known = atom_entry_state(truth, atom_entry, a);
    for (int64_t seg = s0; seg < s1; ++seg) {
        entry[seg] = pack(truth);
        const int64_t r0 = atom_offsets[a] + (seg - s0) * kSeg;
        const int n = (int)min((int64_t)kSeg, row_end - r0);
        const SegSpec sp = spec[seg];
        JumpMachine guess(minimal_residence);
        int64_t n_true = 0;
        int n_guess_after = 0;  // jumps of the speculative run in rows [sync, k)
        int k = 0;
        bool merged = false;
        auto emit_true = [&](const Jump &) { ++n_true; };
        for (; k < n; ++k) {
            if (k >= sp.sync && same_state(truth, guess)) {
                merged = true;
                break;
            }
            const bool after = k >= sp.sync;
            auto emit_guess = [&](const Jump &) { n_guess_after += after ? 1 : 0; };
            const Row e = load_row(events, r0 + k);
            truth.step(e.atom, e.s0, e.s1, e.i1, e.t, emit_true);
            guess.step(e.atom, e.s0, e.s1, e.i1, e.t, emit_guess);
        }
        if (merged || same_state(truth, guess)) {
            // from here on the speculative run is the true run
            n_true += sp.n_after - n_guess_after;
            unpack(truth, sp.exit, a);
        }
        // (otherwise the whole segment was replayed and `truth` already is its exit state)
        if (seg + 1 == s1) {
            write_atom_exit(atom_exit, a, truth, atom_flags[a] & (kHasEvents | kSynced), entry_known);
            if (is_last_shard && truth.have_cand) truth.close_candidate(n_frames_total, emit_true);  // jumps.py:97-105
        }
        counts[seg] = n_true;
    }
}

// Parallel form of the chain (thread per segment).  After its first departure row a
// segment's state no longer depends on how it was entered, so the true entry state of segment
// s is the speculative exit state of segment s-1 -- PROVIDED the true run of s-1 merged with
// its speculative run.  Every thread therefore assumes that entry, replays its own prefix in
// lockstep with the speculative run and reports whether the two merged; by induction from the
// (exactly known, empty) entry of an atom's first segment all assumptions hold iff every
// segment merged.  If one does not (event streams that are not state-consistent), *flag is
// set and the caller falls back to the sequential k_jumps_chain: never wrong, rarely slow.
__global__ void __launch_bounds__(128)
    k_jumps_chain_par(const int32_t *__restrict__ events, const int64_t *__restrict__ atom_offsets,
                      const int64_t *__restrict__ seg_off, int32_t n_atoms, int64_t n_frames_total,
                      int64_t minimal_residence, const SegSpec *__restrict__ spec,
                      PackedState *__restrict__ entry, int64_t *__restrict__ counts,
                      int32_t *__restrict__ flag, const PackedState *__restrict__ atom_entry,
                      PackedState *__restrict__ atom_exit, int32_t *__restrict__ atom_flags,
                      int is_last_shard) {
    const int64_t seg = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int32_t a;
    int64_t r0;
    int n;
    if (!locate_segment(seg, seg_off, atom_offsets, n_atoms, a, r0, n)) return;
    const bool first = seg == seg_off[a], last = seg + 1 == seg_off[a + 1];
    JumpMachine truth(minimal_residence);
    bool entry_known = true;
    if (first) entry_known = atom_entry_state(truth, atom_entry, a);
    else unpack(truth, spec[seg - 1].exit, a);
    entry[seg] = pack(truth);
    const SegSpec sp = spec[seg];
    JumpMachine guess(minimal_residence);
    int64_t n_true = 0;
    int n_guess_after = 0;  // jumps of the speculative run in rows [sync, k)
    bool merged = false;
    auto emit_true = [&](const Jump &) { ++n_true; };
    for (int k = 0; k < n; ++k) {
        if (k >= sp.sync && same_state(truth, guess)) {
            merged = true;
            break;
        }
        const bool after = k >= sp.sync;
        auto emit_guess = [&](const Jump &) { n_guess_after += after ? 1 : 0; };
        const Row e = load_row(events, r0 + k);
        truth.step(e.atom, e.s0, e.s1, e.i1, e.t, emit_true);
        guess.step(e.atom, e.s0, e.s1, e.i1, e.t, emit_guess);
    }
    if (merged || same_state(truth, guess)) {
        n_true += sp.n_after - n_guess_after;
        unpack(truth, sp.exit, a);
    } else if (!last) {
        atomicExch(flag, 1);  // the next segment's assumed entry state is not proven
        atomicOr(atom_flags + a, kUnproven);
    }
    if (last) {
        // entry_known of a later segment: proven by the merge checks of the segments before it
        // (else *flag is set and the sequential chain redoes this); the record's kKnown then
        // follows from the atom's own entry record
        const bool known = first ? entry_known : (!atom_entry || (atom_entry[a].flags & kKnown) != 0);
        write_atom_exit(atom_exit, a, truth, atom_flags[a] & (kHasEvents | kSynced), known);
        if (is_last_shard && truth.have_cand) truth.close_candidate(n_frames_total, emit_true);  // jumps.py:97-105
    }
    counts[seg] = n_true;
}

__global__ void __launch_bounds__(128)
    k_jumps_fill(const int32_t *__restrict__ events, const int64_t *__restrict__ atom_offsets,
                 const int64_t *__restrict__ seg_off, int32_t n_atoms, int64_t n_frames_total,
                 int64_t minimal_residence, const PackedState *__restrict__ entry,
                 const int64_t *__restrict__ counts, int32_t *__restrict__ jumps,
                 const int32_t *__restrict__ shard_flag, int64_t jump_capacity) {
    const int64_t seg = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int32_t atom;
    int64_t r0;
    int n;
    if (!locate_segment(seg, seg_off, atom_offsets, n_atoms, atom, r0, n)) return;
    const int64_t out0 = counts[seg], out1 = counts[seg + 1];  // exclusive offsets after the scan
    if (out1 == out0 || out1 > jump_capacity) return;  // (a too small table is never overrun)
    JumpMachine m(minimal_residence);
    // `atom` is the index inside this table (an atom block after the event exchange); jump rows
    // carry the atom index of the event rows, also for a candidate inherited from the entry state
    unpack(m, entry[seg], n > 0 ? __ldg(events + r0 * 6) : atom);
    int64_t out = out0;
    auto emit = [&](const Jump &j) {
        int2 *o = reinterpret_cast<int2 *>(jumps + out * 6);
        o[0] = make_int2(j.atom, j.start_site);
        o[1] = make_int2(j.dest_site, j.start_time);
        o[2] = make_int2(j.stop_time, j.residence);
        ++out;
    };
    walk_rows(events, r0, n, (reinterpret_cast<uintptr_t>(events) & 15) == 0, [&](int, const Row &e) {
        m.step(e.atom, e.s0, e.s1, e.i1, e.t, emit);
        return out < out1;  // all of this segment's jumps are out: the rest of its rows is moot
    });
    const bool last_of_atom = r0 + n == atom_offsets[atom + 1];
    // shard_flag[2]: this is the trajectory's last frame shard (left by the count call)
    if (last_of_atom && shard_flag[2] != 0 && m.have_cand && out < out1) m.close_candidate(n_frames_total, emit);
}

// exclusive scan, c[0..n) -> offsets, c[n] = total.  Two launches: every block scans its tile of
// kScanTile items in place and leaves the tile total; then every block adds the totals of the tiles
// before it (a few hundred values at most, summed redundantly per block).
constexpr int kScanItems = 4;
constexpr int kScanTile = 1024 * kScanItems;

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t *warp_tot, int64_t &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int64_t wv = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;
    int64_t wincl = wv;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, wincl, o);
        if (lane >= o) wincl += t;
    }
    total = __shfl_sync(0xffffffffu, wincl, 31);
    const int64_t warp_base = __shfl_sync(0xffffffffu, wincl - wv, warp);
    __syncthreads();
    return warp_base + incl - v;
}

__global__ void __launch_bounds__(1024)
    k_scan_tiles(int64_t *__restrict__ c, int64_t n, int64_t *__restrict__ tile_tot) {
    __shared__ int64_t warp_tot[32];
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int64_t v[kScanItems], sum = 0;
#pragma unroll
    for (int j = 0; j < kScanItems; ++j) {
        v[j] = base + j < n ? c[base + j] : 0;
        sum += v[j];
    }
    int64_t total;
    int64_t excl = block_exclusive_scan(sum, warp_tot, total);
#pragma unroll
    for (int j = 0; j < kScanItems; ++j) {
        if (base + j < n) c[base + j] = excl;
        excl += v[j];
    }
    if (threadIdx.x == 0) tile_tot[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024)
    k_scan_add(int64_t *__restrict__ c, int64_t n, const int64_t *__restrict__ tile_tot) {
    __shared__ int64_t warp_tot[32];
    int64_t part = 0;  // totals of the tiles before this one
    for (int64_t b = threadIdx.x; b < blockIdx.x; b += 1024) part += tile_tot[b];
    int64_t before;
    block_exclusive_scan(part, warp_tot, before);
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
#pragma unroll
    for (int j = 0; j < kScanItems; ++j)
        if (base + j < n) c[base + j] += before;
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) c[n] = before + tile_tot[blockIdx.x];
}

// scratch layout inside the caller's d_work (all 16-byte aligned)
struct JumpWork {
    int64_t max_segs;
    int64_t *seg_off;    // [n_atoms + 1]
    int64_t *counts;     // [max_segs + 1]
    SegSpec *spec;       // [max_segs]
    PackedState *entry;  // [max_segs]
    int32_t *flag;       // [8]: [0] parallel chain could not prove a segment's entry state,
                         //      [1] a published shard record differs from the true exit, [2] is_last_shard,
                         //      [3] unresolved entries (gd_jumps_resolve), [4] more rows than the table holds
    int32_t *atom_flags; // [n_atoms]: kHasEvents | kSynced of this shard's events per atom
    int64_t *tile_tot;   // [max_segs / kScanTile + 1] scan tile totals
    size_t bytes;
};
inline JumpWork jump_work(void *base, int64_t n_events, int32_t n_atoms) {
    auto up = [](size_t x) { return (x + 15) & ~size_t(15); };
    JumpWork w;
    w.max_segs = n_events / kSeg + n_atoms + 1;
    unsigned char *p = static_cast<unsigned char *>(base);
    size_t off = 0;
    w.seg_off = reinterpret_cast<int64_t *>(p + off);
    off += up((size_t)(n_atoms + 1) * 8);
    w.counts = reinterpret_cast<int64_t *>(p + off);
    off += up((size_t)(w.max_segs + 1) * 8);
    w.spec = reinterpret_cast<SegSpec *>(p + off);
    off += up((size_t)w.max_segs * sizeof(SegSpec));
    w.entry = reinterpret_cast<PackedState *>(p + off);
    off += up((size_t)w.max_segs * sizeof(PackedState));
    w.flag = reinterpret_cast<int32_t *>(p + off);
    off += 32;
    w.atom_flags = reinterpret_cast<int32_t *>(p + off);
    off += up((size_t)n_atoms * 4);
    w.tile_tot = reinterpret_cast<int64_t *>(p + off);
    off += up((size_t)(w.max_segs / kScanTile + 1) * 8);
    w.bytes = off;
    return w;
}

// ---------------------------------------------------------------- pair matrix

__global__ void __launch_bounds__(256)
    k_pair_count(const int32_t *__restrict__ rows, int64_t n_rows, int32_t stride, int32_t cs,
                 int32_t cd, int32_t n_sites, unsigned long long *__restrict__ scratch) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const int64_t w = (int64_t)n_sites + 1;
    for (; i < n_rows; i += step) {
        const int32_t s = rows[i * stride + cs], d = rows[i * stride + cd];
        if (s < -1 || s >= n_sites || d < -1 || d >= n_sites) continue;
        atomicAdd(scratch + (int64_t)(s + 1) * w + (d + 1), 1ULL);
    }
}

// numpy semantics of `m[start_idx, stop_idx] = counts` with unique rows sorted
// ascending: index -1 aliases n-1 and the lexicographically last source pair wins.
__global__ void __launch_bounds__(256)
    k_pair_finalize(const unsigned long long *__restrict__ scratch, int32_t n_sites,
                    long long *__restrict__ matrix) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_sites;
    if (idx >= n * n) return;
    const int64_t i = idx / n, j = idx - i * n;
    const int64_t w = n + 1;
    unsigned long long v = scratch[(i + 1) * w + (j + 1)];
    if (v == 0 && j == n - 1) v = scratch[(i + 1) * w + 0];
    if (v == 0 && i == n - 1) v = scratch[0 * w + (j + 1)];
    if (v == 0 && i == n - 1 && j == n - 1) v = scratch[0];
    matrix[idx] = (long long)v;
}

}  // namespace

extern "C" int64_t gd_jumps_work_bytes(int64_t n_events, int32_t n_atoms) {
    if (n_events < 0 || n_atoms <= 0) return 0;
    return (int64_t)jump_work(nullptr, n_events, n_atoms).bytes;
}

namespace {

// After a parallel chain whose proof is not read back (first pass of a frame shard): the exit
// record of an atom with an unproven segment says "not known", so the resolve step asks for
// another (checked) round instead of trusting it.  Clears the work bit for the next pass.
__global__ void __launch_bounds__(256)
    k_exit_finish(PackedState *__restrict__ atom_exit, int32_t *__restrict__ atom_flags, int32_t n_atoms) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_atoms) return;
    const int32_t fl = atom_flags[a];
    if (fl & kUnproven) {
        atom_flags[a] = fl & ~kUnproven;
        if (atom_exit) atom_exit[a].flags &= ~(kKnown | kSynced);
    }
}

// per-atom exclusive jump offsets from the scanned per-segment counts
__global__ void __launch_bounds__(256)
    k_atom_jump_offsets(const int64_t *__restrict__ counts, const int64_t *__restrict__ seg_off,
                        int32_t n_atoms, int64_t *__restrict__ out, const int32_t *__restrict__ flag,
                        const int64_t *__restrict__ atom_offsets, int write_tail) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a <= n_atoms) out[a] = counts[seg_off[a]];
    if (a == 0 && write_tail)  // status word of a pass whose flags are not read back by the host
        out[n_atoms + 1] = (atom_offsets[n_atoms] << 8) | (flag[1] ? GD_JUMPS_TAIL_MISMATCH : 0) |
                           (flag[0] ? GD_JUMPS_TAIL_UNPROVEN : 0) | (flag[3] ? GD_JUMPS_TAIL_UNRESOLVED : 0) |
                           (flag[4] ? GD_JUMPS_TAIL_ROWS_OVERFLOW : 0);
}

// a record published as known (entry independent, or run from a known entry) must equal the
// true exit state of the final pass
__global__ void __launch_bounds__(256)
    k_check_published(const PackedState *__restrict__ exit_now, const PackedState *__restrict__ published,
                      int32_t n_atoms, int32_t *__restrict__ flag) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_atoms) return;
    const PackedState pub = published[a];
    if (!(pub.flags & kKnown)) return;
    JumpMachine x(0), y(0);
    unpack(x, exit_now[a], a);
    unpack(y, pub, a);
    if (!same_state(x, y)) atomicExch(flag + 1, 1);
}

// Entry state of every atom for shard `rank` from the records of all shards (thread per
// (shard, atom): every rank evaluates all shards so that all ranks agree on `unresolved`).
__global__ void __launch_bounds__(256)
    k_jumps_resolve(const PackedState *__restrict__ all, int32_t world, int32_t rank, int32_t n_atoms,
                    PackedState *__restrict__ entry, int32_t *__restrict__ unresolved) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)world * n_atoms) return;
    const int32_t q = (int32_t)(idx / n_atoms), a = (int32_t)(idx - (int64_t)q * n_atoms);
    PackedState e;
    e.have = 0;
    e.f_start = e.f_dest = e.f_time = 0;
    e.c_start = e.c_dest = e.c_inner = e.c_start_time = e.c_stop_time = 0;
    e.flags = kKnown;  // no earlier shard has events of this atom: the empty state
    for (int32_t p = q - 1; p >= 0; --p) {
        const PackedState r = all[(int64_t)p * n_atoms + a];
        if (!(r.flags & kHasEvents)) continue;  // the atom's state passes through shard p unchanged
        e = r;
        e.flags = r.flags & kKnown;
        break;
    }
    if (!(e.flags & kKnown)) atomicAdd(unresolved, 1);
    if (q == rank) entry[a] = e;
}

}  // namespace

extern "C" gd_status_t gd_jumps_count_shard(const int32_t *d_events, const int64_t *d_atom_offsets,
                                            int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                            int64_t minimal_residence, const void *d_entry,
                                            int is_last_shard, int reuse_work, const void *d_published,
                                            void *d_work,
                                            void *d_exit, int64_t *d_atom_jump_offsets,
                                            int64_t *h_n_jumps, int32_t *h_flags, void *stream) {
    GD_REQUIRE(d_atom_offsets && d_work, GD_ERR_INVALID, "gd_jumps_count: null argument");
    GD_REQUIRE(n_events >= 0 && (n_events == 0 || d_events), GD_ERR_INVALID,
               "gd_jumps_count: bad event table");
    GD_REQUIRE(n_atoms > 0, GD_ERR_INVALID, "gd_jumps_count: n_atoms must be positive");
    GD_REQUIRE((reinterpret_cast<uintptr_t>(d_work) & 15) == 0, GD_ERR_ALIGNMENT,
               "gd_jumps_count: d_work must be 16-byte aligned");
    GD_REQUIRE(!d_published || d_exit, GD_ERR_INVALID, "gd_jumps_count_shard: d_published needs d_exit");
    cudaStream_t st = (cudaStream_t)stream;
    const JumpWork w = jump_work(d_work, n_events, n_atoms);
    const PackedState *entry = static_cast<const PackedState *>(d_entry);
    PackedState *exit_rec = static_cast<PackedState *>(d_exit);
    if (!reuse_work) {  // segment bookkeeping + speculation: independent of the entry records
        GD_CUDA_CHECK(cudaMemsetAsync(w.flag, 0, 32, st));  // incl. flag[3]: no unresolved entries yet
        k_seg_offsets<<<1, 1024, 0, st>>>(d_atom_offsets, n_atoms, w.seg_off, n_events, w.flag);
        GD_LAUNCH_CHECK();
        GD_CUDA_CHECK(cudaMemsetAsync(w.atom_flags, 0, (size_t)n_atoms * 4, st));
        k_jumps_spec<<<(unsigned)gd_ceil_div(w.max_segs, 128), 128, 0, st>>>(
            d_events, d_atom_offsets, w.seg_off, n_atoms, minimal_residence, w.spec, w.atom_flags);
        GD_LAUNCH_CHECK();
    }
    GD_CUDA_CHECK(cudaMemsetAsync(w.counts, 0, (size_t)(w.max_segs + 1) * 8, st));
    // flag[0..2]; flag[3] (unresolved entries, left by gd_jumps_resolve) is kept
    static const int32_t flags_last[3] = {0, 0, 1}, flags_mid[3] = {0, 0, 0};
    GD_CUDA_CHECK(cudaMemcpyAsync(w.flag, is_last_shard ? flags_last : flags_mid, 12,
                                  cudaMemcpyHostToDevice, st));
    // GD_JUMPS_SEQUENTIAL=1 forces the sequential chain (tests / A-B timing)
    const char *env = getenv("GD_JUMPS_SEQUENTIAL");
    bool sequential = env && atoi(env) != 0;
    int32_t h_flag[5] = {0, 0, 0, 0, 0};
    if (!h_n_jumps && !h_flags && !sequential && !d_atom_jump_offsets) {
        // asynchronous pass: only the exit records are wanted.  The chain's proof is not read
        // back; atoms with an unproven segment publish "not known" instead (k_exit_finish).
        GD_REQUIRE(exit_rec && !d_published, GD_ERR_INVALID,
                   "gd_jumps_count_shard: an asynchronous pass only produces d_exit");
        k_jumps_chain_par<<<(unsigned)gd_ceil_div(w.max_segs, 128), 128, 0, st>>>(
            d_events, d_atom_offsets, w.seg_off, n_atoms, n_frames_total, minimal_residence,
            w.spec, w.entry, w.counts, w.flag, entry, exit_rec, w.atom_flags, is_last_shard);
        GD_LAUNCH_CHECK();
        k_exit_finish<<<(unsigned)gd_ceil_div(n_atoms, 256), 256, 0, st>>>(exit_rec, w.atom_flags, n_atoms);
        GD_LAUNCH_CHECK();
        return GD_OK;
    }
    if (!h_n_jumps && !h_flags && !sequential) {
        // asynchronous FINAL pass: counts, per-atom offsets and a status word (d_atom_jump_offsets
        // [n_atoms + 1]) are left on the device; the caller reads them when it reads the jump
        // offsets anyway and repeats with a checked pass if a status bit is set.
        k_jumps_chain_par<<<(unsigned)gd_ceil_div(w.max_segs, 128), 128, 0, st>>>(
            d_events, d_atom_offsets, w.seg_off, n_atoms, n_frames_total, minimal_residence,
            w.spec, w.entry, w.counts, w.flag, entry, exit_rec, w.atom_flags, is_last_shard);
        GD_LAUNCH_CHECK();
        const unsigned n_tiles = (unsigned)gd_ceil_div(w.max_segs, kScanTile);
        k_scan_tiles<<<n_tiles, 1024, 0, st>>>(w.counts, w.max_segs, w.tile_tot);
        GD_LAUNCH_CHECK();
        k_scan_add<<<n_tiles, 1024, 0, st>>>(w.counts, w.max_segs, w.tile_tot);
        GD_LAUNCH_CHECK();
        if (d_published) {
            k_check_published<<<(unsigned)gd_ceil_div(n_atoms, 256), 256, 0, st>>>(
                exit_rec, static_cast<const PackedState *>(d_published), n_atoms, w.flag);
            GD_LAUNCH_CHECK();
        }
        k_atom_jump_offsets<<<(unsigned)gd_ceil_div(n_atoms + 1, 256), 256, 0, st>>>(
            w.counts, w.seg_off, n_atoms, d_atom_jump_offsets, w.flag, d_atom_offsets, 1);
        GD_LAUNCH_CHECK();
        k_exit_finish<<<(unsigned)gd_ceil_div(n_atoms, 256), 256, 0, st>>>(nullptr, w.atom_flags, n_atoms);
        GD_LAUNCH_CHECK();
        return GD_OK;
    }
    for (;;) {
        if (sequential) {
            k_jumps_chain<<<(unsigned)gd_ceil_div(n_atoms, 64), 64, 0, st>>>(
                d_events, d_atom_offsets, w.seg_off, n_atoms, n_frames_total, minimal_residence,
                w.spec, w.entry, w.counts, entry, exit_rec, w.atom_flags, is_last_shard);
        } else {
            k_jumps_chain_par<<<(unsigned)gd_ceil_div(w.max_segs, 128), 128, 0, st>>>(
                d_events, d_atom_offsets, w.seg_off, n_atoms, n_frames_total, minimal_residence,
                w.spec, w.entry, w.counts, w.flag, entry, exit_rec, w.atom_flags, is_last_shard);
        }
        GD_LAUNCH_CHECK();
        const unsigned n_tiles = (unsigned)gd_ceil_div(w.max_segs, kScanTile);
        k_scan_tiles<<<n_tiles, 1024, 0, st>>>(w.counts, w.max_segs, w.tile_tot);
        GD_LAUNCH_CHECK();
        k_scan_add<<<n_tiles, 1024, 0, st>>>(w.counts, w.max_segs, w.tile_tot);
        GD_LAUNCH_CHECK();
        if (d_atom_jump_offsets) {
            k_atom_jump_offsets<<<(unsigned)gd_ceil_div(n_atoms + 1, 256), 256, 0, st>>>(
                w.counts, w.seg_off, n_atoms, d_atom_jump_offsets, w.flag, d_atom_offsets, 0);
            GD_LAUNCH_CHECK();
        }
        if (d_published) {
            k_check_published<<<(unsigned)gd_ceil_div(n_atoms, 256), 256, 0, st>>>(
                exit_rec, static_cast<const PackedState *>(d_published), n_atoms, w.flag);
            GD_LAUNCH_CHECK();
        }
        if (h_n_jumps)
            GD_CUDA_CHECK(cudaMemcpyAsync(h_n_jumps, w.counts + w.max_segs, sizeof(int64_t),
                                          cudaMemcpyDeviceToHost, st));
        GD_CUDA_CHECK(cudaMemcpyAsync(h_flag, w.flag, 5 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        GD_CUDA_CHECK(cudaStreamSynchronize(st));
        if (sequential || h_flag[0] == 0) break;
        // some segment's entry state could not be proven: redo with the sequential chain
        sequential = true;
        GD_CUDA_CHECK(cudaMemsetAsync(w.counts, 0, (size_t)(w.max_segs + 1) * 8, st));
        GD_CUDA_CHECK(cudaMemsetAsync(w.flag, 0, 8, st));
    }
    // the parallel chain may have left kUnproven bits behind; the next pass must not see them
    if (h_flag[0]) {
        k_exit_finish<<<(unsigned)gd_ceil_div(n_atoms, 256), 256, 0, st>>>(nullptr, w.atom_flags, n_atoms);
        GD_LAUNCH_CHECK();
    }
    if (h_flags) *h_flags = (sequential ? 1 : 0) | (h_flag[1] ? 2 : 0) | (h_flag[3] ? 4 : 0) | (h_flag[4] ? 8 : 0);
    return GD_OK;
}

extern "C" gd_status_t gd_jumps_count(const int32_t *d_events, const int64_t *d_atom_offsets,
                                      int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                      int64_t minimal_residence, void *d_work, int64_t *h_n_jumps,
                                      void *stream) {
    GD_REQUIRE(h_n_jumps, GD_ERR_INVALID, "gd_jumps_count: null argument");
    return gd_jumps_count_shard(d_events, d_atom_offsets, n_events, n_atoms, n_frames_total,
                                minimal_residence, nullptr, 1, 0, nullptr, d_work, nullptr, nullptr,
                                h_n_jumps, nullptr, stream);
}

extern "C" gd_status_t gd_jumps_resolve(const void *d_all_records, int32_t world, int32_t rank,
                                        int32_t n_atoms, void *d_entry, int32_t *d_unresolved,
                                        int32_t *h_unresolved, void *stream) {
    GD_REQUIRE(d_all_records && d_entry && d_unresolved, GD_ERR_INVALID,
               "gd_jumps_resolve: null argument");
    GD_REQUIRE(world > 0 && rank >= 0 && rank < world && n_atoms > 0, GD_ERR_INVALID,
               "gd_jumps_resolve: bad shard geometry");
    cudaStream_t st = (cudaStream_t)stream;
    GD_CUDA_CHECK(cudaMemsetAsync(d_unresolved, 0, sizeof(int32_t), st));
    k_jumps_resolve<<<(unsigned)gd_ceil_div((int64_t)world * n_atoms, 256), 256, 0, st>>>(
        static_cast<const PackedState *>(d_all_records), world, rank, n_atoms,
        static_cast<PackedState *>(d_entry), d_unresolved);
    GD_LAUNCH_CHECK();
    if (h_unresolved) {
        GD_CUDA_CHECK(cudaMemcpyAsync(h_unresolved, d_unresolved, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        GD_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    return GD_OK;
}

extern "C" int64_t gd_jumps_work_unresolved_offset(int64_t n_events, int32_t n_atoms) {
    if (n_events < 0 || n_atoms <= 0) return -1;
    const JumpWork w = jump_work(nullptr, n_events, n_atoms);
    return (int64_t)(reinterpret_cast<const unsigned char *>(w.flag + 3) - static_cast<const unsigned char *>(nullptr));
}

extern "C" gd_status_t gd_jumps_fill_capped(const int32_t *d_events, const int64_t *d_atom_offsets,
                                            int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                            int64_t minimal_residence, const void *d_work,
                                            int32_t *d_jumps, int64_t jump_capacity, void *stream) {
    GD_REQUIRE(d_atom_offsets && d_work && d_jumps, GD_ERR_INVALID, "gd_jumps_fill: null argument");
    GD_REQUIRE(n_atoms > 0 && n_events >= 0 && (n_events == 0 || d_events) && jump_capacity >= 0,
               GD_ERR_INVALID, "gd_jumps_fill: bad size");
    const JumpWork w = jump_work(const_cast<void *>(d_work), n_events, n_atoms);
    k_jumps_fill<<<(unsigned)gd_ceil_div(w.max_segs, 128), 128, 0, (cudaStream_t)stream>>>(
        d_events, d_atom_offsets, w.seg_off, n_atoms, n_frames_total, minimal_residence, w.entry,
        w.counts, d_jumps, w.flag, jump_capacity);
    GD_LAUNCH_CHECK();
    return GD_OK;
}

extern "C" gd_status_t gd_jumps_fill(const int32_t *d_events, const int64_t *d_atom_offsets,
                                     int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                     int64_t minimal_residence, const void *d_work,
                                     int32_t *d_jumps, void *stream) {
    return gd_jumps_fill_capped(d_events, d_atom_offsets, n_events, n_atoms, n_frames_total,
                                minimal_residence, d_work, d_jumps, INT64_MAX, stream);
}

extern "C" gd_status_t gd_pair_matrix(const int32_t *d_rows, int64_t n_rows, int32_t row_stride,
                                      int32_t col_start, int32_t col_dest, int32_t n_sites,
                                      unsigned long long *d_scratch, long long *d_matrix,
                                      void *stream) {
    GD_REQUIRE(d_scratch && d_matrix && n_sites > 0 && n_rows >= 0, GD_ERR_INVALID,
               "gd_pair_matrix: bad argument");
    GD_REQUIRE(n_rows == 0 || d_rows, GD_ERR_INVALID, "gd_pair_matrix: null rows");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_rows > 0) {
        int64_t blocks = gd_ceil_div(n_rows, 256);
        if (blocks > 4096) blocks = 4096;
        k_pair_count<<<(unsigned)blocks, 256, 0, st>>>(d_rows, n_rows, row_stride, col_start,
                                                      col_dest, n_sites, d_scratch);
        GD_LAUNCH_CHECK();
    }
    const int64_t cells = (int64_t)n_sites * n_sites;
    k_pair_finalize<<<(unsigned)gd_ceil_div(cells, 256), 256, 0, st>>>(d_scratch, n_sites, d_matrix);
    GD_LAUNCH_CHECK();
    return GD_OK;
}
