/*
 * gemdat_b200 -- C ABI of the B200-native GEMDAT hot path (libgemdat_b200.so).
 *
 * GEMDAT (the reference, pure Python) has no FFI of its own: the seams these
 * entry points sit behind are private Python functions and the third-party
 * native code they reach.  Each entry point below names the reference
 * interface it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *  - every pointer named d_* is DEVICE memory owned by the caller; h_* is host
 *    memory.  The library never allocates device memory and never frees
 *    anything; variable-length tables are two-phase (count -> caller
 *    allocates -> fill).
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *    Calls are asynchronous on that stream unless stated otherwise.
 *  - return value: GD_OK (0) or a GD_ERR_* code; gd_last_error() gives a
 *    thread-local human readable message for the last failing call.
 *  - NOSITE is -1 everywhere (src/gemdat/transitions.py:27).
 *  - no CPU fallback exists: without a CUDA device every compute entry point
 *    returns GD_ERR_CUDA.
 */
#ifndef GEMDAT_B200_H
#define GEMDAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int gd_status_t;

enum {
    GD_OK = 0,
    GD_ERR_CUDA = 1,           /* CUDA runtime error (message has the detail)            */
    GD_ERR_INVALID = 2,        /* bad argument (null pointer, negative size, ...)         */
    GD_ERR_ALIGNMENT = 3,      /* a device pointer is not aligned as documented           */
    GD_ERR_TRICLINIC = 4,      /* lattice angles are not all 90 deg in float32  */
    GD_ERR_CAPACITY = 5,       /* caller-provided buffer too small; needed size returned  */
    GD_ERR_RADIUS_GT_CUTOFF = 6, /* radius*fraction > max radius: MDAnalysis RuntimeError   */
    GD_ERR_INNER_CONSTANT = 7, /* transitions.py:451 IndexError (outer changes, inner not) */
    GD_ERR_NO_EVENTS = 8,      /* transitions.py:470 ValueError (np.vstack of nothing)    */
    GD_ERR_NO_JUMPS = 9,       /* jumps.py:107-108 ValueError('No jumps found')           */
    GD_ERR_RANGE = 10          /* volume.py:476-477 AssertionError (coords not in [0,1))  */
};

const char *gd_last_error(void);
int gd_version(void);
/* number of SMs of the current device (148 on B200); <0 on error */
int gd_sm_count(void);

/* ------------------------------------------------------------------------- *
 * K1  site-state assignment
 * replaces: _calculate_atom_states, src/gemdat/transitions.py:736-809, i.e.
 *   MDAnalysis PeriodicKDTree.set_coords/search_tree (float32 wrap + periodic
 *   images + scipy cKDTree.query_ball_tree) and the last-write-wins scatter
 *   (transitions.py:802-807), for the outer (fraction 1.0, :160-165) and the
 *   inner (:167-172) call in ONE pass.
 * ------------------------------------------------------------------------- */

/* Build the site search structure ("site grid") on the host.
 *
 *  h_lattice      3x3 row-major lattice matrix (Angstrom); must be orthorhombic
 *                 in the sense MDAnalysis uses (all angles == 90 as float32).
 *  h_sites_frac   [n_sites,3] fractional site coordinates (Structure.frac_coords)
 *  h_site_group   [n_sites] index of the site_radius dict entry the site belongs
 *                 to, in dict order (later groups overwrite earlier ones,
 *                 transitions.py:785-807); -1 = site carries no listed label.
 *  h_group_radius [n_groups] site_radius[label] (Angstrom)
 *  h_group_inner  [n_groups] site_inner_fraction[label]
 *  h_blob/cap     output buffer (host) and its capacity in bytes.  On
 *                 GD_ERR_CAPACITY *needed holds the required size; call again.
 *
 * The blob is position independent: copy it to the device verbatim
 * (256-byte aligned) and hand that pointer to gd_site_states.
 */
gd_status_t gd_sitegrid_build(const double *h_lattice, const double *h_sites_frac,
                              const int32_t *h_site_group, int32_t n_sites,
                              const double *h_group_radius, const double *h_group_inner,
                              int32_t n_groups, void *h_blob, size_t cap, size_t *needed);

/* One pass over all atom-frames of a frame shard, fused with the event COUNT of K2
 * (gd_events_count) so that the states are never re-read for counting.
 *
 *  d_pos_frac     [n_frames,n_atoms,3] float64 fractional positions, C order
 *                 (Trajectory.positions); 16-byte aligned.
 *  d_next_pos_row [n_atoms,3] positions of the first frame of the FOLLOWING frame
 *                 shard (multi-GPU halo; its state is recomputed here), or NULL when
 *                 this shard ends the trajectory; 16-byte aligned.
 *  d_blob/h_blob  device / host copy of the gd_sitegrid_build blob
 *  d_states_outer [n_frames,n_atoms] int32 site index or -1  (Transitions.states)
 *  d_states_inner [n_frames,n_atoms] int32                   (Transitions.inner_states)
 *  d_visited      [2*n_sites] uint8, zero-initialised by the caller: set to 1 for
 *                 every site that contains at least one atom-frame (outer: first
 *                 n_sites bytes, inner: next n_sites).  Feeds the reference's
 *                 "No floating species in range" warning (transitions.py:798-800)
 *                 and the integer_remap quirk (gd_remap_states).  May be NULL.
 *  d_band_counts  [4] uint64, zero-initialised: atom-frames with | |d|-r | < 1e-12 A
 *                 (outer, inner) and < 4e-6 A (outer, inner) for some site; the
 *                 float32 staging of the reference makes the latter band the one
 *                 where a pure-fp64 method would disagree.  May be NULL.
 *  d_event_work   gd_events_work_elems(n_frames,n_atoms) int32: receives what
 *                 gd_events_count would write (per (atom, frame chunk) event counts);
 *  d_changed      [2*n_atoms] uint8, zero-initialised: as in gd_events_count.
 *                 Both NULL = states only.  NOTE: counts are taken on the states as
 *                 written here; if gd_remap_states changes them afterwards (only
 *                 when a labelled group has never-visited sites) the event pattern is
 *                 unchanged because the remap is injective.
 */
gd_status_t gd_site_states(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                           const double *d_next_pos_row, const void *d_blob, const void *h_blob,
                           int32_t *d_states_outer, int32_t *d_states_inner, uint8_t *d_visited,
                           unsigned long long *d_band_counts, int32_t *d_event_work,
                           uint8_t *d_changed, void *stream);

/* gd_site_states that also writes the event ROWS as it finds them, so that K2 never re-reads
 * the 8 B/atom-frame of states (gd_events_compact instead of gd_events_fill).
 *  d_stage     gd_events_stage_elems(n_frames, n_atoms, stage_cap) int32, 8-byte aligned,
 *              uninitialised: [frame chunk, atom, stage_cap] rows of 6 int32 in the reference's
 *              column order (transitions.py:455-468), time relative to this shard's first frame;
 *              only the first (event count) rows of a segment are written.
 *  stage_cap   rows per (frame chunk, atom) segment; gd_events_stage_cap gives the default
 *              (a quarter of the chunk's frames).
 *  d_overflow  one int32, zero-initialised: set when some segment holds more events than
 *              stage_cap (then use gd_events_fill, which is always valid).
 * The staged rows carry the states as written here: when gd_remap_states changes them
 * afterwards, use gd_events_fill. */
int32_t gd_events_stage_cap(int64_t n_frames, int32_t n_atoms);
int64_t gd_events_stage_elems(int64_t n_frames, int32_t n_atoms, int32_t stage_cap);
gd_status_t gd_site_states_staged(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                                  const double *d_next_pos_row, const void *d_blob,
                                  const void *h_blob, int32_t *d_states_outer,
                                  int32_t *d_states_inner, uint8_t *d_visited,
                                  unsigned long long *d_band_counts, int32_t *d_event_work,
                                  uint8_t *d_changed, int32_t *d_stage, int32_t stage_cap,
                                  int32_t *d_overflow, void *stream);

/* integer_remap quirk (src/gemdat/utils.py:101-125 via transitions.py:804-805):
 * d_lut [n_sites] int32 maps the natural site index to the index the reference
 * reports; applied in place.  Only needed when a labelled group has a site that
 * was never visited. */
gd_status_t gd_remap_states(int32_t *d_states, int64_t n, const int32_t *d_lut, int32_t n_sites,
                            void *stream);

/* int32 -> int64 widening for the host-facing arrays (the reference returns int64). */
gd_status_t gd_widen_i32_i64(const int32_t *d_in, int64_t *d_out, int64_t n, void *stream);

/* ------------------------------------------------------------------------- *
 * K2  transition events, forward / backward fill
 * replaces: _calculate_transition_events, src/gemdat/transitions.py:413-482;
 *           ffill / bfill, src/gemdat/utils.py:59-98 (Transitions.states_prev/next,
 *           transitions.py:254-278)
 * ------------------------------------------------------------------------- */

/* Scratch sizing: number of int32 elements gd_events_count needs in d_work. */
int64_t gd_events_work_elems(int64_t n_frames, int32_t n_atoms);

/* Phase 1.  d_states_* are [n_frames,n_atoms] int32.  d_next_outer/inner
 * ([n_atoms], may be NULL) is the first row of the FOLLOWING frame shard
 * (multi-GPU halo); NULL means this shard ends the trajectory.
 * d_changed [2*n_atoms] uint8 (zero-initialised by caller): outer / inner state of
 * the atom changes somewhere in this shard.  Asynchronous. */
gd_status_t gd_events_count(const int32_t *d_states_outer, const int32_t *d_states_inner,
                            int64_t n_frames, int32_t n_atoms, const int32_t *d_next_outer,
                            const int32_t *d_next_inner, int32_t *d_work, uint8_t *d_changed,
                            void *stream);

/* Phase 2.  d_changed_global is d_changed OR-reduced over all shards (same
 * buffer on one GPU).  Computes row offsets, returns the number of events of
 * this shard in *h_n_events and the status of the reference's error paths
 * (GD_ERR_INNER_CONSTANT, GD_ERR_NO_EVENTS is left to the caller because it
 * is a whole-trajectory property).  d_atom_offsets [n_atoms+1] int64 receives
 * the first row of every atom.  Synchronises the stream -- unless h_n_events is NULL:
 * then the total stays in d_atom_offsets[n_atoms] and the GD_ERR_INNER_CONSTANT condition
 * in d_work[gd_events_work_elems() - 16] (non-zero) for the caller to read later. */
gd_status_t gd_events_scan(int32_t *d_work, int64_t n_frames, int32_t n_atoms,
                           const uint8_t *d_changed_global, int64_t *d_atom_offsets,
                           int64_t *h_n_events, void *stream);

/* Phase 3.  d_events [n_events,6] int32, reference column order
 * (atom index, start site, destination site, start inner site,
 *  destination inner site, time), sorted by atom then time; `frame0` is added
 * to the time column (global frame index of the shard's first frame). */
gd_status_t gd_events_fill(const int32_t *d_states_outer, const int32_t *d_states_inner,
                           int64_t n_frames, int32_t n_atoms, const int32_t *d_next_outer,
                           const int32_t *d_next_inner, const int32_t *d_work,
                           const uint8_t *d_changed_global, const int64_t *d_atom_offsets,
                           int64_t frame0, int32_t *d_events, void *stream);

/* Frame-sharded runs: place one shard's rows (sorted by atom, then time) into the
 * merged table.  d_dst_start [n_atoms] int64 = first merged row for this shard's
 * rows of each atom (rows of earlier shards of the same atom come first). */
gd_status_t gd_events_merge(const int32_t *d_src_rows, const int64_t *d_src_atom_offsets,
                            const int64_t *d_dst_start, int32_t n_atoms, int32_t *d_dst_rows,
                            void *stream);

/* gd_events_merge with offsets the host has not validated yet (optimistic pass): atoms whose
 * rows would leave the source (src_capacity rows) or the destination (dst_capacity rows) are
 * skipped instead of copied. */
gd_status_t gd_events_merge_capped(const int32_t *d_src_rows, const int64_t *d_src_atom_offsets,
                                   const int64_t *d_dst_start, int32_t n_atoms, int32_t *d_dst_rows,
                                   int64_t src_capacity, int64_t dst_capacity, void *stream);

/* All shards' jump tables -> the merged table in ONE launch.  d_gathered: the all-gather of every
 * shard's buffer, `stride_words` int32 apart; each buffer starts with `head_words` int32 whose
 * first n_atoms + 1 int64 are the shard's per-atom row offsets, followed by its rows (6 int32
 * each, src_capacity rows).  The placement (rows of atoms < a, then earlier shards' rows of a) is
 * derived from the offsets inside the kernel; offsets that would leave a table are skipped like
 * in gd_events_merge_capped. */
gd_status_t gd_jumps_merge_shards(const int32_t *d_gathered, int64_t stride_words, int32_t head_words,
                                  int32_t world, int32_t n_atoms, int64_t src_capacity,
                                  int32_t *d_dst_rows, int64_t dst_capacity, void *stream);

/* The fill for rows staged by gd_site_states_staged: copies every (frame chunk, atom)
 * segment's rows to their place in the table (same layout and order as gd_events_fill) and
 * adds frame0 to the time column.  d_work / d_changed_global / d_atom_offsets as left by
 * gd_events_scan.  Moves 2 x 24 B per event instead of re-reading 8 B per atom-frame.
 * row_capacity > 0: d_events holds that many rows (a size remembered from an earlier run, the
 * count not yet read back); rows beyond it are not written. */
gd_status_t gd_events_compact(const int32_t *d_stage, int32_t stage_cap, int64_t n_frames,
                              int32_t n_atoms, const int32_t *d_work,
                              const uint8_t *d_changed_global, const int64_t *d_atom_offsets,
                              int64_t frame0, int32_t *d_events, int64_t row_capacity, void *stream);

/* ffill (direction=+1, states_prev) / bfill (direction=-1, states_next) along the
 * frame axis.  d_carry_in [n_atoms] (may be NULL) is the last valid state before
 * (after, for bfill) this shard; d_carry_out [n_atoms] (may be NULL) receives the
 * last valid state at the end of the shard.  d_work: gd_events_work_elems ints. */
gd_status_t gd_fill_states(const int32_t *d_states, int64_t n_frames, int32_t n_atoms, int direction,
                           const int32_t *d_carry_in, int32_t *d_carry_out, int32_t *d_work,
                           int32_t *d_out, void *stream);

/* gd_fill_states without a caller-provided d_work and without any allocation.  NO COUNTERPART
 * IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The per-(frame chunk, atom) carries live in a file-scope `static __device__` array of
 * GD_FILL_NOALLOC_MAX_WORK_ELEMS int32 (16 MiB, held by every GPU that loads the library until
 * exit; one BASELINE config-3 trajectory, 10^6 frames x 1008 atoms, needs 3 938 256 slots).  The
 * call makes no cudaMalloc / cudaFree and calls no allocating routine; inputs and outputs are
 * the caller's.  Shapes whose gd_events_work_elems(n_frames, n_atoms) - 16 exceeds the maximum
 * are rejected with GD_ERR_CAPACITY (use gd_fill_states).  All calls share the one array per
 * device, which keeps the last call's carries (every slot read is overwritten first, nothing
 * needs clearing): calls that overlap on the GPU -- different streams, or two threads -- RACE;
 * order them on one stream or with events.  Same arguments and results as gd_fill_states. */
#define GD_FILL_NOALLOC_MAX_WORK_ELEMS (1 << 22)
gd_status_t gd_fill_states_noalloc(const int32_t *d_states, int64_t n_frames, int32_t n_atoms,
                                   int direction, const int32_t *d_carry_in, int32_t *d_carry_out,
                                   int32_t *d_out, void *stream);

/* The same fill in two steps, for frame shards (one pass over the states less than calling
 * gd_fill_states twice):
 *   1. gd_fill_last_valid: d_last_valid [n_atoms] = last valid state of this shard in scan
 *      direction (NOSITE if it has none); leaves the per-chunk carries in d_work;
 *   2. the caller all-gathers d_last_valid and picks, per atom, the nearest valid entry of
 *      the shards before (ffill) / after (bfill) this one -> d_carry_in (NULL: none);
 *   3. gd_fill_apply writes d_out from d_work and d_carry_in. */
gd_status_t gd_fill_last_valid(const int32_t *d_states, int64_t n_frames, int32_t n_atoms,
                               int direction, int32_t *d_work, int32_t *d_last_valid, void *stream);
gd_status_t gd_fill_apply(const int32_t *d_states, int64_t n_frames, int32_t n_atoms, int direction,
                          const int32_t *d_carry_in, const int32_t *d_work, int32_t *d_out,
                          void *stream);

/* ------------------------------------------------------------------------- *
 * K3  transitions -> jumps
 * replaces: _generic_transitions_to_jumps, src/gemdat/jumps.py:24-119 (the
 *           pandas groupby/iterrows state machine)
 * ------------------------------------------------------------------------- */

/* Scratch (bytes, 16-byte aligned) the two calls below need in d_work. */
int64_t gd_jumps_work_bytes(int64_t n_events, int32_t n_atoms);

/* d_events [n_events,6] int32 sorted by atom then time, d_atom_offsets
 * [n_atoms+1] int64.  n_frames_total = len(transitions.trajectory) (jumps.py:39).
 * The per-atom state machine runs time-parallel over segments of 512 events
 * (speculate, chain, scan).  Writes *h_n_jumps; synchronises. */
gd_status_t gd_jumps_count(const int32_t *d_events, const int64_t *d_atom_offsets, int64_t n_events,
                           int32_t n_atoms, int64_t n_frames_total, int64_t minimal_residence,
                           void *d_work, int64_t *h_n_jumps, void *stream);

/* d_jumps [n_jumps,6] int32: atom index, start site, destination site,
 * start time, stop time, residence_time (jumps.py:110-119).  d_work as left by
 * gd_jumps_count for the same arguments. */
gd_status_t gd_jumps_fill(const int32_t *d_events, const int64_t *d_atom_offsets, int64_t n_events,
                          int32_t n_atoms, int64_t n_frames_total, int64_t minimal_residence,
                          const void *d_work, int32_t *d_jumps, void *stream);

/* --- K3 over frame shards (no event exchange) ------------------------------------
 * The reference runs the machine of jumps.py:45-105 over an atom's whole event list.
 * With the frame axis sharded, every shard runs it over its own time slice of every
 * atom; what crosses a shard boundary is the machine's state per atom, one opaque
 * record of GD_JUMP_STATE_BYTES (pending `fromevent`, pending `candidate_jump`, and
 * flags: shard has events of the atom / one of them is a departure, after which the
 * state no longer depends on how the shard was entered / the state is the true one).
 *
 *   1. gd_jumps_count_shard(h_n_jumps = NULL, h_flags = NULL) -- asynchronous -- leaves in
 *      d_exit [n_atoms] every atom's record when run from the empty state: d_entry = NULL on
 *      the shard that starts the trajectory (empty IS its true entry), all-zero records on the
 *      others ("empty, not known").  An atom whose time-parallel chain could not be proven
 *      publishes "not known";
 *   2. the caller all-gathers the records of all shards ([world, n_atoms]);
 *   3. gd_jumps_resolve gives this shard's true entry records and the number of
 *      (shard, atom) pairs over ALL shards whose entry is not yet known (the same number on
 *      every rank).  While it is > 0 (an atom with events but no departure in a shard --
 *      rare): call gd_jumps_count_shard with the current d_entry and h_flags (checked pass),
 *      gather, resolve again; at most world-1 rounds;
 *   4. gd_jumps_count_shard(d_entry, d_published = this shard's records of the FIRST gather)
 *      counts, gd_jumps_fill writes this shard's jumps; d_atom_jump_offsets [n_atoms+1] places
 *      them per atom when the shards' tables are merged (gd_events_merge takes any 6-int32 rows).
 *      Steps 3 and 4 need only one synchronisation when gd_jumps_resolve leaves its count at
 *      d_work + gd_jumps_work_unresolved_offset() (h_unresolved = NULL): step 4 reports it in
 *      *h_flags bit 2 and the caller repeats from step 3's "while" if it is set.
 * is_last_shard: close pending candidates at n_frames_total (jumps.py:97-105) -- only the
 * shard holding the trajectory's last frame does.  *h_flags: bit 0 the sequential chain
 * ran, bit 1 a record published as known differs from the true exit state (event tables
 * that are not state-consistent: fall back to exchanging the events), bit 2 see above,
 * bit 3 d_atom_offsets says more rows than n_events (nothing was done).
 * reuse_work != 0: d_work still holds the segment speculation of an earlier call on the same
 * table (passes 2.. of the protocol skip it).  Calls with h_n_jumps or h_flags synchronise
 * (the time-parallel chain's proof flag is read back). */
#define GD_JUMP_STATE_BYTES 40
/* Asynchronous FINAL pass (h_n_jumps = h_flags = NULL, d_atom_jump_offsets given, [n_atoms + 2]):
 * nothing is read back; d_atom_jump_offsets[n_atoms] is the jump count and [n_atoms + 1] a status
 * word = (rows of this shard << 8) | GD_JUMPS_TAIL_* bits.  n_events is then the CAPACITY of
 * d_events (a size remembered from an earlier run): with more rows than that the pass does
 * nothing and sets GD_JUMPS_TAIL_ROWS_OVERFLOW.  Any bit set: repeat with a checked pass. */
#define GD_JUMPS_TAIL_MISMATCH 1      /* a record published as known differs from the truth */
#define GD_JUMPS_TAIL_UNPROVEN 2      /* the time-parallel chain could not prove a segment  */
#define GD_JUMPS_TAIL_UNRESOLVED 4    /* gd_jumps_resolve left entries unknown              */
#define GD_JUMPS_TAIL_ROWS_OVERFLOW 8 /* more event rows than n_events                      */
gd_status_t gd_jumps_count_shard(const int32_t *d_events, const int64_t *d_atom_offsets,
                                 int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                 int64_t minimal_residence, const void *d_entry, int is_last_shard,
                                 int reuse_work, const void *d_published, void *d_work, void *d_exit,
                                 int64_t *d_atom_jump_offsets, int64_t *h_n_jumps, int32_t *h_flags,
                                 void *stream);
/* d_all_records [world, n_atoms] records; d_entry [n_atoms] records out; d_unresolved one
 * int32 of device memory receiving the count.  Synchronises unless h_unresolved is NULL. */
int64_t gd_jumps_work_unresolved_offset(int64_t n_events, int32_t n_atoms);
gd_status_t gd_jumps_resolve(const void *d_all_records, int32_t world, int32_t rank, int32_t n_atoms,
                             void *d_entry, int32_t *d_unresolved, int32_t *h_unresolved,
                             void *stream);

/* gd_jumps_fill into a table of jump_capacity rows (a size remembered from an earlier run, the
 * count not read back yet): segments whose jumps would not fit are not written. */
gd_status_t gd_jumps_fill_capped(const int32_t *d_events, const int64_t *d_atom_offsets,
                                 int64_t n_events, int32_t n_atoms, int64_t n_frames_total,
                                 int64_t minimal_residence, const void *d_work, int32_t *d_jumps,
                                 int64_t jump_capacity, void *stream);

/* (start, destination) count matrix: _calculate_transitions_matrix,
 * src/gemdat/transitions.py:812-834.  d_rows [n_rows,row_stride] int32 with the
 * start site in column col_start and the destination in col_dest; d_scratch
 * [(n_sites+1)^2] uint64 zero-initialised by the caller; d_matrix [n_sites,n_sites]
 * int64 (fully written).  Index -1 (NOSITE) aliases n_sites-1 and, as with
 * numpy's fancy assignment of np.unique rows, the last source pair wins. */
gd_status_t gd_pair_matrix(const int32_t *d_rows, int64_t n_rows, int32_t row_stride,
                           int32_t col_start, int32_t col_dest, int32_t n_sites,
                           unsigned long long *d_scratch, long long *d_matrix, void *stream);

/* ------------------------------------------------------------------------- *
 * N1  occupancy, residence times, split
 * replaces: Transitions.occupancy (src/gemdat/transitions.py:280-306: np.unique over the
 *           [T, A] states), Transitions.residence_time (:339-373: Python loop over atoms,
 *           np.diff / np.nonzero per column), Transitions.split and
 *           _split_transitions_events (:375-410, :837-879: boolean masks per part)
 * ------------------------------------------------------------------------- */

/* d_counts [n_sites + 1] uint64, ADDED to (zero it first): d_counts[0] = number of NOSITE
 * entries, d_counts[s + 1] = number of entries equal to site s, over the n int32 entries of
 * d_states (any shape; frame shards add up).  occupancy of site s = d_counts[s + 1] / n_frames. */
gd_status_t gd_occupancy(const int32_t *d_states, int64_t n, int32_t n_sites,
                         unsigned long long *d_counts, void *stream);

/* Residence times from the event table (d_events [n_events, 6] int32, sorted by atom, then
 * time).  A boundary of the reference's np.diff(states[:, atom]) is exactly a row whose start
 * site differs from its destination site; the visit it opens lasts until the next such row of
 * the same atom.  Reported: (atom index, site, time) int32 per visit in (atom, time) order,
 * without visits of NOSITE and without an atom's first and last visit (transitions.py:359-367).
 * Two-phase: gd_residence_count returns the number of rows (synchronises), the caller
 * allocates d_rows [n_rows, 3], gd_residence_fill writes them.  d_work:
 * gd_residence_work_bytes(n_events) bytes, 8-byte aligned. */
int64_t gd_residence_work_bytes(int64_t n_events);
gd_status_t gd_residence_count(const int32_t *d_events, int64_t n_events, void *d_work,
                               int64_t *h_n_rows, void *stream);
gd_status_t gd_residence_fill(const int32_t *d_events, int64_t n_events, const void *d_work,
                              int32_t *d_rows, void *stream);

/* _split_transitions_events: part p holds the rows with d_bins[p] <= time < d_bins[p + 1]
 * (d_bins [n_parts + 1] int64 on the device, np.linspace(0, n_states + 1, n_parts + 1,
 * dtype=int)), in their (atom, time) order, with d_bins[p] subtracted from the time column.
 *   gd_events_split_count: d_part_offsets [n_parts, n_atoms + 1] int64 = per-atom row offsets
 *     inside every part (last column = rows of the part), d_src_lo [n_parts, n_atoms] int64 =
 *     first source row of (part, atom).  Asynchronous.
 *   gd_events_split_fill: d_part_base [n_parts] int64 = first row of part p in d_rows (all
 *     parts back to back, [sum of rows, 6] int32); d_index (may be NULL) receives the source
 *     row number of every row -- the index labels the reference's boolean mask keeps. */
gd_status_t gd_events_split_count(const int32_t *d_events, const int64_t *d_atom_offsets,
                                  int32_t n_atoms, const int64_t *d_bins, int32_t n_parts,
                                  int64_t *d_part_offsets, int64_t *d_src_lo, void *stream);
gd_status_t gd_events_split_fill(const int32_t *d_events, int32_t n_atoms, const int64_t *d_bins,
                                 int32_t n_parts, const int64_t *d_part_offsets,
                                 const int64_t *d_src_lo, const int64_t *d_part_base,
                                 int32_t *d_rows, int64_t *d_index, void *stream);

/* ------------------------------------------------------------------------- *
 * N3  collective jumps
 * replaces: Collective._compute, src/gemdat/collective.py:61-106 (the pandas
 * double loop over jump pairs), called from Jumps.collective, jumps.py:223-251.
 * ------------------------------------------------------------------------- */

/* Scratch bytes for the four calls below (sort keys, counts, offsets, flags). */
int64_t gd_collective_work_bytes(int64_t n_jumps);

/* d_jumps [n_jumps,6] int32 (the K3 table) -> d_sorted, rows ordered by
 * (stop time, start time), equal keys in their incoming order
 * (sort_values(['stop time','start time']), collective.py:68). */
gd_status_t gd_collective_sort(const int32_t *d_jumps, int64_t n_jumps, void *d_work,
                               int32_t *d_sorted, void *stream);

/* d_close[a,b] = d_dist[a,b] < max_dist over the [n_sites,n_sites] fp64 distance
 * matrix from gd_pbc_distances (np.any(dists < max_dist), collective.py:86-88). */
gd_status_t gd_collective_close(const double *d_dist, int32_t n_sites, double max_dist,
                                uint8_t *d_close, void *stream);

/* Walks every jump i over its partners j > i of the sorted table with the
 * reference's break / continue rules (collective.py:76-81).  Returns the number of
 * collective pairs and the number of jumps that are in no pair (n_solo_jumps,
 * :100-101); leaves the per-jump pair offsets in d_work for the fill.
 * Synchronises the stream. */
gd_status_t gd_collective_count(const int32_t *d_sorted, int64_t n_jumps, const uint8_t *d_close,
                                int32_t n_sites, int32_t max_steps, void *d_work,
                                int64_t *h_n_pairs, int64_t *h_n_solo, void *stream);

/* d_pairs [n_pairs,2] int32: row numbers (i, j) in the sorted table, in the order the
 * reference appends them (i ascending, then j ascending). */
gd_status_t gd_collective_fill(const int32_t *d_sorted, int64_t n_jumps, const uint8_t *d_close,
                               int32_t n_sites, int32_t max_steps, const void *d_work,
                               int32_t *d_pairs, void *stream);

/* ------------------------------------------------------------------------- *
 * N4  radial distribution functions
 * replaces: radial_distribution, src/gemdat/rdf.py:111-190 (per-frame loop over
 * lattice.get_all_distances + np.digitize + np.bincount per transition state) and
 * radial_distribution_between_species, rdf.py:193-268 (np.histogram per frame).
 * ------------------------------------------------------------------------- */

/* Transition state of every floating atom in every frame (_get_states_array,
 * rdf.py:53-61).  d_states / _prev / _next [n_frames,n_floating] int32 (site index or
 * -1; prev = ffill, next = bfill, gd_fill_states).  d_site_label [n_sites]: label
 * index of each site, n_labels < 1000.  _uniqify_labels' off-by-one is kept: a state
 * s >= 1 takes the label of site s-1, states 0 and -1 count as "no site" (:19-27).
 * d_lut [(n_labels+1)^3]: class of the dense code ((i+1)(L+1) + (j+1))(L+1) + (k+1),
 * i/j/k = label of now/prev/next or -1 (the caller maps codes to state names,
 * rdf.py:30-50; a negative entry drops the atom-frame).  d_class [n_frames,n_floating]
 * is written; d_first_key [n_class] uint64, initialised by the caller to a large value,
 * receives min((frame << 32) | code) per class: the order in which the reference's
 * result dict gains its keys. */
gd_status_t gd_rdf_classes(const int32_t *d_states, const int32_t *d_states_prev,
                           const int32_t *d_states_next, int64_t n_frames, int32_t n_floating,
                           const int32_t *d_site_label, int32_t n_sites, int32_t n_labels,
                           const int32_t *d_lut, int32_t *d_class,
                           unsigned long long *d_first_key, void *stream);

/* Distance histograms over all frames.  d_pos_frac [n_frames,n_atoms,3] fractional;
 * sources d_idx1 [n1] and targets d_idx2 [n2] are atom indices; d_class [n_frames,n1]
 * (NULL: one class) and d_group [n2] (NULL: one group) select the table the pair is
 * counted in.  Distances follow Lattice.get_all_distances (LLL-reduced basis, both ends wrapped
 * to [0,1) there, shortest of the 27 images; as gd_pbc_distances).  d_edges [n_edges] are the caller's exact bin edges
 * (np.arange(0, max_dist + resolution, resolution)), `resolution` only seeds the
 * search.  mode 0: np.digitize(d, edges, right=True), index = #edges < d;
 * mode 1: np.histogram(d, edges), index = #edges <= d with d == edges[-1] counted in
 * the last bin, i.e. bin b of np.histogram is index b+1.  skip_same_atom drops pairs
 * of an atom with itself.  d_hist [n_class,n_group,n_edges+1] uint64 is ADDED to
 * (zero it first). */
gd_status_t gd_rdf_hist(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                        const double *h_lattice, const int32_t *d_idx1, int32_t n1,
                        const int32_t *d_class, int32_t n_class, const int32_t *d_idx2,
                        const int32_t *d_group, int32_t n2, int32_t n_group,
                        const double *d_edges, int32_t n_edges, double resolution, int32_t mode,
                        int32_t skip_same_atom, unsigned long long *d_hist, void *stream);

/* The same histograms for a trajectory whose lattice changes from frame to frame
 * (radial_distribution_between_species with constant_lattice=False, rdf.py:232-261).
 * h_lattices [n_frames,9] row-major; the reduced cell of every frame is built on the host and
 * copied into d_work (gd_rdf_frames_work_bytes(n_frames) bytes, 8-byte aligned; the call
 * synchronises the stream once for that copy).  The counts of frame t go to their own table:
 * d_hist [n_frames,n_group,n_edges+1] uint64 is ADDED to -- the reference normalises every
 * frame with its own volume, which the caller does on these per-frame counts.  d_skip
 * [n_frames,n1] (NULL: none): a negative entry drops that source in that frame. */
int64_t gd_rdf_frames_work_bytes(int64_t n_frames);
gd_status_t gd_rdf_hist_frames(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                               const double *h_lattices, void *d_work, const int32_t *d_idx1,
                               int32_t n1, const int32_t *d_skip, const int32_t *d_idx2,
                               const int32_t *d_group, int32_t n2, int32_t n_group,
                               const double *d_edges, int32_t n_edges, double resolution,
                               int32_t mode, int32_t skip_same_atom, unsigned long long *d_hist,
                               void *stream);

/* ------------------------------------------------------------------------- *
 * K0  displacement / unwrap
 * replaces: pymatgen Trajectory.to_displacements via Trajectory.displacements
 *           (src/gemdat/trajectory.py:182-193), cumulative_displacements
 *           (:705-713), distances_from_base_position (:715-730), _lengths (:40-60)
 * ------------------------------------------------------------------------- */

/* d_pos_frac [n_frames,n_atoms,3] wrapped fractional positions.
 * d_prev_row [n_atoms,3] last row of the preceding shard (NULL: first shard, d[0]=0).
 * d_base_winding [n_atoms,3] int32 lattice translations accumulated before this
 *   shard (NULL = 0); d_base_pos [n_atoms,3] positions of GLOBAL frame 0 (NULL:
 *   row 0 of this shard).
 * Outputs (any may be NULL): d_cum_frac [n_frames,n_atoms,3] cumulative fractional
 *   displacement; d_cum_cart same in Angstrom (cum_frac @ lattice); d_dist
 *   [n_atoms,n_frames] distance from the base position (note: atom-major, as the
 *   reference returns it); d_speed [n_frames,n_atoms] = np.diff(distance, prepend=0) along
 *   time (TrajectoryMetrics.speed, src/gemdat/metrics.py:35-47) -- TIME-major, the layout
 *   gd_amplitudes_* stream (gd_transpose_f64 gives the reference's [n_atoms,n_frames]);
 *   d_winding_out [n_atoms,3] int32 winding at shard end.
 * d_work: n_chunks*n_atoms*3 int32, n_chunks = gd_unwrap_chunks(n_frames, n_atoms). */
int64_t gd_unwrap_chunks(int64_t n_frames, int32_t n_atoms);
gd_status_t gd_unwrap(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                      const double *h_lattice, const double *d_prev_row,
                      const int32_t *d_base_winding, const double *d_base_pos, int32_t *d_work,
                      double *d_cum_frac, double *d_cum_cart, double *d_dist, double *d_speed,
                      int32_t *d_winding_out, void *stream);

/* gd_unwrap in two steps, for frame shards (the winding a shard starts from is known only
 * after every shard has counted its own):
 *   1. gd_unwrap_windings: d_winding_out [n_atoms,3] = lattice translations inside this shard
 *      (incl. the step from d_prev_row into its first frame); d_work keeps the per-chunk
 *      prefixes;
 *   2. the caller all-gathers d_winding_out and sums the shards before this one (integers:
 *      exact) -> d_base_winding;
 *   3. gd_unwrap_apply writes the outputs from d_work (as left by step 1) + d_base_winding. */
gd_status_t gd_unwrap_windings(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                               const double *d_prev_row, int32_t *d_work, int32_t *d_winding_out,
                               void *stream);
gd_status_t gd_unwrap_apply(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                            const double *h_lattice, const double *d_prev_row,
                            const int32_t *d_base_winding, const double *d_base_pos,
                            const int32_t *d_work, double *d_cum_frac, double *d_cum_cart,
                            double *d_dist, double *d_speed, void *stream);

/* ------------------------------------------------------------------------- *
 * Per-frame averages over atoms and the drift correction (callers on the input side of K0)
 * replaces: Trajectory.center_of_mass (src/gemdat/trajectory.py:732-749: np.average of
 *           base_positions + cumulative_displacements over the atoms, weights = masses),
 *           Trajectory.drift (:751-791: np.mean of the fixed species' displacements per
 *           frame) and Trajectory.apply_drift_correction (:793-830: displacements - drift,
 *           later summed along the frames and wrapped by to_positions, :121-127).
 * ------------------------------------------------------------------------- */

/* d_out [n_frames,3] = sum_a w[a] (base[a] + x[t,a]) / sum_a w[a].  d_x [n_frames,n_atoms,3];
 * d_base [n_atoms,3] or NULL (nothing added); d_weights [n_atoms] or NULL (plain mean).
 * One CTA per frame, fixed reduction order (deterministic). */
gd_status_t gd_frame_mean(const double *d_x, const double *d_base, const double *d_weights,
                          int64_t n_frames, int32_t n_atoms, double *d_out, void *stream);

/* d_drift [n_frames,3] = mean over the selected atoms of the wrapped step
 * (x[t] - x[t-1]) - rint(x[t] - x[t-1]) of the fractional positions d_pos [n_frames,n_atoms,3].
 * d_selected [n_selected] atom numbers, or NULL with n_selected == n_atoms (all atoms).
 * Frame 0 has step 0 (pymatgen to_displacements) unless d_prev_row [n_atoms,3], the last
 * row of the frame shard before this one, is given. */
gd_status_t gd_drift(const double *d_pos, int64_t n_frames, int32_t n_atoms,
                     const int32_t *d_selected, int32_t n_selected, const double *d_prev_row,
                     double *d_drift, void *stream);

/* inclusive sum along the frames of d_in [n_frames,3] -> d_out [n_frames,3] (block scans of
 * 1024 frames + carries).  d_work: gd_cumsum_frames3_work_bytes(n_frames) bytes, 8-byte
 * aligned; d_out must not alias d_in. */
int64_t gd_cumsum_frames3_work_bytes(int64_t n_frames);
gd_status_t gd_cumsum_frames3(const double *d_in, int64_t n_frames, void *d_work, double *d_out,
                              void *stream);

/* d_out [n_frames,n_atoms,3] = np.mod(d_base[a] + d_cumulative[t,a] - d_cumulative_drift[t], 1):
 * the drift-corrected positions from K0's cumulative displacements (gd_unwrap d_cum_frac) and
 * the summed drift.  d_out may be d_cumulative (in place). */
gd_status_t gd_drift_correct(const double *d_cumulative, const double *d_base,
                             const double *d_cumulative_drift, int64_t n_frames, int32_t n_atoms,
                             double *d_out, void *stream);

/* ------------------------------------------------------------------------- *
 * a4  vibration amplitudes
 * replaces: TrajectoryMetrics.amplitudes (src/gemdat/metrics.py:209-231: Python loop over
 *           atoms, np.sign / np.roll / np.array_split / np.sum) and the np.mean / np.std of
 *           vibration_amplitude (:190-207)
 * ------------------------------------------------------------------------- */

/* d_speed_tm [n_frames,n_atoms] (gd_unwrap d_speed).  Every atom's row is cut after each sign
 * flip except its first and last one and the pieces are summed; amplitudes of atom a are
 * d_amplitudes[d_atom_offsets[a] .. d_atom_offsets[a+1]) in time order, atoms ascending -- the
 * order of the reference's list.  Two-phase: gd_amplitudes_count fills d_atom_offsets
 * [n_atoms+1] int64 and returns the total (synchronises when h_n_amplitudes is given);
 * gd_amplitudes_fill writes d_amplitudes.  d_work: gd_amplitudes_work_bytes() bytes, 16-byte
 * aligned, untouched between the two calls. */
int64_t gd_amplitudes_work_bytes(int64_t n_frames, int32_t n_atoms);
gd_status_t gd_amplitudes_count(const double *d_speed_tm, int64_t n_frames, int32_t n_atoms,
                                void *d_work, int64_t *d_atom_offsets, int64_t *h_n_amplitudes,
                                void *stream);
gd_status_t gd_amplitudes_fill(const double *d_speed_tm, int64_t n_frames, int32_t n_atoms,
                               const void *d_work, const int64_t *d_atom_offsets,
                               double *d_amplitudes, void *stream);

/* np.mean(x) and np.std(x) (population: sqrt(mean((x - mean)^2))) of n doubles in two
 * deterministic reduction passes.  d_work: gd_mean_std_work_bytes() bytes.  Synchronises. */
int64_t gd_mean_std_work_bytes(void);
gd_status_t gd_mean_std(const double *d_x, int64_t n, void *d_work, double *h_mean, double *h_std,
                        void *stream);

/* ------------------------------------------------------------------------- *
 * N2  attempt frequency: mean frequency of every atom's speed
 * replaces: meanfreq, src/gemdat/utils.py:128-160 (the weighting of
 *           scipy.signal.periodogram's one-sided spectrum and the np.dot / np.sum
 *           reductions), called from TrajectoryMetrics.attempt_frequency, metrics.py:167-188
 * ------------------------------------------------------------------------- */

/* d_spectrum [n_atoms, n_samples/2 + 1] complex128 (interleaved re, im): the one-sided DFT of
 * every atom's speed row (the transform itself is the caller's: cuFFT; gd_transpose_f64 turns
 * the time-major speed array of gd_unwrap into rows per atom).  d_meanfreq [n_atoms] =
 * sum_k f_k P_k / sum_k P_k in Hz with P_k = |X_k|^2, doubled except at DC and Nyquist,
 * f_k = k * sampling_frequency / n_samples; the DC bin is skipped (constant detrend). */
gd_status_t gd_meanfreq(const double *d_spectrum, int64_t n_samples, int32_t n_atoms,
                        double sampling_frequency, double *d_meanfreq, void *stream);

/* ------------------------------------------------------------------------- *
 * K4  all-lag mean squared displacement
 * replaces: Trajectory.mean_squared_displacement, src/gemdat/trajectory.py:896-933
 *           (numpy FFT autocorrelation); computed here by direct summation.
 * ------------------------------------------------------------------------- */

/* [n_frames,n_atoms,3] -> component-major per atom [n_atoms,3,n_frames] (the layout
 * gd_msd_direct streams). */
gd_status_t gd_transpose_soa(const double *d_in, int64_t n_frames, int32_t n_atoms, double *d_out,
                             void *stream);

/* out[c, r] = in[r, c] for a row-major [n_rows, n_cols] float64 array. */
gd_status_t gd_transpose_f64(const double *d_in, int64_t n_rows, int64_t n_cols, double *d_out,
                             void *stream);

/* d_soa [n_atoms,3,n_frames] unwrapped Cartesian displacement (gd_unwrap d_cum_cart
 * through gd_transpose_soa).  Lags lag0 .. lag0+n_lags-1.  d_msd [n_atoms,n_lags]
 * float64: msd[a,j] = mean_t |r[t+lag0+j]-r[t]|^2 (0 for lag 0). */
gd_status_t gd_msd_direct(const double *d_soa, int64_t n_frames, int32_t n_atoms, int64_t lag0,
                          int64_t n_lags, double *d_msd, void *stream);

/* ------------------------------------------------------------------------- *
 * K5  density histogram
 * replaces: trajectory_to_volume, src/gemdat/volume.py:449-509
 *           (np.digitize against np.linspace edges + np.unique(axis=0) counting)
 * ------------------------------------------------------------------------- */

/* nedges[3] = int(1 + length // resolution) per axis (volume.py:482-484); the
 * grid has (nedges-1) voxels per axis.  d_grid [(nx-1)*(ny-1)*(nz-1)] uint32,
 * accumulated into (caller zero-initialises).  d_range_flag [1] int32 is set to 1
 * if any coordinate is outside [0,1) (the reference asserts). */
gd_status_t gd_density(const double *d_pos_frac, int64_t n_atom_frames, const int32_t *h_nedges,
                       uint32_t *d_grid, int32_t *d_range_flag, void *stream);

/* The same histogram for positions known as [n_frames,n_atoms,3] (the product path).  A lane
 * follows one atom through consecutive frames and counts runs inside one voxel in registers:
 * one RED per run instead of one per atom-frame; when the whole grid fits the CTA's shared
 * memory (<= 48k voxels) the runs go to a CTA-private copy first (shared-memory atomics). */
gd_status_t gd_density_frames(const double *d_pos_frac, int64_t n_frames, int32_t n_atoms,
                              const int32_t *h_nedges, uint32_t *d_grid, int32_t *d_range_flag,
                              void *stream);

/* ------------------------------------------------------------------------- *
 * helpers
 * ------------------------------------------------------------------------- */

/* np.mod(coords, 1) of Trajectory.to_positions (src/gemdat/trajectory.py:121-127).
 * d_flag [1] int32 (zero-initialised) is set when some coordinate lies outside [0, 1);
 * d_out (may alias d_in; NULL = only check) receives the wrapped coordinates. */
gd_status_t gd_wrap_unit(const double *d_in, double *d_out, int64_t n, int32_t *d_flag,
                         void *stream);

/* minimum-image distance matrix between two fractional coordinate sets
 * (pymatgen Lattice.get_all_distances as used at transitions.py:543,595 and
 * jumps.py:203): both sets are mapped into the LLL-reduced basis and wrapped to [0,1)
 * there, then the shortest of the 27 neighbouring images is taken (pbc_shortest_vectors).
 * d_out [n1,n2] float64. */
/* The basis reduction behind it, host only (no device needed): h_lll [9] = LLL-reduced rows
 * (pymatgen Lattice.lll_matrix, delta = 0.75), h_mapping [9] the unimodular integer matrix
 * with lll = mapping @ lattice (Lattice.lll_mapping). */
gd_status_t gd_lll_reduce_host(const double *h_lattice, double *h_lll, double *h_mapping);
gd_status_t gd_pbc_distances(const double *h_lattice, const double *d_frac1, int32_t n1,
                             const double *d_frac2, int32_t n2, double *d_out, void *stream);

/* Synthetic trajectory generator (bench / tests only; see gemdat_b200/synth.py for
 * the recipe, which numpy replays bit-identically for any frame window).
 * d_centre/d_delta [n_atoms,n_epochs,3], frames [frame0, frame0+n_frames). */
gd_status_t gd_synth_positions(const double *d_centre, const double *d_delta, int32_t n_atoms,
                               int32_t n_epochs, int32_t epoch_len, int32_t knot_len,
                               const double *h_noise_scale, uint64_t seed, int64_t frame0,
                               int64_t n_frames, double *d_pos_frac, void *stream);

/* fp64 pipe micro-benchmark: runs `iters` dependent DFMA per thread on a full
 * grid and returns the achieved DFMA/s in *h_dfma_per_s (roofline denominator
 * for the fp64-bound kernels; MEASURED_PEAKS.json has no fp64 entry). */
gd_status_t gd_fp64_peak(int64_t iters, double *h_dfma_per_s, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GEMDAT_B200_H */
