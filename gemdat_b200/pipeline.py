"""Device-resident, frame-sharded run of the hot path (K1 -> K2 -> K3).

One process per GPU; each rank owns a contiguous block of frames and nothing O(frames) or
O(events) ever leaves it:

* K1 (site states + event counts + staged event rows) needs one halo row of POSITIONS from the
  following shard -- handed in by the caller or all-gathered before the kernel starts;
* the flags K2 needs globally (visited sites, "state ever changes" per atom, staging overflow)
  travel in one small all-gather and are OR-ed locally;
* K3 runs every atom's jump machine on the rank's own time slice; only the machine's per-atom
  state (40 bytes) crosses a shard boundary (`gd_jumps_count_shard` / `gd_jumps_resolve`);
* the jump tables (small) are all-gathered and merged per atom, so every rank ends with the
  whole table in the reference's order; `gather_events` assembles the event table on demand.

`HotPath.run` has a checked pass (reads every count where it needs it) and an optimistic pass
(table sizes remembered from an earlier run, nothing read back until one host read at the end,
replayed as a single CUDA graph); see DESIGN.md section 5.  All exchanges go through
``torch.distributed`` (NCCL over NVLink on the GPU box, gloo in the CPU tests of the host
logic).  The single-GPU public API (``Trajectory.transitions_between_sites``) runs the same
kernels through ``gemdat_b200.transitions``.
"""

from __future__ import annotations

import ctypes as C
import os
import time
import warnings

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .lattice import site_arrays
from .transitions import SiteGrid, _radius_to_dict


def merge_plan(counts: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """Per-atom placement of per-shard event rows in the merged (atom, time)-sorted table.

    ``counts`` [n_shards, n_atoms] rows per shard and atom.  Returns ``(atom_offsets
    [n_atoms + 1], dst_start [n_shards, n_atoms])``: shard ``r``'s rows of atom ``a`` start at
    ``dst_start[r, a]`` -- after all rows of atoms < a and after earlier shards' rows of a.
    """
    counts = np.asarray(counts, dtype=np.int64)
    per_atom = counts.sum(axis=0)
    atom_offsets = np.concatenate([[0], np.cumsum(per_atom)])
    before = np.cumsum(counts, axis=0) - counts
    return atom_offsets, atom_offsets[:-1][None, :] + before


class Comm:
    """The two collectives the path needs, over ``torch.distributed``.

    NCCL (one rank per GPU, NVLink) takes device tensors directly.  Any other backend (gloo:
    the CPU tests of the host logic, and several ranks sharing one GPU in the GPU tests) is
    staged through host memory -- every exchange on this path is O(n_atoms) or O(n_events).
    """

    def __init__(self, world: int = 1, group=None):
        self.world, self.group = world, group
        self.direct = True
        if world > 1:
            import torch.distributed as dist

            self.dist = dist
            self.direct = dist.get_backend(group) == 'nccl'

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        """[world, *t.shape] on the device of ``t`` (every rank contributes the same shape)."""
        if self.world == 1:
            return t.unsqueeze(0)
        src = t.contiguous() if self.direct else t.detach().cpu().contiguous()
        # flat in / flat out: the one layout every backend accepts
        out = torch.empty((self.world * src.numel(),), dtype=src.dtype, device=src.device)
        self.dist.all_gather_into_tensor(out, src.reshape(-1), group=self.group)
        out = out.reshape((self.world,) + tuple(src.shape))
        return out if self.direct else out.to(t.device)

    def all_gather_rows(self, rows: torch.Tensor) -> torch.Tensor:
        """Concatenation over the ranks (in rank order) of row tables with different lengths."""
        if self.world == 1:
            return rows
        n = torch.tensor([rows.shape[0]], dtype=torch.int64, device=rows.device)
        counts = self.all_gather(n).flatten().tolist()
        cap = max(1, max(counts))
        padded = torch.zeros((cap,) + tuple(rows.shape[1:]), dtype=rows.dtype, device=rows.device)
        padded[: rows.shape[0]] = rows
        allp = self.all_gather(padded)
        return torch.cat([allp[r, : counts[r]] for r in range(self.world)], dim=0)

    def all_to_all_rows(self, rows: torch.Tensor, send: list[int], recv: list[int]) -> list[torch.Tensor]:
        """Rank r receives, from every rank q, the slice of q's ``rows`` meant for r.

        ``rows`` holds the slices for ranks 0..W-1 back to back (``send[r]`` rows each);
        ``recv[q]`` is the number of rows coming from rank q.  NCCL: one grouped send/recv
        (``all_to_all_single``), every row crosses the link once.  Other backends (CPU tests,
        ranks sharing a GPU) emulate it with a padded all-gather.
        """
        width = int(rows.shape[1])
        if self.direct:
            out = torch.empty((sum(recv), width), dtype=rows.dtype, device=rows.device)
            self.dist.all_to_all_single(out, rows.contiguous(), output_split_sizes=recv,
                                        input_split_sizes=send, group=self.group)
            return list(torch.split(out, recv, dim=0))
        me = self.dist.get_rank(self.group)
        sends = self.all_gather(torch.tensor(send, dtype=torch.int64, device=rows.device))  # [W, W]
        sends_h = sends.cpu().numpy()
        cap = max(1, int(sends_h.sum(axis=1).max()))
        padded = torch.zeros((cap, width), dtype=rows.dtype, device=rows.device)
        padded[: rows.shape[0]] = rows
        allp = self.all_gather(padded)
        out = []
        for q in range(self.world):
            start = int(sends_h[q, :me].sum())
            out.append(allp[q, start: start + int(sends_h[q, me])].contiguous())
        return out

    def _all_reduce(self, t: torch.Tensor, op) -> torch.Tensor:
        if self.world == 1:
            return t
        if self.direct:
            self.dist.all_reduce(t, op=op, group=self.group)
        else:
            h = t.detach().cpu()
            self.dist.all_reduce(h, op=op, group=self.group)
            t.copy_(h)
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        """In-place elementwise maximum over the ranks."""
        return self._all_reduce(t, self.dist.ReduceOp.MAX if self.world > 1 else None)

    def all_reduce_min(self, t: torch.Tensor) -> torch.Tensor:
        """In-place elementwise minimum over the ranks."""
        return self._all_reduce(t, self.dist.ReduceOp.MIN if self.world > 1 else None)

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        """In-place elementwise sum over the ranks (voxel grids, histograms, per-lag partial sums)."""
        return self._all_reduce(t, self.dist.ReduceOp.SUM if self.world > 1 else None)

    def all_to_all_flat(self, flat: torch.Tensor, send: list[int], recv: list[int]) -> list[torch.Tensor]:
        """Rank r receives, from every rank q, the ``send[r]`` elements of q's 1-D ``flat`` meant
        for r (``flat`` holds the pieces for ranks 0..W-1 back to back); ``recv[q]`` = elements
        coming from rank q.  Pieces may differ in size (ranks own different numbers of frames)."""
        if self.world == 1:
            return [flat]
        if self.direct:
            out = torch.empty((sum(recv),), dtype=flat.dtype, device=flat.device)
            self.dist.all_to_all_single(out, flat.contiguous(), output_split_sizes=recv,
                                        input_split_sizes=send, group=self.group)
            return list(torch.split(out, recv))
        me = self.dist.get_rank(self.group)
        sends_h = self.all_gather(torch.tensor(send, dtype=torch.int64)).cpu().numpy()  # [W, W]
        cap = max(1, int(sends_h.sum(axis=1).max()))
        padded = torch.zeros((cap,), dtype=flat.dtype, device=flat.device)
        padded[: flat.numel()] = flat
        allp = self.all_gather(padded)
        out = []
        for q in range(self.world):
            start = int(sends_h[q, :me].sum())
            out.append(allp[q, start: start + int(sends_h[q, me])].contiguous())
        return out


class HotPath:
    def __init__(self, *, sites, lattice, n_atoms: int, site_radius, site_inner_fraction=1.0,
                 minimal_residence: int = 0, n_frames_total: int, frame0: int = 0,
                 world: int = 1, rank: int = 0, group=None):
        frac, labels, _ = site_arrays(sites)
        r, f = _radius_to_dict(site_radius, site_inner_fraction)
        self.radius = r
        self.grid = SiteGrid(lattice_matrix=lattice, site_frac=frac, site_labels=labels,
                             site_radius=r, site_inner_fraction=f)
        self.n_atoms = n_atoms
        self.n_sites = len(frac)
        # the visited-site flags only feed the integer_remap quirk (Q2), which exists for labelled
        # radius groups; with the single group '' K1 is not asked for them (2.4 % of its time)
        self._visited_needed = any(key is not None for key in self.grid.group_keys)
        self.minimal_residence = int(minimal_residence)
        self.n_frames_total = int(n_frames_total)
        self.frame0 = int(frame0)
        self.world, self.rank, self.group = world, rank, group
        self.comm = Comm(world, group)
        self.launches_per_run = 0
        self._jump_cap = 0  # rows per rank reserved in the final jump gather (grows with what is seen)
        self._cap_events = 0  # event rows of this rank's table, remembered for the optimistic pass
        # GD_OPTIMISTIC=0: always the checked pass (A/B)
        self.optimistic = os.environ.get('GD_OPTIMISTIC', '1') not in ('', '0')
        # GD_FIRST_CALL_BOUND=0: the first call takes the checked pass (A/B)
        self.first_call_bound = os.environ.get('GD_FIRST_CALL_BOUND', '1') not in ('', '0')
        self._caps_are_bounds = False
        self.last_pass = 'none'
        self.last_optimistic_reject = None
        # GD_GRAPH=0: launch the optimistic pass kernel by kernel instead of replaying a CUDA graph
        self.use_graph = os.environ.get('GD_GRAPH', '1') not in ('', '0')
        self._graph = self._graph_key = self._graph_state = self._eager_key = None
        self._next_row, self._own_halo = None, False
        self.last_launch_mode = 'eager'
        self.last_k1_events = None
        # GD_K3_EXCHANGE=1: always exchange the event rows by atom block (the fallback route; A/B)
        self.k3_exchange = os.environ.get('GD_K3_EXCHANGE', '0') not in ('', '0')
        # profile_phases: synchronise at phase boundaries and keep wall-clock ms per phase in
        # `last_phases` (diagnostics only; the timed runs leave it off)
        # GD_EVENTS_STAGED=0: K2 re-reads the state arrays (gd_events_fill) instead of compacting
        # rows staged by K1 (A/B; also the route taken when a staging segment overflows)
        self.stage_events = os.environ.get('GD_EVENTS_STAGED', '1') not in ('', '0')
        self.last_events_route = 'none'
        self.profile_phases = False
        self.last_phases = {}
        self._buf = {}

    # ------------------------------------------------------------------ helpers
    def _tensor(self, name, shape, dtype):
        t = self._buf.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
            t = dv.empty(shape, dtype)
            self._buf[name] = t
        return t

    def _all_gather(self, t: torch.Tensor) -> torch.Tensor:
        return self.comm.all_gather(t)

    def _pinned(self, name: str, n_bytes: int) -> torch.Tensor:
        """Persistent page-locked host buffer (its address is baked into the captured graph)."""
        t = self._buf.get(name)
        if t is None or t.numel() != n_bytes:
            t = torch.empty((n_bytes,), dtype=torch.uint8, pin_memory=True)
            self._buf[name] = t
        return t

    def _needs_remap(self, visited_bytes: np.ndarray) -> bool:
        """integer_remap quirk (Q2) needed for these visited flags?  The flags rarely change from
        one pass to the next, so the answer is remembered per flag pattern."""
        key = visited_bytes.tobytes()
        if getattr(self, '_remap_key', None) != key:
            vis = np.frombuffer(key, dtype=np.uint8).astype(bool).reshape(2, self.n_sites)
            self._remap_key, self._remap_needed = key, any(self.grid.remap_lut(v) is not None for v in vis)
        return self._remap_needed

    def _mark(self, name: str) -> None:
        if not self.profile_phases:
            return
        torch.cuda.synchronize()
        now = time.perf_counter()
        self.last_phases[name] = self.last_phases.get(name, 0.0) + 1e3 * (now - self._t_mark)
        self._t_mark = now

    # ------------------------------------------------------------------ the pass
    def _launch_k1(self, pos: torch.Tensor, timed: bool = True):
        """K1 on this rank's frames (states, event counts, staged event rows, flags); nothing is
        read back.  Returns the tensors the rest of the pass works on."""
        lib = _lib.lib
        s = dv.stream_ptr()
        n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
        assert n_atoms == self.n_atoms
        launches = 0
        outer = self._tensor('outer', (n_frames, n_atoms), torch.int32)
        inner = self._tensor('inner', (n_frames, n_atoms), torch.int32)
        band = self._tensor('band', (4,), torch.int64)
        band.zero_()
        # every flag the host (and, frame-sharded, the other ranks) must see after K1 lives in ONE
        # buffer: visited [2S] | changed [2A] | staging overflow (int32), 16-byte aligned sections
        n_vis, n_chg = 2 * self.n_sites, 2 * n_atoms
        o_chg = (n_vis + 15) // 16 * 16
        o_ovf = o_chg + (n_chg + 15) // 16 * 16
        pack = self._tensor('pack_k1', (o_ovf + 16,), torch.uint8)
        pack.zero_()
        visited, changed = pack[:n_vis], pack[o_chg: o_chg + n_chg]
        overflow = pack[o_ovf: o_ovf + 4].view(torch.int32)

        # ---- K1 (+ the event count of K2, fused).  The event at a shard's last frame needs the
        # state of the next shard's first frame: positions are known up front, so the halo ROW OF
        # POSITIONS is exchanged before K1 and its state recomputed locally -- K1 never waits for
        # a neighbour's kernel.
        if self.profile_phases:
            torch.cuda.synchronize()
            self.last_phases, self._t_mark = {}, time.perf_counter()
        next_pos = self._next_row  # the caller may hand over the following shard's first row itself
        if self.world > 1 and not self._own_halo:
            first_rows = self._all_gather(pos[0].contiguous())  # [W, A, 3]
            if self.rank + 1 < self.world:
                next_pos = first_rows[self.rank + 1].contiguous()
        work = self._tensor('work', (int(lib.gd_events_work_elems(n_frames, n_atoms)),), torch.int32)
        self._mark('halo_positions')
        # K1 also leaves the event rows themselves in a staging area ([chunk, atom, cap] rows), so
        # that K2 only compacts them instead of re-reading 8 B/atom-frame of states
        stage = None
        cap = 0
        if self.stage_events:  # GD_EVENTS_STAGE_CAP overrides the default rows per (chunk, atom) segment
            cap = int(os.environ.get('GD_EVENTS_STAGE_CAP', '0')) or int(lib.gd_events_stage_cap(n_frames, n_atoms))
        if cap > 0:
            stage = self._tensor('stage', (int(lib.gd_events_stage_elems(n_frames, n_atoms, cap)),), torch.int32)
        if timed:  # (timing events cannot be recorded inside a graph capture)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        _lib.check(lib.gd_site_states_staged(dv.ptr(pos), n_frames, n_atoms, dv.ptr(next_pos),
                                             dv.ptr(self.grid.d_blob), self.grid.h_blob.ctypes.data,
                                             dv.ptr(outer), dv.ptr(inner),
                                             dv.ptr(visited) if self._visited_needed else None,
                                             dv.ptr(band),
                                             dv.ptr(work), dv.ptr(changed), dv.ptr(stage), cap,
                                             dv.ptr(overflow) if stage is not None else None, s))
        if timed:
            e1.record()
            self.last_k1_events = (e0, e1)
        launches += 1
        self._mark('k1_site_states')
        from types import SimpleNamespace
        return SimpleNamespace(outer=outer, inner=inner, band=band, pack=pack, visited=visited, changed=changed,
                               overflow=overflow, work=work, stage=stage, cap=cap, n_vis=n_vis, o_chg=o_chg,
                               o_ovf=o_ovf, launches=launches, n_frames=n_frames, n_atoms=n_atoms)

    def run(self, pos: torch.Tensor, next_pos_row: torch.Tensor | None = None) -> dict:
        """One pass of the hot path over this rank's frames.  ``next_pos_row`` [n_atoms, 3]: the
        positions of the frame after this shard, if the caller can read them itself (every rank but
        the last; saves the halo all-gather) -- all ranks must then do the same.  With capacities remembered from an
        earlier run the whole pass is queued without reading anything back (`_run_optimistic`: one
        host synchronisation at the very end); whatever it cannot prove there -- a table that
        outgrew its remembered size, an unresolved jump-machine entry, the Q2 remap, an error
        condition -- is redone by the checked pass below, which reads every count where it needs it."""
        self._next_row = None
        if next_pos_row is not None and self.rank + 1 < self.world:
            self._next_row = next_pos_row.contiguous()
        self._own_halo = next_pos_row is not None or self.world == 1
        if (self.optimistic and self.stage_events and self._cap_events == 0 and self.first_call_bound
                and not self.k3_exchange):
            # very first call: no table sizes are known yet.  Instead of the checked pass (a host
            # synchronisation for every count) the one-synchronisation pass runs with UPPER BOUNDS:
            # the event table gets as many rows as K1's staging area can hold (more would set the
            # overflow flag anyway), the jump table a generous guess that the status word checks.
            n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
            cap = (int(os.environ.get('GD_EVENTS_STAGE_CAP', '0'))
                   or int(_lib.lib.gd_events_stage_cap(n_frames, n_atoms)))  # as in _launch_k1
            self._cap_events = int(_lib.lib.gd_events_stage_elems(n_frames, n_atoms, cap)) // 6
            # (frame-sharded, the jump rows of all ranks travel in the final all-gather: keep it small)
            self._jump_cap = max(65536, 64 * n_atoms) if self.world > 1 else max(1 << 20, 64 * n_atoms)
            self._caps_are_bounds = True
        if self.optimistic and self._cap_events > 0:
            res = self._run_optimistic(pos)
            if res is not None:
                return res
            if self._caps_are_bounds:  # e.g. more jumps than guessed: size everything exactly
                self._cap_events, self._jump_cap, self._caps_are_bounds = 0, 0, False
        return self._run_checked(pos)

    def _run_checked(self, pos: torch.Tensor) -> dict:
        lib = _lib.lib
        s = dv.stream_ptr()
        k1 = self._launch_k1(pos)
        outer, inner, band, pack, visited, changed = k1.outer, k1.inner, k1.band, k1.pack, k1.visited, k1.changed
        work, stage, cap, n_vis, o_chg, o_ovf = k1.work, k1.stage, k1.cap, k1.n_vis, k1.o_chg, k1.o_ovf
        launches, n_frames, n_atoms = k1.launches, k1.n_frames, k1.n_atoms

        # ---- what K2 needs from the other shards, in ONE exchange: the OR of the visited flags
        # (warning + Q2 LUT are global), of the "state ever changes" flags (Q4 "atom is skipped" is
        # a whole-trajectory property) and of the overflow flag (all ranks take the same route).
        if self.world > 1:
            pack.copy_(self.comm.all_gather(pack).amax(dim=0))
        host = dv.to_host(pack)
        vis = host[:n_vis].astype(bool).reshape(2, self.n_sites)
        any_outer_change = bool(host[o_chg: o_chg + n_atoms].any())
        staged_ok = stage is not None and not host[o_ovf: o_ovf + 4].any()
        remapped = False
        for arr, v in ((outer, vis[0]), (inner, vis[1])):
            for label, key in zip(self.grid.group_labels, self.grid.group_keys):
                if not (v.any() if key is None else v[key].any()):
                    radius = self.radius[label]
                    warnings.warn(f'No floating species in range of {label} ({radius=})', stacklevel=2)
            lut = self.grid.remap_lut(v)
            if lut is not None:
                d_lut = dv.to_device(lut)
                _lib.check(lib.gd_remap_states(dv.ptr(arr), arr.numel(), dv.ptr(d_lut), self.n_sites, s))
                launches += 1
                remapped = True
        next_o = next_i = None
        if self.world > 1 and (remapped or not staged_ok):
            # the re-read route needs the following shard's first state rows (destination states of
            # the halo event; after the Q2 remap they carry the remapped indices)
            rows = self._all_gather(torch.stack([outer[0], inner[0]]))  # [W, 2, A]
            if self.rank + 1 < self.world:
                next_o, next_i = rows[self.rank + 1, 0].contiguous(), rows[self.rank + 1, 1].contiguous()

        self._mark('flags_exchange')
        # ---- K2: scan + fill (the fill writes the halo event with the neighbour's first states)
        offsets = self._tensor('offsets', (n_atoms + 1,), torch.int64)
        n_events = C.c_int64(0)
        st = lib.gd_events_scan(dv.ptr(work), n_frames, n_atoms, dv.ptr(changed), dv.ptr(offsets),
                                C.byref(n_events), s)
        if st == _lib.GD_ERR_INNER_CONSTANT:
            raise IndexError('index -1 is out of bounds for axis 0 with size 0')
        _lib.check(st)
        launches += 2
        self.last_n_local_events = n_events.value
        if not any_outer_change:
            raise ValueError('need at least one array to concatenate')
        rows = dv.empty((max(n_events.value, 1), 6), torch.int32)
        self.last_events_route = 'none'
        if n_events.value and staged_ok and not remapped:
            _lib.check(lib.gd_events_compact(dv.ptr(stage), cap, n_frames, n_atoms, dv.ptr(work),
                                             dv.ptr(changed), dv.ptr(offsets), self.frame0, dv.ptr(rows), 0, s))
            launches += 1
            self.last_events_route = 'staged'
        elif n_events.value:
            # a segment overflowed its staging rows, or the Q2 remap changed the states after K1:
            # take the rows from the state arrays
            _lib.check(lib.gd_events_fill(dv.ptr(outer), dv.ptr(inner), n_frames, n_atoms, dv.ptr(next_o),
                                          dv.ptr(next_i), dv.ptr(work), dv.ptr(changed), dv.ptr(offsets),
                                          self.frame0, dv.ptr(rows), s))
            launches += 1
            self.last_events_route = 'reread'
        rows = rows[: n_events.value]
        self._mark('k2_events')

        # ---- K3.  Frame shards hold time slices of every atom while the jump machine is
        # sequential along an atom's history.  Its state forgets the past at every departure
        # event, so each shard runs its own slice and only the per-atom machine STATE crosses the
        # shard boundary (40 B per atom, `_jumps_frame_sharded`); the events never move.  Event
        # tables that are not state-consistent fall back to exchanging the rows by atom block.
        res = {'outer': outer, 'inner': inner, 'visited': vis if self._visited_needed else None, 'band': band}
        if self.world == 1:
            jumps, k3_launches = self._jumps_local(rows, offsets, n_atoms)
            res.update(events=rows, atom_offsets=offsets, events_layout='whole', jumps=jumps,
                       n_events=n_events.value)
        else:
            out = None if self.k3_exchange else self._jumps_frame_sharded(rows, offsets, n_atoms)
            if out is not None:
                jumps, n_events_global, k3_launches = out
                res.update(events=rows, atom_offsets=offsets, events_layout='frame_shard', jumps=jumps,
                           n_events=n_events_global)
            else:
                rows, offsets, a_lo, a_hi, n_events_global = self._exchange_events(rows, offsets, n_atoms)
                jumps, k3_launches = self._jumps_local(rows, offsets, a_hi - a_lo)
                jumps = self.comm.all_gather_rows(jumps)  # rank order = atom-block order = (atom, time)
                k3_launches += self.world
                res.update(events=rows, atom_offsets=offsets, events_layout='atom_block',
                           atom_block=(a_lo, a_hi), jumps=jumps, n_events=n_events_global)
        launches += k3_launches
        self._mark('k3_jumps')
        res['n_jumps'] = int(res['jumps'].shape[0])
        self.launches_per_run = launches
        self.last_pass = 'checked'
        if res['events_layout'] != 'atom_block':
            self._cap_events = max(self._cap_events, n_events.value + n_events.value // 8 + 1024)
        if self.world == 1:
            self._jump_cap = max(self._jump_cap, res['n_jumps'] + res['n_jumps'] // 4 + 16)
        return res

    def _enqueue_optimistic(self, pos: torch.Tensor, timed: bool):
        """Everything of the optimistic pass that needs no host decision (capturable in a CUDA
        graph): the whole pass queued on the stream with table sizes remembered from an earlier run:
        K1 -> (flag OR over ranks, on the device) -> K2 scan + compact into a table of
        `_cap_events` rows -> K3 first pass -> record all-gather -> resolve -> K3 final pass ->
        fill into `_jump_cap` rows -> ONE all-gather of offsets + status + rows -> ONE host read.
        Returns None (every rank takes the same decision: all status words are gathered and the
        flag buffer is OR-ed over ranks) when anything needs the checked pass."""
        lib = _lib.lib
        s = dv.stream_ptr()
        W, me = self.world, self.rank
        k1 = self._launch_k1(pos, timed=timed)
        n_frames, n_atoms, launches = k1.n_frames, k1.n_atoms, k1.launches
        assert k1.stage is not None
        if W > 1:
            k1.pack.copy_(self.comm.all_gather(k1.pack).amax(dim=0))
        # ---- K2: nothing read back; the row count stays in offsets[n_atoms]
        offsets = self._tensor('offsets', (n_atoms + 1,), torch.int64)
        _lib.check(lib.gd_events_scan(dv.ptr(k1.work), n_frames, n_atoms, dv.ptr(k1.changed), dv.ptr(offsets),
                                      None, s))
        ecap = self._cap_events
        rows = self._tensor('rows_cap', (ecap, 6), torch.int32)
        _lib.check(lib.gd_events_compact(dv.ptr(k1.stage), k1.cap, n_frames, n_atoms, dv.ptr(k1.work),
                                         dv.ptr(k1.changed), dv.ptr(offsets), self.frame0, dv.ptr(rows), ecap, s))
        launches += 3
        self._mark('k2_events')
        # ---- K3 on the capacity-sized table
        last = 1 if me + 1 == W else 0
        nbytes = _lib.GD_JUMP_STATE_BYTES
        jwork = self._tensor('jwork', (int(lib.gd_jumps_work_bytes(ecap, n_atoms)),), torch.uint8)
        exit_rec = self._tensor('jexit', (n_atoms, nbytes), torch.uint8)
        entry = self._tensor('jentry', (n_atoms, nbytes), torch.uint8)
        args = (dv.ptr(rows), dv.ptr(offsets), ecap, n_atoms, self.n_frames_total, self.minimal_residence)
        head = 2 * (n_atoms + 3)  # int64: jump offsets [A + 1] | status word | K2 error flag
        jcap = self._jump_cap
        jbuf = self._tensor('jbuf_opt', (head + jcap * 6,), torch.int32)
        atom_off = jbuf[:head].view(torch.int64)
        published = None
        if W > 1:
            entry.zero_()
            _lib.check(lib.gd_jumps_count_shard(*args, dv.ptr(entry) if me else None, last, 0, None,
                                                dv.ptr(jwork), dv.ptr(exit_rec), None, None, None, s))
            allrec = self.comm.all_gather(exit_rec)  # [W, A, 40]
            published = allrec[me].contiguous()
            d_unresolved = jwork.data_ptr() + int(lib.gd_jumps_work_unresolved_offset(ecap, n_atoms))
            _lib.check(lib.gd_jumps_resolve(dv.ptr(allrec), W, me, n_atoms, dv.ptr(entry), d_unresolved, None, s))
            launches += 7
        _lib.check(lib.gd_jumps_count_shard(*args, dv.ptr(entry) if W > 1 else None, last, 1 if W > 1 else 0,
                                            dv.ptr(published), dv.ptr(jwork), dv.ptr(exit_rec), dv.ptr(atom_off),
                                            None, None, s))
        jumps = jbuf[head:].view(jcap, 6)
        if jcap:
            _lib.check(lib.gd_jumps_fill_capped(*args, dv.ptr(jwork), dv.ptr(jumps), jcap, s))
        # the INNER_CONSTANT flag of the scan travels with the status words
        n_work = int(lib.gd_events_work_elems(n_frames, n_atoms))
        jbuf[head - 2: head - 1].copy_(k1.work[n_work - 16: n_work - 15])
        launches += 8
        allb = self.comm.all_gather(jbuf) if W > 1 else jbuf.unsqueeze(0)
        heads = allb[:, :head].contiguous().view(torch.int64)  # [W, A + 3]
        merged = None
        if W > 1:
            # per-atom placement of every shard's jump rows, derived inside ONE kernel from offsets
            # the host has not seen yet (rows that would leave a table are skipped)
            merged = self._tensor('jmerged', (max(W * jcap, 1), 6), torch.int32)
            _lib.check(lib.gd_jumps_merge_shards(dv.ptr(allb), int(allb.shape[1]), head, W, n_atoms, jcap,
                                                 dv.ptr(merged), W * jcap, s))
            launches += 1
        # everything the host decides on -- every rank's offsets + status words, and the OR-ed flag
        # buffer -- leaves in ONE copy into a persistent pinned buffer (part of the graph)
        tail = torch.cat([heads.reshape(-1).view(torch.uint8), k1.pack])
        host_tail = self._pinned('host_tail', tail.numel())
        host_tail.copy_(tail, non_blocking=True)
        launches += 1
        from types import SimpleNamespace
        return SimpleNamespace(k1=k1, offsets=offsets, rows=rows, jbuf=jbuf, jumps=jumps, allb=allb, heads=heads,
                               head=head, jcap=jcap, launches=launches, merged=merged, host_tail=host_tail,
                               n_head_bytes=int(heads.numel()) * 8)

    def _run_optimistic(self, pos: torch.Tensor):
        """Optimistic pass = `_enqueue_optimistic` (eagerly the first time, then replayed as ONE
        CUDA graph launch: the ~60 small launches and collectives of a step are launch bound when
        a rank's shard is small) + one host read + the per-atom merge of the gathered jump rows."""
        lib = _lib.lib
        W, me = self.world, self.rank
        if not self.stage_events:
            return None
        key = (pos.data_ptr(), tuple(pos.shape), self._cap_events, self._jump_cap,
               None if self._next_row is None else self._next_row.data_ptr(), self._own_halo)
        graphable = self.use_graph and not self.profile_phases and (W == 1 or self.comm.direct)
        if graphable and self._graph_key == key and self._graph is not None:
            self._graph.replay()
            st = self._graph_state
            self.last_launch_mode = 'graph'
        elif graphable and self._eager_key == key:
            # second pass with the same buffers and sizes: capture it
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                st = self._enqueue_optimistic(pos, timed=False)
            g.replay()
            self._graph, self._graph_key, self._graph_state = g, key, st
            self.last_launch_mode = 'graph'
        else:
            self._graph, self._graph_key, self._graph_state = None, None, None
            st = self._enqueue_optimistic(pos, timed=True)
            self._eager_key = key
            self.last_launch_mode = 'eager'
        s = dv.stream_ptr()
        k1, offsets, rows, jumps, allb, head, jcap = st.k1, st.offsets, st.rows, st.jumps, st.allb, st.head, st.jcap
        n_atoms, launches = k1.n_atoms, st.launches
        torch.cuda.current_stream().synchronize()  # <- the one synchronisation
        raw = st.host_tail.numpy()
        all_off = raw[: st.n_head_bytes].view(np.int64).reshape(W, n_atoms + 3)
        host = raw[st.n_head_bytes:]
        self._mark('k3_jumps')
        status, k2_err = all_off[:, n_atoms + 1], all_off[:, n_atoms + 2]
        n_jumps_rank = all_off[:, n_atoms]
        n_local = int(status[me] >> 8)
        vis = host[: k1.n_vis].astype(bool).reshape(2, self.n_sites)
        why = None  # the first reason to hand the pass to the checked route (same on every rank)
        if (status & 0xFF).any():
            why = f'jump pass status {int(np.bitwise_or.reduce(status & 0xFF))}'
        elif (k2_err & 0xFFFFFFFF).any():
            why = 'outer state changes but inner never does'
        elif int(n_jumps_rank.max()) > jcap:
            why = 'jump table outgrew its remembered size'
        elif host[k1.o_ovf: k1.o_ovf + 4].any():
            why = 'event staging overflow'
        elif not host[k1.o_chg: k1.o_chg + n_atoms].any():
            why = 'no outer state change'
        elif self._needs_remap(host[: k1.n_vis]):
            why = 'Q2 remap'
        self.last_optimistic_reject = why
        if why is not None:
            return None
        for v in vis:
            for label, key in zip(self.grid.group_labels, self.grid.group_keys):
                if not (v.any() if key is None else v[key].any()):
                    radius = self.radius[label]
                    warnings.warn(f'No floating species in range of {label} ({radius=})', stacklevel=3)
        if W > 1:
            jumps_out = st.merged[: int(n_jumps_rank.sum())]
        else:
            jumps_out = jumps[: int(n_jumps_rank[0])]
        self._mark('k3_merge')
        most = int(n_jumps_rank.max())
        if self._caps_are_bounds:
            # sizes are known now: the next pass uses tables that fit (and becomes the graph); the
            # bound-sized tables of this pass live on only as long as the caller keeps its result
            self._cap_events, self._jump_cap, self._caps_are_bounds = 0, 0, False
            for name in ('rows_cap', 'jwork', 'jbuf_opt', 'jmerged'):
                self._buf.pop(name, None)
        self._cap_events = max(self._cap_events, n_local + n_local // 8 + 1024)
        self._jump_cap = max(self._jump_cap, most + most // 4 + 16)
        self.last_n_local_events = n_local
        self.last_events_route = 'staged'
        self.launches_per_run = launches
        self.last_pass = 'optimistic'
        return {'outer': k1.outer, 'inner': k1.inner, 'visited': vis if self._visited_needed else None, 'band': k1.band,
                'events': rows[:n_local], 'atom_offsets': offsets,
                'events_layout': 'frame_shard' if W > 1 else 'whole', 'jumps': jumps_out,
                'n_events': int((status >> 8).sum()), 'n_jumps': int(jumps_out.shape[0])}

    # ------------------------------------------------------------------ pieces of the pass
    def _jumps_local(self, rows, offsets, n_block):
        """K3 on a table holding the whole history of ``n_block`` atoms (count -> allocate -> fill)."""
        lib = _lib.lib
        s = dv.stream_ptr()
        n_local = int(rows.shape[0])
        launches = 0
        jwork = self._tensor('jwork', (int(lib.gd_jumps_work_bytes(n_local, max(n_block, 1))),), torch.uint8)
        n_jumps = C.c_int64(0)
        if n_block > 0:
            _lib.check(lib.gd_jumps_count(dv.ptr(rows), dv.ptr(offsets), n_local, n_block, self.n_frames_total,
                                          self.minimal_residence, dv.ptr(jwork), C.byref(n_jumps), s))
            launches += 4
        jumps = dv.empty((max(n_jumps.value, 1), 6), torch.int32)
        if n_jumps.value:
            _lib.check(lib.gd_jumps_fill(dv.ptr(rows), dv.ptr(offsets), n_local, n_block, self.n_frames_total,
                                         self.minimal_residence, dv.ptr(jwork), dv.ptr(jumps), s))
            launches += 1
        return jumps[: n_jumps.value], launches

    def _jumps_frame_sharded(self, rows, offsets, n_atoms):
        """K3 on this rank's time slice of every atom; only machine states cross the boundary.

        Returns ``(jumps [J, 6] whole table on every rank, global event count, launches)`` or
        ``None`` when a published record turned out not to be the true one (caller falls back to
        the event exchange; all ranks take the same decision)."""
        lib = _lib.lib
        s = dv.stream_ptr()
        W, me = self.world, self.rank
        n_local = int(rows.shape[0])
        last = 1 if me + 1 == W else 0
        nbytes = _lib.GD_JUMP_STATE_BYTES
        jwork = self._tensor('jwork', (int(lib.gd_jumps_work_bytes(n_local, n_atoms)),), torch.uint8)
        exit_rec = self._tensor('jexit', (n_atoms, nbytes), torch.uint8)
        entry = self._tensor('jentry', (n_atoms, nbytes), torch.uint8)
        scratch = self._tensor('jscratch', (4,), torch.int32)
        args = (dv.ptr(rows), dv.ptr(offsets), n_local, n_atoms, self.n_frames_total, self.minimal_residence)
        launches = 0
        # pass 1 (asynchronous): every atom's exit record when the slice is run from the empty state
        # (rank 0 really starts from the empty state; for the others all-zero records say "empty, not known")
        entry.zero_()
        _lib.check(lib.gd_jumps_count_shard(*args, dv.ptr(entry) if me else None, last, 0, None, dv.ptr(jwork),
                                            dv.ptr(exit_rec), None, None, None, s))
        launches += 5
        self._mark('k3a_pass1')
        d_unresolved = jwork.data_ptr() + int(lib.gd_jumps_work_unresolved_offset(n_local, n_atoms))
        # one buffer for the final gather: per-atom jump offsets [A + 1] + (mismatch, event count)
        # word as int64, then room for `cap` jump rows -- cap remembered from earlier runs, so that
        # in steady state offsets and rows travel in ONE all-gather
        head = 2 * (n_atoms + 2)
        cap = self._jump_cap
        jbuf = self._tensor('jbuf', (head + cap * 6,), torch.int32)
        atom_off = jbuf[:head].view(torch.int64)
        n_jumps, flags = C.c_int64(0), C.c_int32(0)
        published = None
        for _ in range(W):
            allrec = self.comm.all_gather(exit_rec)  # [W, A, 40]
            self._mark('k3b_gather_records')
            if published is None:
                published = allrec[me].contiguous()
            # entry records from all shards' exits, then (optimistically, same synchronisation)
            # the final pass from them
            _lib.check(lib.gd_jumps_resolve(dv.ptr(allrec), W, me, n_atoms, dv.ptr(entry), d_unresolved, None, s))
            _lib.check(lib.gd_jumps_count_shard(*args, dv.ptr(entry), last, 1, dv.ptr(published), dv.ptr(jwork),
                                                dv.ptr(exit_rec), dv.ptr(atom_off), C.byref(n_jumps),
                                                C.byref(flags), s))
            launches += 6
            self._mark('k3c_final_count')
            if not flags.value & 4:
                break
            # some shard has events of an atom but no departure among them (or an unproven chain):
            # its exit depends on its entry.  The pass above ran from what is known so far (checked,
            # so its records are right where the entry was); every round settles one more shard.
        else:
            raise RuntimeError('jump machine states did not resolve in world rounds')
        fits = n_jumps.value <= cap
        jumps = jbuf[head:].view(cap, 6) if fits else dv.empty((n_jumps.value, 6), torch.int32)
        if n_jumps.value:
            _lib.check(lib.gd_jumps_fill(*args, dv.ptr(jwork), dv.ptr(jumps), s))
            launches += 1
        atom_off[n_atoms + 1] = (n_local << 8) | (_lib.GD_JUMPS_TAIL_MISMATCH if flags.value & 2 else 0)
        self._mark('k3d_fill')
        allb = self.comm.all_gather(jbuf)  # [W, head + cap * 6]
        all_off = dv.to_host(allb[:, :head].contiguous().view(torch.int64))  # [W, A + 2]
        self._mark('k3e_gather_jumps')
        tail = all_off[:, -1]
        if (tail & 1).any():
            return None
        n_events_global = int((tail >> 8).sum())
        all_off = all_off[:, : n_atoms + 1]
        counts = np.diff(all_off, axis=1)
        most = int(all_off[:, -1].max())
        self._jump_cap = max(self._jump_cap, most + most // 4 + 16)
        if most <= cap:  # every rank's rows came along
            allj = [allb[r, head:] for r in range(W)]
        else:            # first run (or a burst of jumps): second gather with the exact capacity
            padded = self._tensor('jpad', (most, 6), torch.int32)
            padded[: n_jumps.value] = jumps[: n_jumps.value]
            allj = self.comm.all_gather(padded)  # [W, most, 6]
        total_off, dst_start = merge_plan(counts)
        merged = dv.empty((max(int(total_off[-1]), 1), 6), torch.int32)
        d_plan = dv.to_device(np.ascontiguousarray(np.concatenate([all_off, dst_start], axis=1)))  # [W, 2A+1]
        for r in range(W):
            if all_off[r, -1]:
                _lib.check(lib.gd_events_merge(dv.ptr(allj[r]), dv.ptr(d_plan[r]),
                                               dv.ptr(d_plan[r, n_atoms + 1:]), n_atoms, dv.ptr(merged), s))
                launches += 1
        return merged[: int(total_off[-1])], n_events_global, launches

    def atom_block(self, rank: int, n_atoms: int) -> tuple[int, int]:
        return n_atoms * rank // self.world, n_atoms * (rank + 1) // self.world

    def _exchange_events(self, rows, offsets, n_atoms):
        """All-to-all by atom block + per-atom merge of the pieces.

        Every shard's rows are sorted by (atom, time), so the rows for one atom block are one
        contiguous slice.  Returns this rank's block: rows sorted by (atom, time) over ALL frames,
        block-local offsets [n_block + 1], the block bounds and the global event count.
        """
        lib = _lib.lib
        s = dv.stream_ptr()
        W, me = self.world, self.rank
        all_off = dv.to_host(self.comm.all_gather(offsets))  # [W, A + 1] row offsets of every shard
        counts = np.diff(all_off, axis=1)                     # [W, A]
        bounds = [self.atom_block(r, n_atoms) for r in range(W)]
        a_lo, a_hi = bounds[me]
        send = [int(all_off[me, hi] - all_off[me, lo]) for lo, hi in bounds]
        recv = [int(all_off[r, a_hi] - all_off[r, a_lo]) for r in range(W)]
        pieces = self.comm.all_to_all_rows(rows, send, recv)  # list of [recv[r], 6], piece r from shard r
        # place piece r's rows of atom a after the rows of earlier shards of a
        block_counts = counts[:, a_lo:a_hi]
        atom_offsets, dst_start = merge_plan(block_counts)
        n_local = int(atom_offsets[-1])
        merged = dv.empty((max(n_local, 1), 6), torch.int32)
        d_dst = dv.to_device(np.ascontiguousarray(dst_start))
        for r in range(W):
            if recv[r] and a_hi > a_lo:
                src_off = all_off[r, a_lo:a_hi + 1] - all_off[r, a_lo]
                d_src_off = dv.to_device(np.ascontiguousarray(src_off.astype(np.int64)))
                _lib.check(lib.gd_events_merge(dv.ptr(pieces[r]), dv.ptr(d_src_off), dv.ptr(d_dst[r]),
                                               a_hi - a_lo, dv.ptr(merged), s))
        return (merged[:n_local], dv.to_device(atom_offsets.astype(np.int64)), a_lo, a_hi,
                int(all_off[:, -1].sum()))

    def gather_events(self, res: dict) -> torch.Tensor:
        """The whole event table [E, 6] (sorted by atom, time) on every rank."""
        layout = res.get('events_layout', 'whole')
        if self.world == 1 or layout == 'whole':
            return res['events']
        if layout == 'atom_block':
            return self.comm.all_gather_rows(res['events'])
        # frame shards: every shard's rows are sorted by (atom, time); place shard r's rows of atom
        # a after the rows of atoms < a and of earlier shards of a
        lib = _lib.lib
        s = dv.stream_ptr()
        rows, n_atoms = res['events'], self.n_atoms
        all_off = dv.to_host(self.comm.all_gather(res['atom_offsets']))  # [W, A + 1]
        cap = max(1, int(all_off[:, -1].max()))
        padded = torch.zeros((cap, 6), dtype=torch.int32, device=rows.device)
        padded[: rows.shape[0]] = rows
        allr = self.comm.all_gather(padded)
        total_off, dst_start = merge_plan(np.diff(all_off, axis=1))
        merged = dv.empty((max(int(total_off[-1]), 1), 6), torch.int32)
        d_plan = dv.to_device(np.ascontiguousarray(np.concatenate([all_off, dst_start], axis=1)))
        for r in range(self.world):
            if all_off[r, -1]:
                _lib.check(lib.gd_events_merge(dv.ptr(allr[r]), dv.ptr(d_plan[r]),
                                               dv.ptr(d_plan[r, n_atoms + 1:]), n_atoms, dv.ptr(merged), s))
        return merged[: int(total_off[-1])]
