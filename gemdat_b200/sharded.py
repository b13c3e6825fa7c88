"""The rest of the path over contiguous frame shards (one process per GPU).

``pipeline.HotPath`` shards K1 -> K2 -> K3.  This module shards what sits beside them, each with
the one exchange the path really has (``torch.distributed``: NCCL over NVLink on the GPU box,
gloo in the CPU tests of the host logic):

=========  ==========================================================  =====================
stage      what crosses a shard boundary                               collective
=========  ==========================================================  =====================
K0 unwrap  last position row of the previous shard (halo), the first   all-gather [W, 2, A, 3]
           row of shard 0 (base), every shard's winding total          f64 + [W, A, 3] int32
           (integers: the exclusive sum over shards is exact)
ffill      last valid state of the shards before (after, for bfill)    all-gather [W, A] int32
K5         nothing until the end: local uint32 grids                   all-reduce SUM
N1         local site counts                                           all-reduce SUM
K4 MSD     every lag pairs distant frames, so the frame shards are      all-to-all (24 B per
           re-cut by ATOM once; the atom-averaged curve is the sum of   atom-frame, once) +
           per-lag partial sums over the ranks' atom blocks             all-reduce SUM [lags]
N4 RDF     local [class, group, bin] tables, first-occurrence keys      all-reduce SUM / MIN
=========  ==========================================================  =====================

The functions named ``shard_*`` are the host-side rules (pure tensor code, tested on CPU with
gloo); ``FrameShards`` runs the kernels around them.
"""

from __future__ import annotations

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .pipeline import Comm

NOSITE = -1


# ------------------------------------------------------------------ host-side rules


def shard_bounds(n_total: int, world: int) -> list[tuple[int, int]]:
    """Contiguous blocks of ``n_total`` items (frames or atoms), one per rank."""
    return [(n_total * r // world, n_total * (r + 1) // world) for r in range(world)]


def shard_base_winding(all_totals: torch.Tensor, rank: int) -> torch.Tensor:
    """``all_totals`` [W, A, 3] int32 winding totals of every shard -> lattice translations
    accumulated before shard ``rank`` (exclusive sum over the shards: integers, exact)."""
    return all_totals[:rank].sum(dim=0, dtype=torch.int32).contiguous()


def shard_fill_carry(all_last: torch.Tensor, rank: int, direction: int) -> torch.Tensor:
    """``all_last`` [W, A] int32: last valid state of every shard in scan direction (NOSITE if
    none).  Returns the state carried into shard ``rank``: that of the nearest shard before it
    (ffill, direction +1) or after it (bfill, -1) that has one."""
    world = all_last.shape[0]
    carry = torch.full_like(all_last[0], NOSITE)
    order = range(rank) if direction > 0 else range(world - 1, rank, -1)
    for r in order:  # walk towards `rank`: nearer shards overwrite farther ones
        carry = torch.where(all_last[r] != NOSITE, all_last[r], carry)
    return carry.contiguous()


# ------------------------------------------------------------------ the sharded kernels


class FrameShards:
    def __init__(self, *, lattice, n_atoms: int, n_frames_total: int, world: int = 1, rank: int = 0,
                 group=None):
        self.lattice = np.ascontiguousarray(np.asarray(lattice, dtype=np.float64).reshape(3, 3))
        self.n_atoms = int(n_atoms)
        self.n_frames_total = int(n_frames_total)
        self.world, self.rank = int(world), int(rank)
        self.comm = Comm(world, group)
        self.frame_bounds = shard_bounds(self.n_frames_total, self.world)
        self.atom_bounds = shard_bounds(self.n_atoms, self.world)

    # -------------------------------------------------------------- K0
    def unwrap(self, pos: torch.Tensor, *, want_frac=False, want_cart=False, want_dist=False,
               want_speed=False) -> dict:
        """K0 on this rank's frames [n, A, 3]: outputs as ``gd_unwrap`` for the same frames of the
        whole trajectory (bit-identical to one shard; ``dist`` is [A, n], ``speed`` [n, A])."""
        lib, s = _lib.lib, dv.stream_ptr()
        n, A = int(pos.shape[0]), int(pos.shape[1])
        work = dv.empty((int(lib.gd_unwrap_chunks(n, A)) * A * 3,), torch.int32)
        total = dv.empty((A, 3), torch.int32)
        frac = dv.empty((n, A, 3), torch.float64) if want_frac else None
        cart = dv.empty((n, A, 3), torch.float64) if want_cart else None
        dist = dv.empty((A, n), torch.float64) if want_dist else None
        speed = dv.empty((n, A), torch.float64) if want_speed else None
        if self.world == 1:  # one shard: the fused single-read kernel
            _lib.check(lib.gd_unwrap(dv.ptr(pos), n, A, self.lattice.ctypes.data, None, None, None,
                                     dv.ptr(work), dv.ptr(frac), dv.ptr(cart), dv.ptr(dist), dv.ptr(speed),
                                     dv.ptr(total), s))
            return {'frac': frac, 'cart': cart, 'dist': dist, 'speed': speed, 'winding': total}
        prev = base_pos = base_wind = None
        if self.world > 1:
            ends = self.comm.all_gather(torch.stack([pos[0], pos[-1]]).contiguous())  # [W, 2, A, 3]
            base_pos = ends[0, 0].contiguous()
            if self.rank:
                prev = ends[self.rank - 1, 1].contiguous()
        _lib.check(lib.gd_unwrap_windings(dv.ptr(pos), n, A, dv.ptr(prev), dv.ptr(work), dv.ptr(total), s))
        all_totals = self.comm.all_gather(total)  # [W, A, 3]
        if self.world > 1:
            base_wind = shard_base_winding(all_totals, self.rank)
        _lib.check(lib.gd_unwrap_apply(dv.ptr(pos), n, A, self.lattice.ctypes.data, dv.ptr(prev),
                                       dv.ptr(base_wind), dv.ptr(base_pos), dv.ptr(work), dv.ptr(frac),
                                       dv.ptr(cart), dv.ptr(dist), dv.ptr(speed), s))
        return {'frac': frac, 'cart': cart, 'dist': dist, 'speed': speed,
                'winding': all_totals.sum(dim=0, dtype=torch.int32)}

    # -------------------------------------------------------------- drift / centre of mass
    def drift_correction(self, pos: torch.Tensor, selected: np.ndarray | None = None) -> torch.Tensor:
        """Drift-corrected, wrapped positions of this rank's frames [n, A, 3]
        (``Trajectory.apply_drift_correction``): the local drift needs the row before the shard
        (halo), its running sum the totals of the shards before (all-gather of [W, 3], added in
        shard order), K0 the usual winding exchange.  Equal to one shard up to the rounding of
        the summed drift (the sums are grouped by shard)."""
        lib, s = _lib.lib, dv.stream_ptr()
        n, A = int(pos.shape[0]), int(pos.shape[1])
        prev, base_pos = None, pos[0].contiguous()
        if self.world > 1:
            ends = self.comm.all_gather(torch.stack([pos[0], pos[-1]]).contiguous())  # [W, 2, A, 3]
            base_pos = ends[0, 0].contiguous()
            if self.rank:
                prev = ends[self.rank - 1, 1].contiguous()
        d_sel = None if selected is None else dv.to_device(np.ascontiguousarray(selected, dtype=np.int32))
        n_sel = A if selected is None else int(len(selected))
        drift = dv.empty((n, 3), torch.float64)
        _lib.check(lib.gd_drift(dv.ptr(pos), n, A, dv.ptr(d_sel), n_sel, dv.ptr(prev), dv.ptr(drift), s))
        summed = dv.empty((n, 3), torch.float64)
        scan_work = dv.empty((int(lib.gd_cumsum_frames3_work_bytes(n)),), torch.uint8)
        _lib.check(lib.gd_cumsum_frames3(dv.ptr(drift), n, dv.ptr(scan_work), dv.ptr(summed), s))
        if self.world > 1:
            totals = self.comm.all_gather(summed[-1].contiguous())  # [W, 3]
            for r in range(self.rank):  # shard order: the same grouping on every rank
                summed += totals[r]
        cum = self.unwrap(pos, want_frac=True)['frac']
        _lib.check(lib.gd_drift_correct(dv.ptr(cum), dv.ptr(base_pos), dv.ptr(summed), n, A, dv.ptr(cum), s))
        return cum

    def center_of_mass(self, pos: torch.Tensor, weights: np.ndarray | None = None) -> torch.Tensor:
        """[n, 3] weighted mean of the unwrapped positions of this rank's frames
        (``Trajectory.center_of_mass``): K0 + one reduction per frame, no further exchange."""
        n, A = int(pos.shape[0]), int(pos.shape[1])
        base_pos = pos[0].contiguous()
        if self.world > 1:
            base_pos = self.comm.all_gather(base_pos)[0].contiguous()
        cum = self.unwrap(pos, want_frac=True)['frac']
        d_w = None if weights is None else dv.to_device(np.ascontiguousarray(weights, dtype=np.float64))
        out = dv.empty((n, 3), torch.float64)
        _lib.check(_lib.lib.gd_frame_mean(dv.ptr(cum), dv.ptr(base_pos), dv.ptr(d_w), n, A, dv.ptr(out),
                                          dv.stream_ptr()))
        return out

    # -------------------------------------------------------------- ffill / bfill
    def fill(self, states: torch.Tensor, direction: int) -> torch.Tensor:
        """states_prev (direction +1) / states_next (-1) of this rank's frames [n, A] int32."""
        lib, s = _lib.lib, dv.stream_ptr()
        n, A = int(states.shape[0]), int(states.shape[1])
        work = dv.empty((int(lib.gd_events_work_elems(n, A)),), torch.int32)
        last = dv.empty((A,), torch.int32)
        _lib.check(lib.gd_fill_last_valid(dv.ptr(states), n, A, direction, dv.ptr(work), dv.ptr(last), s))
        carry = None
        if self.world > 1:
            carry = shard_fill_carry(self.comm.all_gather(last), self.rank, direction)
        out = dv.empty((n, A), torch.int32)
        _lib.check(lib.gd_fill_apply(dv.ptr(states), n, A, direction, dv.ptr(carry), dv.ptr(work),
                                     dv.ptr(out), s))
        return out

    # -------------------------------------------------------------- K5 / N1
    def density(self, pos: torch.Tensor, nedges: np.ndarray) -> torch.Tensor:
        """Voxel counts of ALL frames on every rank: local uint32 grid + all-reduce(SUM)."""
        from .volume import density_device

        grid, flag = density_device(pos, nedges)
        self.comm.all_reduce_max(flag)
        assert int(flag.item()) == 0  # volume.py:476-477
        return self.comm.all_reduce_sum(grid)

    def site_counts(self, states: torch.Tensor, n_sites: int) -> torch.Tensor:
        """[n_sites + 1] int64 entries per site ([0] = NOSITE) over all frames, on every rank."""
        counts = dv.zeros((n_sites + 1,), torch.int64)
        _lib.check(_lib.lib.gd_occupancy(dv.ptr(states), states.numel(), n_sites, dv.ptr(counts),
                                         dv.stream_ptr()))
        return self.comm.all_reduce_sum(counts)

    # -------------------------------------------------------------- K4
    def atoms_of_rank(self) -> tuple[int, int]:
        return self.atom_bounds[self.rank]

    def regroup_by_atom(self, cart: torch.Tensor) -> torch.Tensor:
        """Frame shard [n, A, 3] -> this rank's ATOM block over all frames, component-major
        [A_r, 3, T] (the layout ``gd_msd_direct`` streams): one all-to-all, 24 B per atom-frame."""
        lib, s = _lib.lib, dv.stream_ptr()
        n, A = int(cart.shape[0]), int(cart.shape[1])
        soa = dv.empty((A, 3, n), torch.float64)
        _lib.check(lib.gd_transpose_soa(dv.ptr(cart), n, A, dv.ptr(soa), s))
        return self.exchange_atom_blocks(soa)

    def exchange_atom_blocks(self, soa: torch.Tensor) -> torch.Tensor:
        """[A, 3, n] (all atoms, this rank's frames) -> [A_r, 3, T] (this rank's atoms, all
        frames).  An atom block of a frame shard is contiguous in the component-major layout, so
        the pieces go out as they lie; the received pieces are time blocks of the result."""
        if self.world == 1:
            return soa
        n = int(soa.shape[2])
        a_lo, a_hi = self.atoms_of_rank()
        send = [(hi - lo) * 3 * n for lo, hi in self.atom_bounds]
        recv = [(a_hi - a_lo) * 3 * (f1 - f0) for f0, f1 in self.frame_bounds]
        pieces = self.comm.all_to_all_flat(soa.reshape(-1), send, recv)
        out = soa.new_empty((a_hi - a_lo, 3, self.n_frames_total))
        for (f0, f1), piece in zip(self.frame_bounds, pieces):
            out[:, :, f0:f1] = piece.view(a_hi - a_lo, 3, f1 - f0)
        return out

    def msd(self, pos: torch.Tensor, lag0: int = 0, n_lags: int | None = None):
        """All-lag MSD from frame-sharded wrapped positions.  Returns ``(msd, (a_lo, a_hi),
        curve)``: ``msd`` [A_r, n_lags] rows of this rank's atom block (the reference's [A, T]
        array is their concatenation in rank order), ``curve`` [n_lags] the atom-averaged MSD on
        every rank (all-reduce of the per-lag partial sums)."""
        lib, s = _lib.lib, dv.stream_ptr()
        cart = self.unwrap(pos, want_cart=True)['cart']
        soa = self.regroup_by_atom(cart)
        del cart
        T = self.n_frames_total
        n_lags = T - lag0 if n_lags is None else int(n_lags)
        a_lo, a_hi = self.atoms_of_rank()
        msd = dv.empty((a_hi - a_lo, n_lags), torch.float64)
        if a_hi > a_lo:
            _lib.check(lib.gd_msd_direct(dv.ptr(soa), T, a_hi - a_lo, lag0, n_lags, dv.ptr(msd), s))
        curve = msd.sum(dim=0)
        self.comm.all_reduce_sum(curve)
        return msd, (a_lo, a_hi), curve / self.n_atoms

    # -------------------------------------------------------------- N4
    def species_rdf_counts(self, pos: torch.Tensor, idx1, idx2, bins: np.ndarray, resolution: float,
                           same: bool) -> torch.Tensor:
        """np.histogram counts of rdf.py:232-262 summed over ALL frames, [len(bins) + 1] int64 on
        every rank (bin b of np.histogram is entry b + 1)."""
        from .rdf import _hist_device

        hist = _hist_device(pos, self.lattice, idx1, None, 1, idx2, None, 1, bins, resolution, 1, same)
        return self.comm.all_reduce_sum(hist)[0, 0]

    def reduce_rdf(self, hist: torch.Tensor, first_key: torch.Tensor | None = None, frame0: int = 0):
        """Local [class, group, bin] uint64 tables -> totals on every rank; the first-occurrence
        keys ((frame << 32) | code, frames local to the shard) get the shard's first frame added
        before the all-reduce(MIN), so the state order is that of the whole trajectory."""
        self.comm.all_reduce_sum(hist)
        if first_key is not None:
            if frame0:
                never = first_key == torch.iinfo(torch.int64).max
                first_key += (frame0 << 32)
                first_key[never] = torch.iinfo(torch.int64).max
            self.comm.all_reduce_min(first_key)
        return hist, first_key
