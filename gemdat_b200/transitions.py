"""Transitions between sites: host-side mirror of reference src/gemdat/transitions.py.

Same names, kwargs, return types, warnings and exceptions as the reference for the
hot path: ``Transitions.from_trajectory`` (:96-186), ``Transitions.jumps`` (:188-202),
``matrix`` / ``states_next`` / ``states_prev`` (:243-278), ``occupancy*`` /
``atom_locations`` / ``residence_time`` / ``split`` (:280-410), ``SiteRadius``
(:485-733) and the three private seams the CUDA kernels sit behind:

* ``_calculate_atom_states``        (:736-809)  -> ``gd_sitegrid_build`` + ``gd_site_states`` (K1)
* ``_calculate_transition_events``  (:413-482)  -> ``gd_events_count/scan/fill`` (K2)
* ``_calculate_transitions_matrix`` (:812-834)  -> ``gd_pair_matrix``

All arithmetic on the path runs in ``libgemdat_b200.so``; this file only prepares
small host-side tables, maps status codes to the reference's exceptions and wraps
results in numpy / pandas.
"""

from __future__ import annotations

import ctypes as C
import os
import typing
from dataclasses import dataclass
from itertools import combinations_with_replacement
from warnings import warn

import numpy as np
import pandas as pd
import torch

from . import _device as dv
from . import _lib
from .caching import weak_lru_cache
from .lattice import Sites, site_arrays

if typing.TYPE_CHECKING:
    from .jumps import Jumps
    from .trajectory import Trajectory

NOSITE = -1

EVENT_COLUMNS = [
    'atom index',
    'start site',
    'destination site',
    'start inner site',
    'destination inner site',
    'time',
]


# --------------------------------------------------------------------------- K1 host side


class SiteGrid:
    """Host + device copy of the search structure built by ``gd_sitegrid_build``."""

    _blob_guess = 8 << 20  # bytes tried first (C3's 2160 sites need 2 MB); grows when it was too small

    def __init__(self, *, lattice_matrix, site_frac, site_labels, site_radius: dict,
                 site_inner_fraction: dict):
        labels = np.asarray(list(site_labels), dtype=object)
        self.group_labels = list(site_radius.keys())
        self.n_sites = len(site_frac)
        group = np.full(self.n_sites, -1, dtype=np.int32)
        self.group_keys: list[np.ndarray | None] = []
        for g, label in enumerate(self.group_labels):
            if label:
                key = np.nonzero(labels == label)[0]
                if key.size == 0:
                    # the reference does `key, site_group = zip(*empty)` -> ValueError
                    raise ValueError('not enough values to unpack (expected 2, got 0)')
                group[key] = g
                self.group_keys.append(key)
            else:
                group[:] = g
                self.group_keys.append(None)
        radius = np.array([site_radius[k] for k in self.group_labels], dtype=np.float64)
        inner = np.array([site_inner_fraction[k] for k in self.group_labels], dtype=np.float64)
        lat = dv.host_f64(lattice_matrix)
        frac = dv.host_f64(site_frac)
        needed = C.c_size_t(0)
        args = (lat.ctypes.data, frac.ctypes.data, group.ctypes.data, self.n_sites,
                radius.ctypes.data, inner.ctypes.data, len(radius))
        # one pass in the common case: the size query of gd_sitegrid_build does the whole
        # construction, so a buffer that is large enough for most site sets is tried first (its
        # untouched pages cost nothing) and the exact-size query only runs when it was too small
        guess = np.zeros(SiteGrid._blob_guess, dtype=np.uint8)
        st = _lib.lib.gd_sitegrid_build(*args, guess.ctypes.data, guess.nbytes, C.byref(needed))
        if st == _lib.GD_OK:
            self.h_blob = guess[:needed.value].copy()
        elif st == _lib.GD_ERR_CAPACITY:
            SiteGrid._blob_guess = max(SiteGrid._blob_guess, 2 * int(needed.value))
            self.h_blob = np.zeros(needed.value, dtype=np.uint8)
            st = _lib.lib.gd_sitegrid_build(*args, self.h_blob.ctypes.data, self.h_blob.nbytes,
                                            C.byref(needed))
        if st == _lib.GD_ERR_RADIUS_GT_CUTOFF:
            raise RuntimeError('Set cutoff greater or equal to the radius.')
        if st == _lib.GD_ERR_TRICLINIC:
            raise NotImplementedError(_lib.last_error())
        _lib.check(st)
        self.radius = radius
        self.group = group
        self._d_blob = None

    @property
    def d_blob(self) -> torch.Tensor:
        if self._d_blob is None:
            self._d_blob = dv.to_device(self.h_blob)
        return self._d_blob

    def upload_async(self) -> torch.Tensor:
        """Device copy of the table queued on the current stream from a page-locked staging copy,
        without blocking the host (the upload pipeline has part copies in flight at that point)."""
        if self._d_blob is None:
            self._pinned_blob = torch.empty((self.h_blob.nbytes,), dtype=torch.uint8, pin_memory=True)
            self._pinned_blob.numpy()[:] = self.h_blob
            d = dv.empty((self.h_blob.nbytes,), torch.uint8)
            d.copy_(self._pinned_blob, non_blocking=True)
            self._d_blob = d
        return self._d_blob

    def remap_lut(self, visited: np.ndarray) -> np.ndarray | None:
        """LUT natural index -> reported index for the integer_remap quirk, None if identity.

        transitions.py:804-805 maps a group's local site number through
        ``integer_remap(siteno, key, palette=np.unique(siteno))``: the palette holds only
        the sites that were hit, so the i-th *visited* site of a labelled group is
        reported as ``key[i]``.
        """
        lut = np.arange(self.n_sites, dtype=np.int32)
        for key in self.group_keys:
            if key is None:
                continue
            v = visited[key]
            rank = np.cumsum(v) - 1
            lut[key[v]] = key[rank[v]]
        return None if np.array_equal(lut, np.arange(self.n_sites)) else lut


def _site_states_device(*, sites, trajectory, site_radius: dict, site_inner_fraction: dict,
                        which: str = 'both', count_events: bool = False):
    """K1 on the device.  Returns (outer int32 [T,A], inner int32 [T,A], info dict).

    With ``count_events`` the same pass also leaves the per-(atom, frame chunk) event counts
    and the "state ever changes" flags K2 needs (``info['event_work']``, ``info['changed']``).
    """
    site_frac, site_labels, _ = site_arrays(sites)

    def make_grid():
        return SiteGrid(
            lattice_matrix=trajectory.get_lattice().matrix,
            site_frac=site_frac,
            site_labels=site_labels,
            site_radius=site_radius,
            site_inner_fraction=site_inner_fraction,
        )

    host_future = None
    resident = True
    pipelined = _wants_pipeline(trajectory, count_events)
    grid = None if pipelined else make_grid()
    if pipelined:
        # host-resident input: upload, K1 and the download of the int64 states run as a
        # three-stage pipeline over frame parts (PCIe is full duplex); events are counted by the
        # stand-alone pass afterwards (it re-reads 8 B/atom-frame from HBM, noise next to PCIe)
        pos, outer, inner, visited, band, host_future, outside, grid = _pipelined_states(trajectory, make_grid)
        n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
        work = changed = None
        # some coordinate was outside [0, 1): the device copy was wrapped afterwards (np.mod of
        # to_positions) and K1 runs again over it (rare: inputs are normally wrapped)
        resident = outside
        if outside:
            host_future = None
        else:
            trajectory._unit_wrapped = True
    if resident:
        pos = trajectory._device_positions()
        n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
        outer = dv.empty((n_frames, n_atoms), torch.int32)
        inner = dv.empty((n_frames, n_atoms), torch.int32)
        visited = dv.zeros((2 * grid.n_sites,), torch.uint8)
        band = dv.zeros((4,), torch.int64)
        work = changed = None
        if count_events and n_frames > 0:
            work = dv.empty((int(_lib.lib.gd_events_work_elems(n_frames, n_atoms)),), torch.int32)
            changed = dv.zeros((2 * n_atoms,), torch.uint8)
        st = _lib.lib.gd_site_states(
            dv.ptr(pos), n_frames, n_atoms, None, dv.ptr(grid.d_blob), grid.h_blob.ctypes.data,
            dv.ptr(outer), dv.ptr(inner), dv.ptr(visited), dv.ptr(band), dv.ptr(work), dv.ptr(changed),
            dv.stream_ptr(),
        )
        _lib.check(st)
    visited_h = dv.to_host(visited).astype(bool).reshape(2, grid.n_sites)
    band_h = dv.to_host(band)
    for which_i, (arr, vis) in enumerate(((outer, visited_h[0]), (inner, visited_h[1]))):
        if which == 'outer' and which_i == 1 or which == 'inner' and which_i == 0:
            continue
        for label, key in zip(grid.group_labels, grid.group_keys):
            hit = vis.any() if key is None else vis[key].any()
            if not hit:
                radius = site_radius[label]
                warn(f'No floating species in range of {label} ({radius=})', stacklevel=3)
        lut = grid.remap_lut(vis)
        if lut is not None:
            # injective on the visited sites, so the event pattern counted by K1 is unchanged
            d_lut = dv.to_device(lut)
            _lib.check(_lib.lib.gd_remap_states(dv.ptr(arr), arr.numel(), dv.ptr(d_lut),
                                                grid.n_sites, dv.stream_ptr()))
            host_future = None  # the prefetched host copy predates the remap: fetch again lazily
    info = {
        'visited_outer': visited_h[0],
        'visited_inner': visited_h[1],
        'band_1e-12': (int(band_h[0]), int(band_h[1])),
        'band_4e-6': (int(band_h[2]), int(band_h[3])),
        'event_work': work,
        'changed': changed,
        'host_future': host_future,
    }
    return outer, inner, info


PIPELINE_MIN_BYTES = 64 << 20
# parts whose upload is queued before the site grid is built (and kept in front of the compute).
# Measured on B200 / PCIe 5 (profiles/r2zl_e2e_ahead.txt): 1 / 2 / 4 / 6 ahead = 56.1 / 55.2 / 52.8 /
# 51.1 ms per 1e5 C3 frames with the 7 ms grid build of 2160 sites, but with a fast build (200 sites)
# 50.2 / 48.8 / 73.4 / 73.8 ms: once several uploads are still queued when the first downloads are
# issued, the two directions stop overlapping (48 + 25 ms).  2 is safe in both regimes.
PIPELINE_AHEAD = 2
PIPELINE_PARTS = 32  # 16: 59.2 ms, 32: 57.6, 48: 57.0 per 1e5 C3 frames (profiles/r2zh_e2e_parts.txt)


def _wants_pipeline(trajectory, count_events: bool) -> bool:
    coords = trajectory.coords
    if dv.is_device(coords) or not count_events or trajectory.coords_are_displacement:
        return False
    holder = trajectory.__dict__.get('_dev_holder') or {}
    cached = holder.get('pos')
    if cached is not None and cached[0] is coords:
        return False  # already on the device
    return coords.nbytes >= PIPELINE_MIN_BYTES and coords.flags.c_contiguous


_STAGING: dict = {}
_COPY_POOL = None
COPY_THREADS = 8
COPY_MIN_BYTES = 4 << 20  # per thread


def _staging_buffers(n_elems: int):
    """Two page-locked float64 buffers of at least ``n_elems``, kept for the process (page
    locking costs more than the copy it serves)."""
    have = _STAGING.get('buf')
    if have is None or have[0].numel() < n_elems:
        have = [torch.empty((n_elems,), dtype=torch.float64, pin_memory=True) for _ in range(2)]
        _STAGING['buf'] = have
    return have


def _threaded_copy(dst: np.ndarray, src: np.ndarray) -> None:
    """dst[...] = src over the leading axis with several host threads (numpy releases the GIL
    inside the copy loop; one thread does not fill the host memory bandwidth)."""
    global _COPY_POOL
    n = len(src)
    n_threads = min(COPY_THREADS, os.cpu_count() or 1, max(1, src.nbytes // max(1, COPY_MIN_BYTES)))
    if n_threads <= 1:
        np.copyto(dst, src)
        return
    if _COPY_POOL is None:
        from concurrent.futures import ThreadPoolExecutor

        _COPY_POOL = ThreadPoolExecutor(max_workers=COPY_THREADS, thread_name_prefix='gd-stage')
    cuts = [n * i // n_threads for i in range(n_threads + 1)]
    jobs = [_COPY_POOL.submit(np.copyto, dst[a:b], src[a:b]) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]
    for j in jobs:
        j.result()


class HostStates:
    """int64 host copies of the two state arrays, arriving through a side stream."""

    def __init__(self, outer: torch.Tensor, inner: torch.Tensor, done: torch.cuda.Event):
        self._outer, self._inner, self._done = outer, inner, done

    def outer(self) -> np.ndarray:
        self._done.synchronize()
        return self._outer.numpy()

    def inner(self) -> np.ndarray:
        self._done.synchronize()
        return self._inner.numpy()


def _pipelined_states(trajectory, grid, trace: list | None = None):
    """K1 over frame parts with upload (copy stream), compute (current stream) and the
    download of the widened states (second copy stream) overlapped.

    ``grid`` is a ``SiteGrid`` or a callable that builds one: the build runs on the host
    (~7 ms for the 2160 sites of C3) and is started AFTER the first uploads are queued, so it hides
    behind the copies instead of delaying them."""
    lib = _lib.lib
    dev = dv.require_cuda()
    coords = trajectory.coords
    n_frames, n_atoms = int(coords.shape[0]), int(coords.shape[1])
    src = torch.from_numpy(coords)
    pos = dv.empty((n_frames, n_atoms, 3), torch.float64)
    outer = dv.empty((n_frames, n_atoms), torch.int32)
    inner = dv.empty((n_frames, n_atoms), torch.int32)
    band = dv.zeros((4,), torch.int64)
    outside = dv.zeros((1,), torch.int32)  # set when a coordinate lies outside [0, 1)
    host_o = torch.empty((n_frames, n_atoms), dtype=torch.int64, pin_memory=True)
    host_i = torch.empty((n_frames, n_atoms), dtype=torch.int64, pin_memory=True)
    n_parts = min(PIPELINE_PARTS, n_frames)
    # frame boundaries in multiples of 4 keep every part 16-byte aligned whatever n_atoms is
    bounds = sorted({min(n_frames, (n_frames * k // n_parts + 3) & ~3) for k in range(n_parts)} | {0, n_frames})
    n_parts = len(bounds) - 1
    part_max = max(b - a for a, b in zip(bounds[:-1], bounds[1:]))
    stage = [(dv.empty((part_max, n_atoms), torch.int64), dv.empty((part_max, n_atoms), torch.int64))
             for _ in range(2)]
    main = torch.cuda.current_stream()
    up, down = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    up.wait_stream(main)
    down.wait_stream(main)
    uploaded = [torch.cuda.Event(enable_timing=trace is not None) for _ in range(n_parts)]
    staged_free = [None, None]

    # pageable source: the driver would bounce every part through its own small staging buffer
    # on one host thread; copy the part into one of two page-locked buffers with several host
    # threads instead, then let the copy engine take it from there
    pageable = not src.is_pinned()
    bounce = _staging_buffers(part_max * n_atoms * 3) if pageable else None
    bounce_free = [None, None]

    def upload(k):
        f0, f1 = bounds[k], bounds[k + 1]
        part = src[f0:f1]
        if pageable:
            if bounce_free[k % 2] is not None:
                bounce_free[k % 2].synchronize()
            part = bounce[k % 2][: (f1 - f0) * n_atoms * 3].view(f1 - f0, n_atoms, 3)
            _threaded_copy(part.numpy(), coords[f0:f1])
        with torch.cuda.stream(up):
            pos[f0:f1].copy_(part, non_blocking=True)
            uploaded[k].record(up)
            if pageable:
                bounce_free[k % 2] = uploaded[k]

    # page-locked source: uploads are queued `ahead` parts in front of the compute and the
    # host-side build of the site grid runs while they are in flight; its small table then queues
    # behind those parts only.  Pageable source: every upload costs host time, one part ahead.
    ahead = 1 if pageable or not callable(grid) else max(1, min(n_parts, PIPELINE_AHEAD))
    for k in range(ahead):
        upload(k)
    if callable(grid):
        try:
            grid = grid()
        except Exception:
            up.synchronize()  # queued copies still write into `pos`: let them finish before it is freed
            raise
    d_blob = grid.upload_async()
    visited = dv.zeros((2 * grid.n_sites,), torch.uint8)
    done = torch.cuda.Event()
    for k in range(n_parts):
        if k + ahead < n_parts and not pageable:
            upload(k + ahead)
        f0, f1 = bounds[k], bounds[k + 1]
        main.wait_event(uploaded[k])
        _lib.check(lib.gd_wrap_unit(dv.ptr(pos[f0:f1]), None, (f1 - f0) * n_atoms * 3, dv.ptr(outside),
                                    main.cuda_stream))
        _lib.check(lib.gd_site_states(
            dv.ptr(pos[f0:f1]), f1 - f0, n_atoms, None, dv.ptr(d_blob), grid.h_blob.ctypes.data,
            dv.ptr(outer[f0:f1]), dv.ptr(inner[f0:f1]), dv.ptr(visited), dv.ptr(band), None, None,
            main.cuda_stream))
        buf = stage[k % 2]
        if staged_free[k % 2] is not None:
            main.wait_event(staged_free[k % 2])
        n = (f1 - f0) * n_atoms
        _lib.check(lib.gd_widen_i32_i64(dv.ptr(outer[f0:f1]), dv.ptr(buf[0]), n, main.cuda_stream))
        _lib.check(lib.gd_widen_i32_i64(dv.ptr(inner[f0:f1]), dv.ptr(buf[1]), n, main.cuda_stream))
        ready = torch.cuda.Event(enable_timing=trace is not None)
        ready.record(main)
        with torch.cuda.stream(down):
            down.wait_event(ready)
            host_o[f0:f1].copy_(buf[0][: f1 - f0], non_blocking=True)
            host_i[f0:f1].copy_(buf[1][: f1 - f0], non_blocking=True)
            free = torch.cuda.Event(enable_timing=trace is not None)
            free.record(down)
            staged_free[k % 2] = free
        if trace is not None:
            trace.append((uploaded[k], ready, free))
        if k + 1 < n_parts and pageable:
            upload(k + 1)  # the host-side copy runs while the device works on part k
    done.record(down)
    # later work on the current stream (and the allocator's reuse of the staging buffers) is
    # ordered after both side streams
    main.wait_stream(up)
    main.wait_stream(down)
    was_outside = bool(int(outside.item()))
    if was_outside:
        _lib.check(lib.gd_wrap_unit(dv.ptr(pos), dv.ptr(pos), pos.numel(), dv.ptr(outside), main.cuda_stream))
    holder = trajectory.__dict__.setdefault('_dev_holder', {})
    holder['pos'] = (coords, pos)
    return pos, outer, inner, visited, band, HostStates(host_o, host_i, done), was_outside, grid


def widen_to_host(t: torch.Tensor) -> np.ndarray:
    """int32 device array -> int64 numpy (the reference's dtype), widened on the device."""
    if t.numel() == 0:
        return np.zeros(tuple(t.shape), dtype=np.int64)
    out = dv.empty(tuple(t.shape), torch.int64)
    _lib.check(_lib.lib.gd_widen_i32_i64(dv.ptr(t), dv.ptr(out), t.numel(), dv.stream_ptr()))
    return dv.to_host(out)


def _calculate_atom_states(
    sites,
    trajectory: 'Trajectory',
    site_radius: dict[str, float],
    site_inner_fraction: dict[str, float],
) -> np.ndarray:
    """Drop-in for transitions.py:736-809: int64 [time, atom] site index, -1 = no site."""
    _, inner, _ = _site_states_device(
        sites=sites, trajectory=trajectory, site_radius=site_radius,
        site_inner_fraction=site_inner_fraction, which='inner',
    )
    return widen_to_host(inner)


# --------------------------------------------------------------------------- K2 host side


class EventTable:
    """Device-resident event table: int32 [E, 6] rows sorted by (atom, time) + per-atom offsets."""

    def __init__(self, rows: torch.Tensor, atom_offsets: torch.Tensor, n_atoms: int):
        self.rows = rows
        self.atom_offsets = atom_offsets
        self.n_atoms = n_atoms

    def __len__(self):
        return int(self.rows.shape[0])

    def to_dataframe(self, index: torch.Tensor | None = None) -> pd.DataFrame:
        # widened to the reference's int64 on the device, one DMA to pinned memory, no host copy
        labels = None if index is None else dv.to_host(index)
        return pd.DataFrame(data=widen_to_host(self.rows), columns=EVENT_COLUMNS, index=labels, copy=False)

    @classmethod
    def from_dataframe(cls, events: pd.DataFrame, n_atoms: int | None = None, return_order: bool = False):
        data = np.ascontiguousarray(events[EVENT_COLUMNS].to_numpy(dtype=np.int64))
        order = np.argsort(data[:, 0], kind='stable')
        data = data[order]
        if n_atoms is None:
            n_atoms = int(data[:, 0].max()) + 1 if len(data) else 1
        counts = np.bincount(data[:, 0], minlength=n_atoms)
        offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        rows = dv.to_device(data.astype(np.int32)) if len(data) else dv.empty((0, 6), torch.int32)
        table = cls(rows, dv.to_device(offsets), n_atoms)
        return (table, order) if return_order else table


def _events_device(outer: torch.Tensor, inner: torch.Tensor, work: torch.Tensor | None = None,
                   changed: torch.Tensor | None = None) -> EventTable:
    """K2: scan + fill; the count pass is skipped when K1 already produced ``work`` / ``changed``."""
    n_frames, n_atoms = int(outer.shape[0]), int(outer.shape[1])
    lib = _lib.lib
    s = dv.stream_ptr()
    if work is None or changed is None:
        work = dv.empty((int(lib.gd_events_work_elems(n_frames, n_atoms)),), torch.int32)
        changed = dv.zeros((2 * n_atoms,), torch.uint8)
        _lib.check(lib.gd_events_count(dv.ptr(outer), dv.ptr(inner), n_frames, n_atoms, None, None,
                                       dv.ptr(work), dv.ptr(changed), s))
    offsets = dv.empty((n_atoms + 1,), torch.int64)
    n_events = C.c_int64(0)
    st = lib.gd_events_scan(dv.ptr(work), n_frames, n_atoms, dv.ptr(changed), dv.ptr(offsets),
                            C.byref(n_events), s)
    if st == _lib.GD_ERR_INNER_CONSTANT:
        raise IndexError('index -1 is out of bounds for axis 0 with size 0')
    _lib.check(st)
    if not bool(changed[:n_atoms].any()):
        # np.vstack([]) in the reference (transitions.py:470)
        raise ValueError('need at least one array to concatenate')
    rows = dv.empty((n_events.value, 6), torch.int32)
    if n_events.value:
        _lib.check(lib.gd_events_fill(dv.ptr(outer), dv.ptr(inner), n_frames, n_atoms, None, None,
                                      dv.ptr(work), dv.ptr(changed), dv.ptr(offsets), 0,
                                      dv.ptr(rows), s))
    return EventTable(rows, offsets, n_atoms)


def _calculate_transition_events(*, atom_sites, atom_inner_sites) -> pd.DataFrame:
    """Drop-in for transitions.py:413-482 (accepts numpy or CUDA int arrays)."""
    outer = dv.to_device(atom_sites, torch.int32)
    inner = dv.to_device(atom_inner_sites, torch.int32)
    return _events_device(outer, inner).to_dataframe()


def _calculate_transitions_matrix(events, n_sites: int) -> np.ndarray:
    """Drop-in for transitions.py:812-834 (``events`` = DataFrame with start/destination site)."""
    pairs = np.ascontiguousarray(events[['start site', 'destination site']].to_numpy(dtype=np.int32))
    return _pair_matrix_device(dv.to_device(pairs) if len(pairs) else None, len(pairs), 2, 0, 1, n_sites)


def _pair_matrix_device(rows, n_rows: int, stride: int, cs: int, cd: int, n_sites: int) -> np.ndarray:
    scratch = dv.zeros(((n_sites + 1) * (n_sites + 1),), torch.int64)
    out = dv.empty((n_sites, n_sites), torch.int64)
    _lib.check(_lib.lib.gd_pair_matrix(dv.ptr(rows), n_rows, stride, cs, cd, n_sites,
                                       dv.ptr(scratch), dv.ptr(out), dv.stream_ptr()))
    return dv.to_host(out).astype(int)


def _fill_device(states: torch.Tensor, direction: int) -> torch.Tensor:
    n_frames, n_atoms = int(states.shape[0]), int(states.shape[1])
    work = dv.empty((int(_lib.lib.gd_events_work_elems(n_frames, n_atoms)),), torch.int32)
    out = dv.empty((n_frames, n_atoms), torch.int32)
    _lib.check(_lib.lib.gd_fill_states(dv.ptr(states), n_frames, n_atoms, direction, None, None,
                                       dv.ptr(work), dv.ptr(out), dv.stream_ptr()))
    return out


def ffill(arr, fill_val: int = -1, axis=-1) -> np.ndarray:
    """Drop-in for utils.py:59-84 on the GPU (2-D integer arrays, fill value -1)."""
    return _fill_host(arr, fill_val, axis, +1)


def bfill(arr, fill_val: int = -1, axis=-1) -> np.ndarray:
    """Drop-in for utils.py:87-98."""
    return _fill_host(arr, fill_val, axis, -1)


def _fill_host(arr, fill_val, axis, direction):
    arr = np.asarray(arr)
    if arr.ndim > 2:
        raise ValueError
    if fill_val != NOSITE:
        raise NotImplementedError('the device kernel fills the NOSITE (-1) value only')
    time_major = arr if axis == 0 else arr.T  # kernel scans along axis 0
    out = widen_to_host(_fill_device(dv.to_device(np.ascontiguousarray(time_major), torch.int32), direction))
    out = out if axis == 0 else out.T
    return out.astype(arr.dtype, copy=False)


# --------------------------------------------------------------------------- Transitions


class Transitions:
    """Container for transitions between sites (reference transitions.py:30-410).

    ``states`` / ``inner_states`` / ``events`` are materialised on first access from
    the device-resident int32 arrays (``d_states`` ...), which the downstream
    kernels (jumps, fills, matrices) consume directly.
    """

    DISORDER_ERROR_MSG = (
        'Input `sites` has partial occupancies. In GEMDAT, '
        '`sites` is treated as the set of all possible sites '
        'for the floating species, and partial occupancies '
        'can lead to ambiguous site assignments. '
        'Remove partial occupancies manually, or set '
        '`remove_part_occup_from_structure=True` '
        'to do it automatically.'
    )

    def __init__(self, *, trajectory, diff_trajectory, sites, events, states, inner_states,
                 site_radius, _device=None):
        if not getattr(sites, 'is_ordered', True):
            warn(self.DISORDER_ERROR_MSG, stacklevel=2)
        self.sites = sites
        self.trajectory = trajectory
        self.diff_trajectory = diff_trajectory
        self.site_radius = site_radius
        self._states = states
        self._inner_states = inner_states
        self._events = events
        self._dev = _device or {}
        self.info = self._dev.get('info', {})

    # -- lazily materialised host views
    @property
    def states(self) -> np.ndarray:
        if self._states is None:
            fut = self.info.get('host_future')
            self._states = fut.outer() if fut is not None else widen_to_host(self._dev['outer'])
        return self._states

    @property
    def inner_states(self) -> np.ndarray:
        if self._inner_states is None:
            fut = self.info.get('host_future')
            self._inner_states = fut.inner() if fut is not None else widen_to_host(self._dev['inner'])
        return self._inner_states

    @property
    def events(self) -> pd.DataFrame:
        if self._events is None:
            self._events = self._dev['events'].to_dataframe(self._dev.get('events_index'))
        return self._events

    @property
    def d_states(self) -> torch.Tensor:
        if 'outer' not in self._dev:
            self._dev['outer'] = dv.to_device(self._states, torch.int32)
        return self._dev['outer']

    @property
    def d_inner_states(self) -> torch.Tensor:
        if 'inner' not in self._dev:
            self._dev['inner'] = dv.to_device(self._inner_states, torch.int32)
        return self._dev['inner']

    @property
    def d_events(self) -> EventTable:
        if 'events' not in self._dev:
            self._dev['events'] = EventTable.from_dataframe(self._events, self.n_floating)
        return self._dev['events']

    @classmethod
    def from_trajectory(
        cls,
        *,
        trajectory,
        sites,
        floating_specie: str,
        site_radius: float | dict[str, float] | None = None,
        site_inner_fraction: float | dict[str, float] = 1.0,
        remove_part_occup_from_structure: bool = False,
        fraction_of_overlap: float = 0.0,
    ) -> 'Transitions':
        """transitions.py:96-186."""
        from .metrics import TrajectoryMetrics

        if not getattr(sites, 'is_ordered', True):
            if remove_part_occup_from_structure:
                sites = remove_partial_occupancies_from_structure(sites)
            else:
                warn(cls.DISORDER_ERROR_MSG, stacklevel=2)

        diff_trajectory = trajectory.filter(floating_specie)

        if site_radius is None:
            vibration_amplitude = TrajectoryMetrics(diff_trajectory).vibration_amplitude()
            site_radius_obj = SiteRadius.from_vibration_amplitude(
                trajectory=trajectory, sites=sites, vibration_amplitude=vibration_amplitude,
                inner_fraction=site_inner_fraction,
            )
        else:
            site_radius_obj = SiteRadius.from_given_radius(
                trajectory=trajectory, sites=sites, radius=site_radius,
                inner_fraction=site_inner_fraction, fraction_of_overlap=fraction_of_overlap,
            )

        # the reference runs _calculate_atom_states twice (outer fraction 1.0, then inner);
        # K1 produces both in one pass over the positions
        outer, inner, info = _site_states_device(
            sites=sites, trajectory=diff_trajectory, site_radius=site_radius_obj.radius,
            site_inner_fraction=site_radius_obj.inner_fraction, count_events=True,
        )
        events = _events_device(outer, inner, info.pop('event_work'), info.pop('changed'))

        return cls(
            sites=sites, trajectory=trajectory, diff_trajectory=diff_trajectory,
            events=None, states=None, inner_states=None, site_radius=site_radius_obj,
            _device={'outer': outer, 'inner': inner, 'events': events, 'info': info},
        )

    def jumps(self, **kwargs) -> 'Jumps':
        """transitions.py:188-202."""
        from .jumps import Jumps

        return Jumps(self, **kwargs)

    def radial_distribution(self, **kwargs):
        """transitions.py:204-223."""
        from .rdf import radial_distribution

        return radial_distribution(transitions=self, **kwargs)

    @property
    def n_floating(self) -> int:
        return len(self.diff_trajectory.species)

    @property
    def n_states(self) -> int:
        if self._states is not None:
            return len(self._states)
        return int(self._dev['outer'].shape[0])

    @property
    def n_events(self) -> int:
        if self._events is not None:
            return len(self._events)
        return len(self._dev['events'])

    @property
    def n_sites(self) -> int:
        return len(self.sites)

    @weak_lru_cache()
    def matrix(self) -> np.ndarray:
        """transitions.py:243-252: [n_sites, n_sites] counts of (start, destination) events."""
        ev = self.d_events
        return _pair_matrix_device(ev.rows if len(ev) else None, len(ev), 6, 1, 2, self.n_sites)

    @weak_lru_cache()
    def states_next(self) -> np.ndarray:
        """transitions.py:254-265 (bfill along time)."""
        return widen_to_host(_fill_device(self.d_states, -1))

    @weak_lru_cache()
    def states_prev(self) -> np.ndarray:
        """transitions.py:267-278 (ffill along time)."""
        return widen_to_host(_fill_device(self.d_states, +1))

    # -- N1: reductions over the device-resident tables (gd_sitestats.cu)
    def site_counts(self) -> np.ndarray:
        """int64 [n_sites + 1]: number of (frame, atom) entries at NOSITE ([0]) and at every
        site ([s + 1]), counted by ``gd_occupancy`` on the int32 device states."""
        st = self.d_states
        counts = dv.zeros((self.n_sites + 1,), torch.int64)
        _lib.check(_lib.lib.gd_occupancy(dv.ptr(st), st.numel(), self.n_sites, dv.ptr(counts),
                                         dv.stream_ptr()))
        return dv.to_host(counts)

    def occupancy(self) -> Sites:
        """Drop-in for transitions.py:280-306: the sites with the fraction of frames each one is
        occupied as their occupancy (the reference's ``np.unique`` count / ``len(states)``)."""
        frac, labels, _ = site_arrays(self.sites)
        visits = self.site_counts()[1:]
        symbols = getattr(self.sites, 'symbols', None) or [
            str(getattr(getattr(s, 'specie', None), 'name', 'X')) for s in self.sites
        ]
        return Sites(getattr(self.sites, 'lattice'), symbols, frac, labels,
                     occupancies=visits / self.n_states)

    def _occupancy_by_label(self) -> dict[str, list[float]]:
        occ = self.occupancy()
        groups: dict[str, list[float]] = {}
        for label, value in zip(occ.labels, occ.occupancies.tolist()):
            groups.setdefault(label, []).append(value)
        return groups

    def occupancy_by_site_type(self) -> dict[str, float]:
        """Drop-in for transitions.py:308-321: mean occupancy of the sites of each label."""
        return {label: sum(vals) / len(vals) for label, vals in self._occupancy_by_label().items()}

    def atom_locations(self) -> dict[str, float]:
        """Drop-in for transitions.py:323-337: fraction of time the atoms spend on each label."""
        n_floating = self.n_floating
        return {label: sum(vals) / n_floating for label, vals in self._occupancy_by_label().items()}

    def _outer_change_table(self) -> torch.Tensor | None:
        """Event rows to derive the visits from: the K2 table when it describes ``states`` (the
        parts made by ``split`` cut states and events at different frames, like the reference),
        else the outer changes re-extracted from the states (K2 with inner = outer)."""
        if 'events' in self._dev and self._dev.get('events_match_states', True):
            return self._dev['events'].rows
        st = self.d_states
        try:
            return _events_device(st, st).rows
        except ValueError:  # no atom ever changes site
            return None

    def residence_time(self) -> pd.DataFrame:
        """Drop-in for transitions.py:339-373: one row per completed visit of a site
        (``gd_residence_count`` / ``gd_residence_fill`` on the event table: a state change of the
        reference's ``np.diff`` is an event row whose start and destination site differ)."""
        _, labels, _ = site_arrays(self.sites)
        columns = ['atom index', 'site', 'label', 'time']
        rows = self._outer_change_table()
        n_rows = C.c_int64(0)
        if rows is not None and len(rows):
            lib, s = _lib.lib, dv.stream_ptr()
            n_events = int(rows.shape[0])
            work = dv.empty((int(lib.gd_residence_work_bytes(n_events)),), torch.uint8)
            _lib.check(lib.gd_residence_count(dv.ptr(rows), n_events, dv.ptr(work), C.byref(n_rows), s))
        if n_rows.value == 0:
            return pd.DataFrame([], columns=columns)
        visits = dv.empty((n_rows.value, 3), torch.int32)
        _lib.check(lib.gd_residence_fill(dv.ptr(rows), n_events, dv.ptr(work), dv.ptr(visits), s))
        v = dv.to_host(visits).astype(np.int64)
        site_labels = np.asarray(labels, dtype=object)
        return pd.DataFrame({'atom index': v[:, 0], 'site': v[:, 1], 'label': site_labels[v[:, 1]],
                             'time': v[:, 2]}, columns=columns)

    def split(self, n_parts: int = 10) -> list['Transitions']:
        """Drop-in for transitions.py:375-410.  The parts share this object's device arrays:
        states are frame slices (``np.array_split`` bounds), every part's event table is cut out
        of the (atom, time)-sorted table by ``gd_events_split_count/fill``."""
        n_states = self.n_states
        event_parts = _split_events_device(self.d_events, n_states, n_parts)
        outer, inner = self.d_states, self.d_inner_states
        trajectories = self.trajectory.split(n_parts)
        diff_trajectories = self.diff_trajectory.split(n_parts)
        big, small = n_states % n_parts, n_states // n_parts
        parts, lo = [], 0
        for i in range(n_parts):
            hi = lo + small + (1 if i < big else 0)
            table, index = event_parts[i]
            parts.append(self.__class__(
                sites=self.sites, trajectory=trajectories[i], diff_trajectory=diff_trajectories[i],
                states=None, inner_states=None, events=None, site_radius=self.site_radius,
                _device={'outer': outer[lo:hi], 'inner': inner[lo:hi], 'events': table,
                         'events_index': index, 'events_match_states': False},
            ))
            lo = hi
        return parts


def _split_events_device(table: 'EventTable', n_states: int, n_parts: int):
    """[(EventTable, source row numbers int64 device tensor)] per part of transitions.py:837-879."""
    if len(table) < n_parts:
        raise ValueError(f'Not enough transitions per part to split into {n_parts}')
    lib, s = _lib.lib, dv.stream_ptr()
    n_atoms = table.n_atoms
    bins = np.linspace(0, n_states + 1, n_parts + 1, dtype=int).astype(np.int64)
    d_bins = dv.to_device(bins)
    part_off = dv.empty((n_parts, n_atoms + 1), torch.int64)
    src_lo = dv.empty((n_parts, n_atoms), torch.int64)
    _lib.check(lib.gd_events_split_count(dv.ptr(table.rows), dv.ptr(table.atom_offsets), n_atoms,
                                         dv.ptr(d_bins), n_parts, dv.ptr(part_off), dv.ptr(src_lo), s))
    sizes = dv.to_host(part_off[:, n_atoms].contiguous())
    base = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    total = int(base[-1])
    rows = dv.empty((max(total, 1), 6), torch.int32)
    index = dv.empty((max(total, 1),), torch.int64)
    if total:
        d_base = dv.to_device(base[:-1].copy())
        _lib.check(lib.gd_events_split_fill(dv.ptr(table.rows), n_atoms, dv.ptr(d_bins), n_parts,
                                            dv.ptr(part_off), dv.ptr(src_lo), dv.ptr(d_base),
                                            dv.ptr(rows), dv.ptr(index), s))
    return [(EventTable(rows[base[p]:base[p + 1]], part_off[p], n_atoms), index[base[p]:base[p + 1]])
            for p in range(n_parts)]


def _split_transitions_events(events: pd.DataFrame, n_states: int, n_parts=10, split_key='time',
                              dependent_keys='time') -> list[pd.DataFrame]:
    """Drop-in for transitions.py:837-879 on a DataFrame: rows with ``bins[p] <= time <
    bins[p + 1]`` per part, the time column shifted to the part's first frame, the original
    index labels kept."""
    if split_key != 'time' or dependent_keys != 'time':
        raise NotImplementedError('the device split works on the time column')
    if len(events) < n_parts:
        raise ValueError(f'Not enough transitions per part to split into {n_parts}')
    table, order = EventTable.from_dataframe(events, return_order=True)
    labels = events.index.to_numpy()
    extra = [c for c in events.columns if c not in EVENT_COLUMNS]
    out = []
    for part, src in _split_events_device(table, n_states, n_parts):
        src_h = order[dv.to_host(src)] if len(part) else np.zeros(0, dtype=np.int64)
        data = widen_to_host(part.rows)
        keep = np.argsort(src_h, kind='stable')  # the reference keeps the incoming row order
        df = pd.DataFrame(data[keep], columns=EVENT_COLUMNS, index=labels[src_h[keep]])
        for c in extra:
            df[c] = events[c].to_numpy()[src_h[keep]]
        out.append(df[list(events.columns)])
    return out


def remove_partial_occupancies_from_structure(structure):
    """utils.py:310-330 for the ``Sites`` stand-in: set every occupancy to 1."""
    new = structure.copy()
    if hasattr(new, 'occupancies'):
        new.occupancies = np.ones(len(new))
        return new
    raise NotImplementedError('pass a gemdat_b200.Sites object or an ordered pymatgen Structure')


# --------------------------------------------------------------------------- SiteRadius


@dataclass
class SiteRadius:
    """transitions.py:485-711 (scalar host prep feeding K1; the S x S distance matrix is a kernel)."""

    radius: dict[str, float]
    inner_fraction: dict[str, float]
    pdist: np.ndarray
    site_pairs: dict
    min_dist: dict

    @classmethod
    def from_given_radius(cls, *, trajectory, sites, radius, inner_fraction,
                          fraction_of_overlap: float = 0.0) -> 'SiteRadius':
        if fraction_of_overlap > 1:
            raise ValueError('fraction_of_overlap must be <= 1.0')
        frac, _, _ = site_arrays(sites)
        pdist = trajectory.get_lattice().get_all_distances(frac, frac)
        r, f = _radius_to_dict(radius=radius, inner_fraction=inner_fraction)
        obj = cls(radius=r, inner_fraction=f, pdist=pdist, site_pairs={}, min_dist={})
        obj._site_pairs()
        obj._min_dist(sites)
        factor = 1 + fraction_of_overlap
        if obj.sites_are_overlapping(factor=factor):
            obj.raise_if_overlapping(sites=sites, factor=factor)
        return obj

    @classmethod
    def from_vibration_amplitude(cls, *, trajectory, sites, vibration_amplitude: float,
                                 inner_fraction=1.0) -> 'SiteRadius':
        frac, labels, _ = site_arrays(sites)
        site_radius = 2 * vibration_amplitude
        pdist = trajectory.get_lattice().get_all_distances(frac, frac)
        min_dist = np.min(pdist[np.triu_indices_from(pdist, k=1)])
        if min_dist < 2 * site_radius:
            site_radius = (0.5 * min_dist) - 0.005
            if site_radius * 2 < 0.5:
                idx = np.argwhere(np.triu(pdist, k=1) == min_dist)
                lines = []
                for i, j in idx:
                    lines.append('\nToo close:')
                    lines.append(f'({i}) {labels[i]} {frac[i]}')
                    lines.append(' - ')
                    lines.append(f'({j}) {labels[i]} {frac[j]}')
                msg = ''.join(lines)
                raise ValueError(
                    'Two sites are within half an Angstrom of each other. '
                    'This is not realistic, check/change the given sites. '
                    f'Expected: > {site_radius * 2:.4f}, '
                    f'got: {min_dist:.4f} for {msg}'
                )
        r, f = _radius_to_dict(radius=float(site_radius), inner_fraction=inner_fraction)
        obj = cls(radius=r, pdist=pdist, inner_fraction=f, site_pairs={}, min_dist={})
        obj._site_pairs()
        obj._min_dist(sites)
        return obj

    def outer_states_fraction(self) -> dict[str, float]:
        return {label: 1.0 for label in self.radius}

    def _site_pairs(self):
        pairs = list(combinations_with_replacement(self.radius.keys(), 2))
        self.site_pairs = {(i, j): self.radius[i] + self.radius[j] for (i, j) in pairs}

    def _min_dist(self, sites):
        self.min_dist = {}
        _, labels, _ = site_arrays(sites)
        site_labels = np.array(labels)
        for (i, j), _pair_dist in self.site_pairs.items():
            ii = np.where(site_labels == i)[0]
            jj = np.where(site_labels == j)[0]
            if ii.size == 0 or jj.size == 0:
                self.min_dist[(i, j)] = float('inf')
                continue
            sub = self.pdist[np.ix_(ii, jj)]
            _, a, b = np.intersect1d(ii, jj, return_indices=True)
            if a.size:
                sub[a, b] = np.inf
            self.min_dist[(i, j)] = float(sub.min(initial=np.inf))

    def sites_are_overlapping(self, factor: float = 1.0) -> bool:
        return any(factor * self.min_dist[key] < pair for key, pair in self.site_pairs.items())

    def raise_if_overlapping(self, sites, factor: float = 1.0) -> None:
        frac, labels, _ = site_arrays(sites)
        lines = []
        for key, pair_dist in self.site_pairs.items():
            min_dist = self.min_dist[key]
            if factor * min_dist < pair_dist:
                for i, j in np.argwhere(np.triu(self.pdist, k=1) == min_dist):
                    lines.append('\nOverlapping: ')
                    lines.append(f'({i}) {frac[i]}')
                    lines.append(' - ')
                    lines.append(f'({j}) {frac[j]}, ')
                    lines.append(f'expected < {min_dist / 2:.2f}, ')
                    lines.append(f'got {labels[i]}: {self.radius[labels[i]]} ')
                    lines.append(f'and {labels[j]}: {self.radius[labels[j]]}')
        msg = ''.join(lines)
        raise ValueError(
            'Sites are overlapping with the chosen site_radius '
            'and fraction_of_overlap, make site_radius smaller '
            f'for {msg}'
        )


def _radius_to_dict(radius, inner_fraction):
    """transitions.py:713-733."""
    if isinstance(radius, (int, float)):
        if isinstance(inner_fraction, dict):
            r = {label: radius for label in inner_fraction.keys()}
        elif isinstance(inner_fraction, (int, float)):
            r = {'': radius}
    elif isinstance(radius, dict):
        r = radius
    else:
        raise TypeError(f'Invalid type for `site_radius`: {type(radius)}')
    if isinstance(inner_fraction, (int, float)):
        f = {label: inner_fraction for label in r.keys()}
    elif isinstance(inner_fraction, dict):
        f = inner_fraction
    else:
        raise TypeError(f'Invalid type for `site_inner_fraction`: {type(inner_fraction)}')
    return r, f
