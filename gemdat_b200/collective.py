"""Collective: host-side mirror of reference src/gemdat/collective.py.

``Collective(jumps, sites, lattice, max_steps, max_dist)`` (:30-59) runs the pair
search of ``_compute`` (:61-106) on the device: ``gd_collective_sort`` orders the jump
table by (stop time, start time), ``gd_collective_close`` thresholds the site-site
distance matrix, ``gd_collective_count`` / ``gd_collective_fill`` walk the pairs with the
reference's break / continue rules and write them in its emission order.  The results
carry the reference's attribute names (``collective``, ``coll_jumps``, ``n_solo_jumps``,
``n_coll_jumps``) and methods (``site_pair_count_matrix``, ``..._labels``,
``multiple_collective``); ``pairs`` / ``d_pairs`` / ``events`` are the same data as index
arrays for callers that do not want one pandas row per pair.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import pandas as pd
import torch

from . import _device as dv
from . import _lib
from .caching import weak_lru_cache
from .lattice import as_lattice_matrix, site_arrays

_COLS = ['atom index', 'start site', 'destination site', 'start time', 'stop time', 'residence_time']


def _close_matrix(lattice, frac: np.ndarray, max_dist: float) -> torch.Tensor:
    """uint8 [n_sites, n_sites] on the device: minimum-image distance < max_dist."""
    n = len(frac)
    d_frac = dv.to_device(np.ascontiguousarray(frac, dtype=np.float64))
    dist = dv.empty((n, n), torch.float64)
    matrix = np.ascontiguousarray(as_lattice_matrix(lattice), dtype=np.float64)
    s = dv.stream_ptr()
    _lib.check(_lib.lib.gd_pbc_distances(matrix.ctypes.data, dv.ptr(d_frac), n, dv.ptr(d_frac), n,
                                         dv.ptr(dist), s))
    close = dv.empty((n, n), torch.uint8)
    _lib.check(_lib.lib.gd_collective_close(dv.ptr(dist), n, float(max_dist), dv.ptr(close), s))
    return close


def collective_pairs_device(d_jumps: torch.Tensor, close: torch.Tensor, max_steps: int):
    """(sorted jump table int32 [J, 6], pairs int32 [P, 2], n_solo) on the device."""
    lib = _lib.lib
    n = int(d_jumps.shape[0])
    n_sites = int(close.shape[0])
    s = dv.stream_ptr()
    work = dv.empty((int(lib.gd_collective_work_bytes(n)),), torch.uint8)
    ordered = dv.empty((n, 6), torch.int32)
    _lib.check(lib.gd_collective_sort(dv.ptr(d_jumps), n, dv.ptr(work), dv.ptr(ordered), s))
    n_pairs, n_solo = C.c_int64(0), C.c_int64(0)
    _lib.check(lib.gd_collective_count(dv.ptr(ordered), n, dv.ptr(close), n_sites, int(max_steps),
                                       dv.ptr(work), C.byref(n_pairs), C.byref(n_solo), s))
    pairs = dv.empty((n_pairs.value, 2), torch.int32)
    if n_pairs.value:
        _lib.check(lib.gd_collective_fill(dv.ptr(ordered), n, dv.ptr(close), n_sites, int(max_steps),
                                          dv.ptr(work), dv.ptr(pairs), s))
    return ordered, pairs, int(n_solo.value)


class Collective:
    def __init__(self, jumps, sites, lattice, max_steps: int, max_dist: float = 1):
        self.jumps = jumps
        self.sites = sites
        self.lattice = lattice
        self.max_steps = max_steps
        self.max_dist = max_dist
        self._compute()

    def _compute(self):
        """collective.py:61-106."""
        frac, _, _ = site_arrays(self.sites)
        data = self.jumps.data
        # the device table left by the default conversion method, else the frame's values
        table = getattr(self.jumps, '_d_table', None)
        if table is None or int(table.shape[0]) != len(data):
            table = dv.to_device(np.ascontiguousarray(data[_COLS].to_numpy(dtype=np.int32)).reshape(-1, 6))
        close = _close_matrix(self.lattice, frac, self.max_dist)
        ordered, pairs, n_solo = collective_pairs_device(table, close, self.max_steps)
        self.d_events = ordered
        self.d_pairs = pairs
        self.n_solo_jumps = n_solo
        self.n_coll_jumps = len(data) - n_solo

    # ---- index-array views
    @property
    def events(self) -> pd.DataFrame:
        """The jump table in the order the pairs refer to (collective.py:68)."""
        return self._events()

    @weak_lru_cache()
    def _events(self) -> pd.DataFrame:
        return pd.DataFrame(dv.to_host(self.d_events).astype(np.int64), columns=_COLS, copy=False)

    @property
    def pairs(self) -> np.ndarray:
        """int64 [P, 2]: rows (i, j) of ``events``, in the reference's emission order."""
        return self._pairs()

    @weak_lru_cache()
    def _pairs(self) -> np.ndarray:
        return dv.to_host(self.d_pairs).astype(np.int64).reshape(-1, 2)

    # ---- the reference's attributes
    @property
    def collective(self) -> list:
        """List of (event_i, event_j) rows like collective.py:89 (one pandas row each)."""
        ev = self.events
        rows = [ev.iloc[k] for k in range(len(ev))]
        return [(rows[i], rows[j]) for i, j in self.pairs.tolist()]

    @property
    def coll_jumps(self) -> list:
        """((start_i, dest_i), (start_j, dest_j)) like collective.py:90-96."""
        return [((a, b), (c, d)) for a, b, c, d in self._coll_sites().tolist()]

    @weak_lru_cache()
    def _coll_sites(self) -> np.ndarray:
        sd = self.events[['start site', 'destination site']].to_numpy()
        p = self.pairs
        return np.concatenate([sd[p[:, 0]], sd[p[:, 1]]], axis=1).reshape(-1, 4)

    @weak_lru_cache()
    def site_pair_count_matrix(self) -> np.ndarray:
        """collective.py:108-133."""
        _, labels, _ = site_arrays(self.sites)
        site_pairs = self.site_pair_count_matrix_labels()
        index = {p: k for k, p in enumerate(site_pairs)}
        names = sorted(set(labels))
        code = {name: k for k, name in enumerate(names)}
        lab = np.array([code[x] for x in labels], dtype=np.int64)
        # (label, label) code -> position in site_pairs
        lut = np.array([[index[(a, b)] for b in names] for a in names], dtype=np.int64)
        cs = self._coll_sites()
        out = np.zeros((len(site_pairs), len(site_pairs)), dtype=int)
        if len(cs):
            pi = lut[lab[cs[:, 0]], lab[cs[:, 1]]]
            pj = lut[lab[cs[:, 2]], lab[cs[:, 3]]]
            np.add.at(out, (pi, pj), 1)
        return out

    @weak_lru_cache()
    def site_pair_count_matrix_labels(self) -> list:
        _, labels, _ = site_arrays(self.sites)
        return list({(label1, label2) for label1 in labels for label2 in labels})

    @weak_lru_cache()
    def multiple_collective(self) -> tuple[np.ndarray, np.ndarray]:
        """collective.py:140-166: np.unique over the site pairs, sorted within each pair."""
        cs = self._coll_sites()
        collective = np.empty((len(cs), 2), dtype=[('start', int), ('stop', int)])
        collective['start'][:, 0], collective['stop'][:, 0] = cs[:, 0], cs[:, 1]
        collective['start'][:, 1], collective['stop'][:, 1] = cs[:, 2], cs[:, 3]
        collective = np.sort(collective, axis=1)
        return np.unique(collective, axis=0, return_counts=True)
