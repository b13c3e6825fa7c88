"""Jumps: host-side mirror of reference src/gemdat/jumps.py for the hot path.

``Jumps(transitions, conversion_method=..., minimal_residence=0)`` (:123-151) keeps the
reference's plug-in hook: any ``Callable[[Transitions, minimal_residence=int],
DataFrame]``.  The default method is the CUDA state machine (``gd_jumps_count`` /
``gd_jumps_fill``, K3) standing in for the pandas ``groupby``/``iterrows`` loop of
``_generic_transitions_to_jumps`` (:24-119).  ``matrix`` (:212-221) and
``jump_diffusivity`` (:185-210) follow the reference formulae.  ``collective`` (:223-251)
hands the device table to ``gemdat_b200.collective.Collective``.  ``rates`` (:435-462),
``activation_energies`` (:253-308), ``to_graph`` (:364-410) and
``activation_energy_between_sites`` (:340-362) are the consumers the N1 / N2 rows exist for:
a few scalars per site pair on the host, fed by ``Transitions.split`` / ``atom_locations`` /
``occupancy`` (device reductions), the per-part jump tables (K3) and the attempt frequency (N2).
"""

from __future__ import annotations

import ctypes as C
from collections import Counter
from itertools import product
from math import ceil
from typing import Callable

import numpy as np
import pandas as pd
import torch

from . import _device as dv
from . import _lib
from .caching import weak_lru_cache
from .collective import Collective
from .lattice import FloatWithUnit, site_arrays
from .transitions import Transitions, _pair_matrix_device

ANGSTROM = 1e-10  # scipy.constants.angstrom

JUMP_COLUMNS = [
    'atom index',
    'start site',
    'destination site',
    'start time',
    'stop time',
    'residence_time',
]


def _jumps_device(transitions: Transitions, minimal_residence: int) -> torch.Tensor:
    """int32 [J, 6] device table; raises ValueError('No jumps found') like jumps.py:107-108."""
    ev = transitions.d_events
    n_atoms = ev.n_atoms
    n_frames_total = len(transitions.trajectory)  # jumps.py:39
    lib = _lib.lib
    n_events = len(ev)
    work = dv.empty((int(lib.gd_jumps_work_bytes(n_events, n_atoms)),), torch.uint8)
    n_jumps = C.c_int64(0)
    s = dv.stream_ptr()
    _lib.check(lib.gd_jumps_count(dv.ptr(ev.rows), dv.ptr(ev.atom_offsets), n_events, n_atoms,
                                  n_frames_total, int(minimal_residence), dv.ptr(work),
                                  C.byref(n_jumps), s))
    if n_jumps.value == 0:
        raise ValueError('No jumps found')
    out = dv.empty((n_jumps.value, 6), torch.int32)
    _lib.check(lib.gd_jumps_fill(dv.ptr(ev.rows), dv.ptr(ev.atom_offsets), n_events, n_atoms,
                                 n_frames_total, int(minimal_residence), dv.ptr(work), dv.ptr(out), s))
    return out


def _generic_transitions_to_jumps(transitions: Transitions, *, minimal_residence: int = 0) -> pd.DataFrame:
    """Drop-in for jumps.py:24-119."""
    table = _jumps_device(transitions, minimal_residence)
    transitions._dev['jumps'] = table
    return pd.DataFrame(data=dv.to_host(table).astype(np.int64), columns=JUMP_COLUMNS, copy=False)


class Jumps:
    def __init__(
        self,
        transitions: Transitions,
        *,
        conversion_method: Callable[..., pd.DataFrame] = _generic_transitions_to_jumps,
        minimal_residence: int = 0,
    ):
        self.transitions = transitions
        self.trajectory = transitions.diff_trajectory
        self.sites = transitions.sites
        self.conversion_method = conversion_method
        self.data = conversion_method(transitions, minimal_residence=minimal_residence)
        self.minimal_residence = minimal_residence
        # device copy of the table when the default method made it (used by collective())
        self._d_table = transitions._dev.pop('jumps', None) if hasattr(transitions, '_dev') else None

    @property
    def n_jumps(self) -> int:
        return len(self.data)

    @property
    def n_solo_jumps(self) -> int:
        """jumps.py:158-161."""
        return self.collective().n_solo_jumps

    @property
    def n_floating(self) -> int:
        return len(self.trajectory.species)

    @property
    def solo_fraction(self) -> float:
        """jumps.py:168-171."""
        return self.n_solo_jumps / self.n_jumps

    @property
    def site_pairs(self) -> list[tuple[str, str]]:
        _, labels, _ = site_arrays(self.sites)
        return list(product(labels, repeat=2))

    @property
    def jump_names(self) -> list[str]:
        return ['->'.join(key) for key in self.site_pairs]

    @weak_lru_cache()
    def jump_diffusivity(self, dimensions: int) -> float:
        """jumps.py:185-210: sum(pdist^2 * matrix) * A^2 / (2 d n_floating total_time), m^2/s."""
        frac, _, _ = site_arrays(self.sites)
        pdist = self.trajectory.get_lattice().get_all_distances(frac, frac)
        jump_diff = np.sum(pdist**2 * self.matrix())
        jump_diff *= ANGSTROM**2 / (2 * dimensions * self.n_floating * self.trajectory.total_time)
        return FloatWithUnit(jump_diff, 'm^2 s^-1')

    @weak_lru_cache()
    def matrix(self) -> np.ndarray:
        """jumps.py:212-221: [n_sites, n_sites] jump counts."""
        pairs = np.ascontiguousarray(
            self.data[['start site', 'destination site']].to_numpy(dtype=np.int32)
        )
        rows = dv.to_device(pairs) if len(pairs) else None
        return _pair_matrix_device(rows, len(pairs), 2, 0, 1, self.transitions.n_sites)

    @weak_lru_cache()
    def collective(self, max_dist: float = 1) -> Collective:
        """jumps.py:223-251: window of 1 / attempt frequency, in frames."""
        from .metrics import TrajectoryMetrics

        trajectory = self.trajectory
        attempt_freq, _ = TrajectoryMetrics(trajectory).attempt_frequency()
        max_steps = ceil(1.0 / (attempt_freq * trajectory.time_step))
        return Collective(jumps=self, sites=self.transitions.sites, lattice=trajectory.get_lattice(),
                          max_steps=max_steps, max_dist=max_dist)

    @weak_lru_cache()
    def counter(self) -> Counter:
        """jumps.py:310-324: jump counts keyed by (start label, destination label)."""
        _, labels, _ = site_arrays(self.sites)
        out: Counter = Counter()
        for (i, j), val in self._counter().items():
            out[labels[i], labels[j]] += val
        return out

    @weak_lru_cache()
    def _counter(self) -> Counter:
        return Counter(zip(self.data['start site'], self.data['destination site']))

    def _part_counters(self, n_parts: int) -> list[Counter]:
        return [part.counter() for part in self.split(n_parts)]

    def _unique_site_pairs(self) -> list[tuple[str, str]]:
        """``site_pairs`` without its repetitions, in the same order of first appearance (the
        reference walks all n_sites^2 pairs and overwrites equal keys with equal values)."""
        _, labels, _ = site_arrays(self.sites)
        return list(product(dict.fromkeys(labels), repeat=2))

    @weak_lru_cache()
    def rates(self, n_parts: int = 10) -> pd.DataFrame:
        """jumps.py:435-462: jumps per second and per floating atom between site labels, mean and
        sample standard deviation (ddof=1) over ``n_parts`` consecutive parts of the run."""
        counters = self._part_counters(n_parts)
        denom = self.n_floating * (self.trajectory.total_time / n_parts)
        table = {}
        for pair in self._unique_site_pairs():
            n_jumps = [c[pair] for c in counters]
            table[pair] = float(np.mean(n_jumps) / denom), float(np.std(n_jumps, ddof=1) / denom)
        df = pd.DataFrame(table).T
        df.columns = ('rates', 'std')
        return df

    @weak_lru_cache()
    def activation_energies(self, n_parts: int = 10) -> pd.DataFrame:
        """jumps.py:253-308: E_act = -ln(rate / attempt frequency) k_B T in eV per pair of site
        labels, mean and sample standard deviation over ``n_parts`` parts.  The rate of a part is
        its jump count over (time the atoms spend on the start label: ``atom_locations`` of the
        part x n_floating x part duration); A->A jumps are halved (:293-296)."""
        from scipy.constants import Boltzmann, elementary_charge

        from .metrics import TrajectoryMetrics

        trajectory = self.trajectory
        attempt_freq, _ = TrajectoryMetrics(trajectory).attempt_frequency()
        temperature = trajectory.metadata['temperature']
        locations = [part.atom_locations() for part in self.transitions.split(n_parts)]
        counters = self._part_counters(n_parts)
        n_floating = self.n_floating
        part_time = trajectory.total_time / n_parts
        table = {}
        for start, stop in self._unique_site_pairs():
            n_jumps = np.array([c[start, stop] for c in counters])
            share = np.array([loc[start] for loc in locations])
            eff_rate = n_jumps / (share * n_floating * part_time)
            if start == stop:
                eff_rate /= 2
            e_act = -np.log(eff_rate / attempt_freq) * (Boltzmann * temperature) / elementary_charge
            table[start, stop] = np.mean(e_act), np.std(e_act, ddof=1)
        df = pd.DataFrame(table).T
        df.columns = ('energy', 'std')
        return df

    def activation_energy_between_sites(self, start, stop) -> float:
        """jumps.py:340-362: the ``e_act`` attribute of the graph edge start -> stop."""
        edge = self.to_graph().get_edge_data(start, stop)
        if not edge:
            raise IndexError(f'No jumps between ({start}) and ({stop})')
        return edge['e_act']

    @weak_lru_cache()
    def to_graph(self, min_e_act: float | None = None, max_e_act: float | None = None):
        """jumps.py:364-410: networkx DiGraph, one node per site (attribute ``label``), one edge
        per (start site, destination site) that has jumps, weighted by the activation energy from
        the whole run: rate = jumps / (site occupancy x total time)."""
        import networkx as nx
        from scipy.constants import Boltzmann, elementary_charge

        from .metrics import TrajectoryMetrics

        lo = min_e_act if min_e_act else float('-inf')
        hi = max_e_act if max_e_act else float('inf')
        occupancy = self.transitions.occupancy().occupancies
        attempt_freq, _ = TrajectoryMetrics(self.trajectory).attempt_frequency()
        kBT = Boltzmann * self.trajectory.metadata['temperature']
        total_time = self.trajectory.total_time
        _, labels, _ = site_arrays(self.sites)
        graph = nx.DiGraph()
        for i, label in enumerate(labels):
            graph.add_node(i, label=label)
        for (start, stop), n_jumps in self._counter().items():
            eff_rate = n_jumps / (occupancy[start] * total_time)
            e_act = -np.log(eff_rate / attempt_freq) * kBT
            e_act /= elementary_charge
            if lo <= e_act <= hi:
                graph.add_edge(start, stop, e_act=e_act)
        return graph

    def split(self, n_parts: int) -> list['Jumps']:
        """jumps.py:412-433."""
        return [
            Jumps(part, conversion_method=self.conversion_method,
                  minimal_residence=self.minimal_residence)
            for part in self.transitions.split(n_parts)
        ]
