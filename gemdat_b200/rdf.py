"""Radial distribution functions: host-side mirror of reference src/gemdat/rdf.py.

``radial_distribution(transitions=, floating_specie=, max_dist=5.0, resolution=0.1)``
(rdf.py:111-190) and ``radial_distribution_between_species(trajectory=, specie_1=,
specie_2=, ...)`` (:193-268) with the reference's names, keyword-only arguments and return
types (``dict[str, RDFCollection]`` / ``RDFData``).  The per-frame loop over
``lattice.get_all_distances`` + ``np.digitize`` / ``np.histogram`` + ``np.bincount`` is one
kernel (``gd_rdf_hist``); the transition state of every floating atom in every frame comes
from ``gd_rdf_classes`` on the device-resident ``states`` / ``states_prev`` / ``states_next``.

Kept as the reference has them: ``_uniqify_labels``' off-by-one (site 0 counts as "no site",
site s takes the label of site s-1, rdf.py:19-27), the ``'~>'`` name built from
``unique_labels[-1]`` when the previous site is unknown (:40-45), self distances counted in
bin 0 of the state-resolved RDF, and self pairs removed from the species RDF only when
``specie_1 == specie_2`` compares equal (:250-254).  ``unique_labels`` is ``list(set(labels))``
as in the reference, so the ``'~>'`` names follow this process' string hashing exactly as
they do there.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Collection

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .lattice import as_lattice_matrix, site_arrays
from .trajectory import _symbol
from .transitions import _fill_device


@dataclass
class RDFData:
    """rdf.py:73-98 (plotting is out of scope)."""

    x: np.ndarray
    y: np.ndarray
    label: str
    state: str


class RDFCollection(list):
    """rdf.py:101-108."""


def _state_names(unique_labels: list[str]):
    """rdf.py:30-50 as (dense code -> name index, names); dense code = ((i+1)(L+1) + (j+1))(L+1) + (k+1)."""
    n = len(unique_labels)
    names: list[str] = []
    index: dict[str, int] = {}
    lut = np.empty((n + 1) ** 3, dtype=np.int32)
    r = [-1] + list(range(n))
    for i in r:
        for j in r:
            for k in r:
                if i != -1:
                    name = '@' + unique_labels[i]
                elif j == -1 or k == -1:
                    name = '~>' + unique_labels[j]
                else:
                    name = unique_labels[j] + '->' + unique_labels[k]
                if name not in index:
                    index[name] = len(names)
                    names.append(name)
                lut[((i + 1) * (n + 1) + (j + 1)) * (n + 1) + (k + 1)] = index[name]
    return lut, names


def _hist(pos: torch.Tensor, lattice, idx1, cls, n_class, idx2, group, n_group, bins: np.ndarray,
          resolution: float, mode: int, skip_same: bool) -> np.ndarray:
    """uint64 counts [n_class, n_group, len(bins) + 1] from gd_rdf_hist (host copy)."""
    return dv.to_host(_hist_device(pos, lattice, idx1, cls, n_class, idx2, group, n_group, bins,
                                   resolution, mode, skip_same))


def _hist_device(pos: torch.Tensor, lattice, idx1, cls, n_class, idx2, group, n_group, bins: np.ndarray,
                 resolution: float, mode: int, skip_same: bool) -> torch.Tensor:
    """The same table left on the device (frame shards add theirs up with an all-reduce)."""
    n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
    d_idx1 = dv.to_device(np.ascontiguousarray(idx1, dtype=np.int32))
    d_idx2 = dv.to_device(np.ascontiguousarray(idx2, dtype=np.int32))
    d_group = dv.to_device(np.ascontiguousarray(group, dtype=np.int32)) if group is not None else None
    d_edges = dv.to_device(np.ascontiguousarray(bins, dtype=np.float64))
    hist = dv.zeros((n_class, n_group, len(bins) + 1), torch.int64)
    matrix = np.ascontiguousarray(as_lattice_matrix(lattice), dtype=np.float64)
    _lib.check(_lib.lib.gd_rdf_hist(
        dv.ptr(pos), n_frames, n_atoms, matrix.ctypes.data, dv.ptr(d_idx1), len(idx1),
        dv.ptr(cls) if cls is not None else None, n_class, dv.ptr(d_idx2),
        dv.ptr(d_group) if d_group is not None else None, len(idx2), n_group, dv.ptr(d_edges), len(bins),
        float(resolution), mode, int(skip_same), dv.ptr(hist), dv.stream_ptr()))
    return hist


def radial_distribution(*, transitions, floating_specie: str, max_dist: float = 5.0,
                        resolution: float = 0.1) -> dict[str, RDFCollection]:
    """Drop-in for rdf.py:111-190."""
    trajectory = transitions.trajectory  # note: needs the trajectory with ALL species
    _, labels, _ = site_arrays(transitions.sites)
    labels = list(labels)
    symbols = [_symbol(sp) for sp in trajectory.species]
    symbol_set = tuple(sorted(set(symbols)))  # Structure.symbol_set
    floating = [n for n, s in enumerate(symbols) if s == _symbol(floating_specie)]

    unique_labels = list(set(labels))  # rdf.py:21,32
    lut, names = _state_names(unique_labels)
    site_label = np.array([unique_labels.index(lab) for lab in labels], dtype=np.int32)

    bins = np.arange(0, max_dist + resolution, resolution)
    pos = trajectory._device_positions()
    n_frames = int(pos.shape[0])
    states = transitions.d_states
    assert int(states.shape[1]) == len(floating), 'transitions do not belong to this floating specie'
    cls = dv.empty((n_frames, len(floating)), torch.int32)
    first_key = torch.full((len(names),), torch.iinfo(torch.int64).max, dtype=torch.int64, device=pos.device)
    # named so that the buffers outlive the launch (a temporary would hand its block back to the
    # caching allocator, and the next temporary would overwrite it before the kernel runs)
    prev, nxt = _fill_device(states, +1), _fill_device(states, -1)
    d_site_label, d_lut = dv.to_device(site_label), dv.to_device(lut)
    _lib.check(_lib.lib.gd_rdf_classes(
        dv.ptr(states), dv.ptr(prev), dv.ptr(nxt), n_frames, len(floating), dv.ptr(d_site_label),
        len(labels), len(unique_labels), dv.ptr(d_lut), dv.ptr(cls), dv.ptr(first_key), dv.stream_ptr()))
    del prev, nxt
    group = np.array([symbol_set.index(s) for s in symbols], dtype=np.int32)
    hist = _hist(pos, trajectory.get_lattice(), floating, cls, len(names), np.arange(len(symbols)), group,
                 len(symbol_set), bins, resolution, 0, False)
    # the reference's dict is filled frame by frame, states in ascending code order (:163-173)
    keys = dv.to_host(first_key)
    order = [int(c) for c in np.argsort(keys, kind='stable') if keys[c] != np.iinfo(np.int64).max]
    ret: dict[str, RDFCollection] = {}
    for c in order:
        coll = ret.setdefault(names[c], RDFCollection())
        for g, symbol in enumerate(symbol_set):
            # drop the last element: distance > max_dist (:184)
            coll.append(RDFData(x=bins, y=hist[c, g, :-1].astype(int), label=symbol, state=names[c]))
    return ret


def _hist_frames(pos: torch.Tensor, lattices: np.ndarray, idx1, idx2, bins: np.ndarray, resolution: float,
                 skip_same: bool) -> np.ndarray:
    """uint64 counts [n_frames, len(bins) + 1] of gd_rdf_hist_frames (one cell per frame, np.histogram
    binning, one group), host copy."""
    n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
    lib = _lib.lib
    d_idx1 = dv.to_device(np.ascontiguousarray(idx1, dtype=np.int32))
    d_idx2 = dv.to_device(np.ascontiguousarray(idx2, dtype=np.int32))
    d_edges = dv.to_device(np.ascontiguousarray(bins, dtype=np.float64))
    work = dv.empty((int(lib.gd_rdf_frames_work_bytes(n_frames)),), torch.uint8)
    hist = dv.zeros((n_frames, len(bins) + 1), torch.int64)
    _lib.check(lib.gd_rdf_hist_frames(
        dv.ptr(pos), n_frames, n_atoms, lattices.ctypes.data, dv.ptr(work), dv.ptr(d_idx1), len(idx1), None,
        dv.ptr(d_idx2), None, len(idx2), 1, dv.ptr(d_edges), len(bins), float(resolution), 1, int(skip_same),
        dv.ptr(hist), dv.stream_ptr()))
    return dv.to_host(hist)


def radial_distribution_between_species(*, trajectory, specie_1: str | Collection[str],
                                        specie_2: str | Collection[str], max_dist: float = 5.0,
                                        resolution: float = 0.1) -> RDFData:
    """Drop-in for rdf.py:193-268; with ``constant_lattice=False`` every frame has its own cell
    (``gd_rdf_hist_frames``: per-frame counts, normalised with the frame's volume like :248-261)."""
    if trajectory.lattice is None:
        raise RuntimeError('trajectory.lattice unexpectedly None')
    symbols = [_symbol(sp) for sp in trajectory.species]

    def select(species):
        wanted = {_symbol(s) for s in ([species] if isinstance(species, str) else species)}
        return [n for n, s in enumerate(symbols) if s in wanted]

    idx1, idx2 = select(specie_1), select(specie_2)
    num_atoms_1, num_atoms_2 = len(idx1), len(idx2)
    # the reference reads `.coords` (:222-223); for a trajectory held as positions that is this
    pos = trajectory._device_positions()
    num_time_steps = int(pos.shape[0])
    bins = np.arange(0, max_dist + resolution, resolution)
    same = specie_1 == specie_2
    str1 = specie_1 if isinstance(specie_1, str) else '/'.join(specie_1)
    str2 = specie_2 if isinstance(specie_2, str) else '/'.join(specie_2)
    shell = (bins + resolution) ** 3 - bins**3
    if not getattr(trajectory, 'constant_lattice', True):
        lattices = np.ascontiguousarray(trajectory.lattice, dtype=np.float64).reshape(num_time_steps, 9)
        per_frame = _hist_frames(pos, lattices, idx1, idx2, bins, resolution, same)  # [T, len(bins) + 1]
        hist_t = per_frame[:, 1:len(bins)].astype(float)
        volumes = np.abs(np.linalg.det(lattices.reshape(-1, 3, 3)))
        # hist_t / norm_t with norm_t = (n1 n2 / V_t) (4/3) pi shell, summed over the frames
        particle_vol = num_atoms_2 * num_atoms_1 / volumes
        norm = particle_vol[:, None] * ((4 / 3) * np.pi * shell)[None, :-1]
        counts = (hist_t / norm).sum(axis=0) / num_time_steps
        return RDFData(x=bins[:-1], y=counts, label=f'{str1}-{str2}', state='')
    hist = _hist(pos, trajectory.get_lattice(), idx1, None, 1, idx2, None, 1, bins, resolution, 1, same)
    hist_total = hist[0, 0, 1:len(bins)].astype(float)

    lat = trajectory.get_lattice()
    particle_vol = num_atoms_2 * num_atoms_1 / lat.volume
    norm = (particle_vol * (4 / 3) * np.pi * shell)[:-1]
    # sum over frames of hist_t / norm with one norm for a constant lattice (fp64; the reference
    # divides frame by frame, agreement ~1e-15 relative)
    counts = hist_total / norm / num_time_steps
    return RDFData(x=bins[:-1], y=counts, label=f'{str1}-{str2}', state='')
