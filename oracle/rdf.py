"""Radial distribution functions, CPU restatement (TEST INFRASTRUCTURE: only tests/, smoke()
and the cpu_baseline leg of bench.py may import this).

Follows reference src/gemdat/rdf.py:

* ``uniqify`` (:19-27): site index -> label index through ``np.digitize(arr, arange(n),
  right=True)``.  With ``right=True`` both -1 and 0 fall in bin 0, and a state s >= 1 reads
  ``mapping[s]`` = the label of site s-1: site 0 counts as "no site" and every other site
  takes the label of its predecessor.  Kept as the reference has it.
* ``state_names`` (:30-50): code ``int(i*1e6 + j*1e3 + k)`` over i, j, k in {-1, 0..L-1} for
  (now, previous, next); ``'@' + label_i`` when on a site, ``'~>' + unique_labels[j]`` when
  previous or next is unknown (j = -1 indexes the LAST label), else ``label_j->label_k``.
* ``states_array`` (:53-61).
* ``radial_distribution`` (:111-190): per frame, distances floating -> all atoms,
  ``np.digitize(d, bins, right=True)``, one bincount per (state of the floating atom, symbol
  of the other atom); the overflow bin is dropped at the end (:184).  Keys appear in the
  order the reference's defaultdict creates them: frame, then ascending state code, then
  symbol.
* ``between_species`` (:193-268): ``np.histogram`` per frame over ``bins`` (self pairs dropped
  when the two species arguments compare equal), divided by the shell volume x pair density,
  averaged over frames.

``unique_labels`` is ``list(set(labels))`` in the reference: its order follows the process'
string hashing, so callers pass the order to use.
Pinned by tests/golden/rdf.npz (made by executing the reference, make_golden_rdf.py).
"""

from __future__ import annotations

import numpy as np

from .pmg import Lattice


def uniqify(arr: np.ndarray, labels, unique_labels) -> np.ndarray:
    mapping = np.array([-1] + [unique_labels.index(lab) for lab in labels])
    arr = np.asarray(arr)
    return mapping[np.where(arr >= 1, arr, 0)]


def state_code(i, j, k):
    return int(i * 1e6 + j * 1e3 + k)


def state_names(unique_labels) -> dict:
    names = {}
    r = [-1] + list(range(len(unique_labels)))
    for i in r:
        for j in r:
            for k in r:
                if i != -1:
                    name = '@' + unique_labels[i]
                elif j == -1 or k == -1:
                    name = '~>' + unique_labels[j]
                else:
                    name = unique_labels[j] + '->' + unique_labels[k]
                names[state_code(i, j, k)] = name
    return names


def states_array(states, states_prev, states_next, labels, unique_labels) -> np.ndarray:
    s = uniqify(states, labels, unique_labels)
    p = uniqify(states_prev, labels, unique_labels)
    n = uniqify(states_next, labels, unique_labels)
    return (s * 1e6 + p * 1e3 + n).astype(int)


def radial_distribution(*, positions, symbols, floating_specie, codes, unique_labels, lattice,
                        max_dist=5.0, resolution=0.1) -> dict:
    """{(state name, symbol): int counts[len(bins)]} in the reference's insertion order."""
    positions = np.asarray(positions, dtype=np.float64)
    lat = Lattice(lattice)
    names = state_names(unique_labels)
    symbol_set = tuple(sorted(set(symbols)))
    symbol_idx = {sym: np.array([n for n, s in enumerate(symbols) if s == sym]) for sym in symbol_set}
    floating = np.array([n for n, s in enumerate(symbols) if s == floating_specie])
    bins = np.arange(0, max_dist + resolution, resolution)
    length = len(bins) + 1
    out: dict = {}
    for t in range(len(positions)):
        d = lat.get_all_distances(positions[t, floating], positions[t])
        which = np.digitize(d, bins, right=True)
        for code in np.unique(codes[t]):
            rows = np.nonzero(codes[t] == code)[0]
            for sym in symbol_set:
                key = (names[int(code)], sym)
                if key not in out:
                    out[key] = np.zeros(length, dtype=int)
                out[key] += np.bincount(which[np.ix_(rows, symbol_idx[sym])].ravel(), minlength=length)
    return {key: (bins, val[:-1]) for key, val in out.items()}


def between_species(*, coords_1, coords_2, same: bool, lattice, max_dist=5.0, resolution=0.1):
    """(x, y) of rdf.py:193-268; ``lattice`` is one 3x3 matrix or one per frame [T, 3, 3]
    (``constant_lattice=False``: every frame has its own distances and its own normalisation)."""
    coords_1 = np.asarray(coords_1, dtype=np.float64)
    coords_2 = np.asarray(coords_2, dtype=np.float64)
    n_t, n2, _ = coords_2.shape
    n1 = coords_1.shape[1]
    lattices = np.asarray(lattice, dtype=np.float64)
    if lattices.ndim == 2:
        lattices = np.array([lattices] * n_t)
    bins = np.arange(0, max_dist + resolution, resolution)
    shell = (bins + resolution) ** 3 - bins**3
    counts = np.zeros(len(bins) - 1, dtype=float)
    for t in range(n_t):
        lat = Lattice(lattices[t])
        d = lat.get_all_distances(coords_1[t], coords_2[t])
        d = d[~np.eye(n1, dtype=bool)] if same else d.ravel()
        hist, _ = np.histogram(d, bins=bins, density=False)
        norm = ((n2 * n1 / lat.volume) * (4 / 3) * np.pi * shell)[:-1]
        counts += hist / norm
    counts /= n_t
    return bins[:-1], counts
