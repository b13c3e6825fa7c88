"""Collective jumps, CPU restatement (TEST INFRASTRUCTURE: only tests/, smoke() and the
cpu_baseline leg of bench.py may import this).

Follows reference src/gemdat/collective.py:61-106 (``Collective._compute``) on plain
integer arrays instead of pandas rows:

* jumps sorted by ('stop time', 'start time'), stable on the incoming order (:68);
* for every i, the partners j > i are walked in sorted order; the walk of i STOPS at the
  first j with ``start_j - stop_i > max_steps`` (:76-77, a ``break`` although start
  times are not monotonic in this order; kept as the reference has it), and a j with
  ``start_i - stop_j > max_steps`` (:78-79) or on the same atom (:80-81) is skipped;
* a pair is collective when any of the four distances {start_i, dest_i} x
  {start_j, dest_j} under the minimum-image convention is ``< max_dist`` (:83-88);
* ``n_solo_jumps`` = jumps that are in no pair (:100-101).

Parity pinned by tests/golden/collective.npz (made by executing the reference's
collective.py, tests/golden/make_golden_collective.py).
"""

from __future__ import annotations

import numpy as np

from .pmg import Lattice

# column order of the jump table (reference jumps.py:110-119)
ATOM, START_SITE, DEST_SITE, START_TIME, STOP_TIME = 0, 1, 2, 3, 4


def sort_jumps(jumps: np.ndarray) -> np.ndarray:
    """Rows ordered like ``sort_values(['stop time', 'start time'])`` (collective.py:68).

    pandas sorts several keys with a lexsort, which is stable; np.lexsort is too."""
    jumps = np.asarray(jumps)
    order = np.lexsort((jumps[:, START_TIME], jumps[:, STOP_TIME]))
    return jumps[order]


def close_matrix(site_frac: np.ndarray, lattice: np.ndarray, max_dist: float) -> np.ndarray:
    """[n_sites, n_sites] bool: ``lattice.get_all_distances(...) < max_dist`` (collective.py:86-88)."""
    dist = Lattice(np.asarray(lattice, dtype=np.float64)).get_all_distances(site_frac, site_frac)
    return dist < max_dist


def collective_pairs(jumps: np.ndarray, site_frac: np.ndarray, lattice: np.ndarray, max_steps: int,
                     max_dist: float, close: np.ndarray | None = None):
    """Returns (sorted_jumps, pairs[P, 2] of row numbers in the sorted table, in emission order)."""
    ev = sort_jumps(jumps)
    n = len(ev)
    if close is None:
        close = close_matrix(site_frac, lattice, max_dist)
    start = ev[:, START_TIME].astype(np.int64)
    stop = ev[:, STOP_TIME].astype(np.int64)
    pairs = []
    for i in range(n - 1):
        for j in range(i + 1, n):
            if start[j] - stop[i] > max_steps:
                break
            if start[i] - stop[j] > max_steps:
                continue
            if ev[i, ATOM] == ev[j, ATOM]:
                continue
            a = (ev[i, START_SITE], ev[i, DEST_SITE])
            b = (ev[j, START_SITE], ev[j, DEST_SITE])
            if close[a[0], b[0]] or close[a[0], b[1]] or close[a[1], b[0]] or close[a[1], b[1]]:
                pairs.append((i, j))
    return ev, np.array(pairs, dtype=np.int64).reshape(-1, 2)


def solo_count(n_jumps: int, pairs: np.ndarray) -> int:
    """collective.py:100-101."""
    hit = np.zeros(n_jumps, dtype=bool)
    hit[pairs.reshape(-1)] = True
    return int(n_jumps - hit.sum())


def coll_jumps(ev: np.ndarray, pairs: np.ndarray) -> np.ndarray:
    """[P, 2, 2] ((start_i, dest_i), (start_j, dest_j)) like collective.py:90-96."""
    out = np.empty((len(pairs), 2, 2), dtype=np.int64)
    out[:, 0, 0] = ev[pairs[:, 0], START_SITE]
    out[:, 0, 1] = ev[pairs[:, 0], DEST_SITE]
    out[:, 1, 0] = ev[pairs[:, 1], START_SITE]
    out[:, 1, 1] = ev[pairs[:, 1], DEST_SITE]
    return out


def site_pair_count_matrix(cj: np.ndarray, labels, site_pairs) -> np.ndarray:
    """collective.py:108-133 with the pair list given (the reference builds it from a set,
    whose order is not defined across processes)."""
    index = {p: k for k, p in enumerate(site_pairs)}
    m = np.zeros((len(site_pairs), len(site_pairs)), dtype=int)
    for (si, di), (sj, dj) in cj:
        m[index[(labels[si], labels[di])], index[(labels[sj], labels[dj])]] += 1
    return m


def multiple_collective(cj: np.ndarray):
    """collective.py:140-166: unique (sorted-within-pair) site pairs and their counts."""
    collective = np.array(dtype=[('start', int), ('stop', int)],
                          object=[[(a[0], a[1]), (b[0], b[1])] for a, b in cj.tolist()])
    if len(collective) == 0:
        collective = collective.reshape(0, 2)
    collective = np.sort(collective, axis=1)
    return np.unique(collective, axis=0, return_counts=True)
