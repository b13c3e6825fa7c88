"""Oracle for the N1 rows: src/gemdat/transitions.py:280-410, 837-879.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Plain numpy restatements of
``Transitions.occupancy`` (:280-306), ``occupancy_by_site_type`` (:308-321), ``atom_locations``
(:323-337), ``residence_time`` (:339-373), ``split`` (:375-410) and
``_split_transitions_events`` (:837-879); pinned to ``tests/golden/n1.npz`` (outputs of the
reference's own code, ``tests/golden/make_golden_n1.py``) by ``tests/test_oracle_golden.py``.
"""

from __future__ import annotations

import numpy as np

NOSITE = -1


def occupancy(states: np.ndarray, n_sites: int) -> np.ndarray:
    """:280-306: per site, the count of ``np.unique(states)`` divided by ``len(states)``; 0 for
    sites that never occur."""
    unq, counts = np.unique(states, return_counts=True)
    share = dict(zip(unq.tolist(), (counts / len(states)).tolist()))
    return np.array([share.get(i, 0) for i in range(n_sites)], dtype=float)


def by_label(occ: np.ndarray, labels: list[str], divide_by_count: bool, n_floating: int = 1) -> dict:
    """:308-337: python ``sum`` of the occupancies per label, in site order, over the number of
    sites of the label (``occupancy_by_site_type``) or over ``n_floating`` (``atom_locations``)."""
    groups: dict[str, list[float]] = {}
    for lab, v in zip(labels, occ.tolist()):
        groups.setdefault(lab, []).append(v)
    return {k: sum(v) / (len(v) if divide_by_count else n_floating) for k, v in groups.items()}


def residence_time(states: np.ndarray) -> np.ndarray:
    """:339-373: rows (atom, site, time) of every visit between two state changes of an atom,
    visits of NOSITE skipped; int64 [R, 3]."""
    rows = []
    for atom in range(states.shape[1]):
        col = states[:, atom]
        boundaries = np.nonzero(np.diff(col))[0] + 1
        for start, end in zip(boundaries[:-1], boundaries[1:]):
            if col[start] != NOSITE:
                rows.append((atom, col[start], end - start))
    return np.array(rows, dtype=np.int64).reshape(-1, 3)


def split_bounds(n_states: int, n_parts: int):
    """Frame bounds of ``np.array_split`` (states) and the time bins of the event split."""
    small, big = divmod(n_states, n_parts)
    sizes = [small + 1] * big + [small] * (n_parts - big)
    state_bounds = np.concatenate([[0], np.cumsum(sizes)])
    bins = np.linspace(0, n_states + 1, n_parts + 1, dtype=int)
    return state_bounds, bins


def split_events(events: np.ndarray, n_states: int, n_parts: int):
    """:837-879 on an int array [E, 6] (time in column 5): list of (rows, source row numbers)."""
    if len(events) < n_parts:
        raise ValueError(f'Not enough transitions per part to split into {n_parts}')
    _, bins = split_bounds(n_states, n_parts)
    out = []
    for start, stop in zip(bins[:-1], bins[1:]):
        keep = np.nonzero((events[:, 5] >= start) & (events[:, 5] < stop))[0]
        part = events[keep].copy()
        part[:, 5] -= start
        out.append((part, keep))
    return out
