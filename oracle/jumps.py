"""Oracle for transitions -> jumps: src/gemdat/jumps.py:24-119, 185-221 and
src/gemdat/transitions.py:812-834.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference walks a pandas DataFrame with ``iterrows``; here the same per-atom
sequential state machine runs over the int64 event table.  Output
``[J, 6]`` int64: ``atom index, start site, destination site, start time, stop time,
residence_time`` (jumps.py:110-119).
"""

from __future__ import annotations

import numpy as np

JUMP_COLUMNS = [
    'atom index',
    'start site',
    'destination site',
    'start time',
    'stop time',
    'residence_time',
]

# event columns
_ATOM, _S0, _S1, _I0, _I1, _T = range(6)


def transitions_to_jumps(events: np.ndarray, n_frames: int, minimal_residence: int = 0) -> np.ndarray:
    """``n_frames`` is ``len(transitions.trajectory)`` (jumps.py:39)."""
    events = np.asarray(events, dtype=np.int64)
    jumps: list[tuple] = []
    # groupby('atom index') iterates atoms in ascending order, rows keep table order
    order = np.argsort(events[:, _ATOM], kind='stable')
    ev = events[order]
    bounds = np.nonzero(np.diff(ev[:, _ATOM]))[0] + 1
    for chunk in np.split(ev, bounds):
        if len(chunk) == 0:
            continue
        have_from = False
        from_start_site = from_dest_site = from_start_time = 0
        cand = None  # [atom, start site, dest site, dest inner, start time, stop time]
        for atom, s0, s1, i0, i1, t in chunk.tolist():
            stop = t + 1
            if cand is not None:
                if s1 == cand[2]:
                    if i1 != -1:
                        cand[3] = i1
                else:
                    residence = t - (cand[5] - 1)
                    if residence >= minimal_residence and cand[3] != -1:
                        jumps.append((cand[0], cand[1], cand[2], cand[4], cand[5], residence))
                    cand = None
            if s0 != -1 and s0 != s1:
                have_from = True
                from_start_site, from_dest_site, from_start_time = s0, s1, t
            if have_from:
                if s1 == from_start_site:
                    have_from = False
                    cand = None
                elif i1 != -1:
                    cand = [atom, from_start_site, s1, i1, from_start_time, stop]
                    have_from = False
                elif s1 != from_dest_site:
                    cand = [atom, from_start_site, s1, i1, from_start_time, stop]
                    have_from = False
        if cand is not None:
            residence = n_frames - (cand[5] - 1)
            if residence >= minimal_residence and cand[3] != -1:
                jumps.append((cand[0], cand[1], cand[2], cand[4], cand[5], residence))
    if len(jumps) == 0:
        raise ValueError('No jumps found')
    out = np.array(jumps, dtype=np.int64).reshape(-1, 6)
    return out[out[:, 1] != out[:, 2]]


def transitions_matrix(start_dest: np.ndarray, n_sites: int) -> np.ndarray:
    """Count matrix of (start site, destination site) rows; negative indices wrap as numpy does."""
    m = np.zeros((n_sites, n_sites), dtype=int)
    idx, counts = np.unique(np.asarray(start_dest).reshape(-1, 2), return_counts=True, axis=0)
    m[idx[:, 0], idx[:, 1]] = counts
    return m


def jump_diffusivity(*, pdist: np.ndarray, matrix: np.ndarray, dimensions: int, n_floating: int, total_time: float) -> float:
    """jumps.py:185-210 (m^2/s); ``scipy.constants.angstrom`` = 1e-10."""
    angstrom = 1e-10
    jump_diff = np.sum(pdist**2 * matrix)
    jump_diff *= angstrom**2 / (2 * dimensions * n_floating * total_time)
    return float(jump_diff)


# ---------------------------------------------------------------------------------------------
# The same machine with its state made explicit, for time slices of an atom's event list.
# Restates what the frame-sharded K3 protocol (include/gemdat_b200.h, gd_jumps_count_shard)
# relies on: the state after an event list is a function of (entry state, events), and after
# a "departure" row (start site != -1 and != destination site) it no longer depends on the
# entry state when the rows are state-consistent.

EMPTY_STATE = (None, None)  # (pending fromevent (start site, dest site, time), pending candidate)


def jump_machine_slice(rows, minimal_residence: int = 0, entry=EMPTY_STATE):
    """Runs jumps.py:45-95 over ``rows`` (one atom, time ordered) from ``entry``.

    Returns ``(jumps, exit_state, has_departure)``; jumps still include start == destination
    rows (jumps.py:113 filters them at the very end) -- ``finish_slices`` applies that filter.
    """
    frm, cand = entry
    cand = None if cand is None else list(cand)
    jumps, has_departure = [], False
    for atom, s0, s1, i0, i1, t in np.asarray(rows, dtype=np.int64).reshape(-1, 6).tolist():
        if cand is not None:
            if s1 == cand[2]:
                if i1 != -1:
                    cand[3] = i1
            else:
                residence = t - (cand[5] - 1)
                if residence >= minimal_residence and cand[3] != -1:
                    jumps.append((cand[0], cand[1], cand[2], cand[4], cand[5], residence))
                cand = None
        if s0 != -1 and s0 != s1:
            frm = (s0, s1, t)
            has_departure = True
        if frm is not None:
            if s1 == frm[0]:
                frm, cand = None, None
            elif i1 != -1 or s1 != frm[1]:
                cand = [atom, frm[0], s1, i1, frm[2], t + 1]
                frm = None
    return jumps, (frm, None if cand is None else tuple(cand)), has_departure


def close_pending(state, n_frames: int, minimal_residence: int = 0):
    """jumps.py:97-105: the candidate still pending after an atom's last event."""
    cand = state[1]
    if cand is None:
        return []
    residence = n_frames - (cand[5] - 1)
    if residence >= minimal_residence and cand[3] != -1:
        return [(cand[0], cand[1], cand[2], cand[4], cand[5], residence)]
    return []


def finish_slices(jumps) -> np.ndarray:
    out = np.array(jumps, dtype=np.int64).reshape(-1, 6)
    return out[out[:, 1] != out[:, 2]]
