"""Oracle for transition events: src/gemdat/transitions.py:413-482.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Returns a plain int64 array ``[E, 6]`` with the reference's column order
``atom index, start site, destination site, start inner site,
destination inner site, time`` (transitions.py:471-481), rows sorted by atom then
time.  Quirks kept: an atom whose outer state never changes is
skipped even if its inner state changes; ``IndexError`` when the outer state
changes but the inner state is constant; ``ValueError`` when no atom changes state.
"""

from __future__ import annotations

import numpy as np

EVENT_COLUMNS = [
    'atom index',
    'start site',
    'destination site',
    'start inner site',
    'destination inner site',
    'time',
]


def calculate_transition_events(atom_sites: np.ndarray, atom_inner_sites: np.ndarray) -> np.ndarray:
    atom_sites = np.asarray(atom_sites)
    atom_inner_sites = np.asarray(atom_inner_sites)
    n_frames = atom_sites.shape[0]
    blocks = []
    for atom in range(atom_sites.shape[1]):
        outer = atom_sites[:, atom]
        inner = atom_inner_sites[:, atom]
        # t where state[t] != state[(t+1) % T]  (np.roll(-1) wraps around)
        t_outer = np.nonzero(outer != np.concatenate([outer[1:], outer[:1]]))[0]
        if t_outer.size < 1:
            continue
        t_inner = np.nonzero(inner != np.concatenate([inner[1:], inner[:1]]))[0]
        if t_outer[-1] == n_frames - 1:
            t_outer = t_outer[:-1]
        if t_inner.size == 0:
            raise IndexError('index -1 is out of bounds for axis 0 with size 0')
        if t_inner[-1] == n_frames - 1:
            t_inner = t_inner[:-1]
        time = np.union1d(t_outer, t_inner)
        rows = np.empty((time.size, 6), dtype=np.int64)
        rows[:, 0] = atom
        rows[:, 1] = outer[time]
        rows[:, 2] = outer[time + 1]
        rows[:, 3] = inner[time]
        rows[:, 4] = inner[time + 1]
        rows[:, 5] = time
        blocks.append(rows)
    if not blocks:
        raise ValueError('need at least one array to concatenate')
    return np.vstack(blocks)
