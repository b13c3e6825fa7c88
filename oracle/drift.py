"""Oracle for the per-frame averages: src/gemdat/trajectory.py:732-830.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Plain numpy restatements of
``Trajectory.drift`` (:751-791), ``apply_drift_correction`` (:793-830, with the ``to_positions``
of :121-127 that turns the corrected steps back into positions) and ``center_of_mass``
(:732-749); pinned to ``tests/golden/drift.npz`` (outputs of the reference's own code,
``tests/golden/make_golden_drift.py``) by ``tests/test_oracle_golden.py``.
"""

from __future__ import annotations

import numpy as np


def displacements(positions: np.ndarray) -> np.ndarray:
    """pymatgen ``to_displacements`` [recalled]: wrapped frame-to-frame steps, frame 0 = 0."""
    d = positions - np.roll(positions, 1, axis=0)
    d[0] = 0.0
    return d - np.round(d)


def drift(positions: np.ndarray, selected) -> np.ndarray:
    """:751-791: mean step of the selected atoms per frame, [T, 1, 3]."""
    return np.mean(displacements(positions[:, selected]), axis=1)[:, None, :]


def corrected_steps(positions: np.ndarray, selected) -> np.ndarray:
    """:793-830: ``displacements - drift`` (the coords of the returned trajectory)."""
    return displacements(positions) - drift(positions, selected)


def corrected_positions(positions: np.ndarray, selected) -> np.ndarray:
    """:121-127 on the result: ``np.mod(base + np.cumsum(steps, axis=0), 1)``."""
    return np.mod(positions[0] + np.cumsum(corrected_steps(positions, selected), axis=0), 1)


def center_of_mass(positions: np.ndarray, masses) -> np.ndarray:
    """:732-749: mass-weighted mean of base + cumulative displacements, [T, 1, 3] (not wrapped)."""
    unwrapped = positions[0] + np.cumsum(displacements(positions), axis=0)
    return np.average(unwrapped, axis=1, weights=masses).reshape(-1, 1, 3)
