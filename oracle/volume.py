"""Oracle for the density histogram: src/gemdat/volume.py:449-509.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""

from __future__ import annotations

import numpy as np

from .pmg import Lattice


def grid_shape(lattice_matrix, resolution: float) -> tuple[int, int, int]:
    """Number of *edges* per axis: n = int(1 + length // resolution) (volume.py:482-484)."""
    lengths = Lattice(lattice_matrix).lengths
    return tuple(int(1 + length // resolution) for length in lengths)


def trajectory_to_volume(positions: np.ndarray, lattice_matrix, resolution: float = 0.2) -> np.ndarray:
    coords = np.asarray(positions, dtype=np.float64).reshape(-1, 3)
    assert coords.min() >= 0
    assert coords.max() < 1
    nx, ny, nz = grid_shape(lattice_matrix, resolution)
    digitized = np.vstack(
        [
            np.digitize(coords[:, 0], bins=np.linspace(0, 1, nx)[1:]),
            np.digitize(coords[:, 1], bins=np.linspace(0, 1, ny)[1:]),
            np.digitize(coords[:, 2], bins=np.linspace(0, 1, nz)[1:]),
        ]
    ).T
    idx, counts = np.unique(digitized, return_counts=True, axis=0)
    data = np.zeros((nx - 1, ny - 1, nz - 1), dtype=int)
    data[idx[:, 0], idx[:, 1], idx[:, 2]] = counts
    return data
