"""Restatement of the pymatgen 2024.1.27 pieces the hot path reaches  [recalled].

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

* ``Lattice``: ``matrix``, ``lengths``, ``angles``, ``parameters``, ``metric_tensor``,
  ``volume``, ``get_cartesian_coords`` (= ``frac @ matrix``), ``get_all_distances``
  (minimum-image distance matrix, pymatgen ``pbc_shortest_vectors``: wrap both
  fractional sets to [0,1), try the 27 neighbouring images, keep the shortest).
* ``to_displacements`` / ``to_positions`` of ``pymatgen.core.trajectory.Trajectory``:
  ``d[t] = x[t]-x[t-1]`` with ``d[0]=0``, wrapped by ``d - np.around(d)``; the inverse is
  ``base + cumsum(d)``.  GEMDAT then applies ``mod 1`` (src/gemdat/trajectory.py:121-127).
"""

from __future__ import annotations

import numpy as np


class Lattice:
    def __init__(self, matrix):
        self._matrix = np.array(matrix, dtype=np.float64).reshape(3, 3)

    @property
    def matrix(self) -> np.ndarray:
        return self._matrix

    @property
    def lengths(self):
        return tuple(np.sqrt(np.sum(self._matrix**2, axis=1)).tolist())

    @property
    def abc(self):
        return self.lengths

    @property
    def angles(self):
        m = self._matrix
        lengths = self.lengths
        ang = np.zeros(3)
        for i in range(3):
            j = (i + 1) % 3
            k = (i + 2) % 3
            c = np.dot(m[j], m[k]) / (lengths[j] * lengths[k])
            ang[i] = max(min(c, 1.0), -1.0)
        ang = np.arccos(ang) * 180.0 / np.pi
        return tuple(ang.tolist())

    @property
    def parameters(self):
        return (*self.lengths, *self.angles)

    @property
    def metric_tensor(self) -> np.ndarray:
        return np.dot(self._matrix, self._matrix.T)

    @property
    def volume(self) -> float:
        m = self._matrix
        return float(abs(np.dot(np.cross(m[0], m[1]), m[2])))

    def get_cartesian_coords(self, frac) -> np.ndarray:
        return np.dot(np.asarray(frac, dtype=np.float64), self._matrix)

    def get_fractional_coords(self, cart) -> np.ndarray:
        return np.dot(np.asarray(cart, dtype=np.float64), np.linalg.inv(self._matrix))

    def get_all_distances(self, f1, f2) -> np.ndarray:
        f1 = np.atleast_2d(np.asarray(f1, dtype=np.float64))
        f2 = np.atleast_2d(np.asarray(f2, dtype=np.float64))
        f1 = f1 - np.floor(f1)
        f2 = f2 - np.floor(f2)
        images = np.array(
            [[i, j, k] for i in (-1, 0, 1) for j in (-1, 0, 1) for k in (-1, 0, 1)],
            dtype=np.float64,
        )
        best = np.full((len(f1), len(f2)), np.inf)
        for img in images:
            v = (f2[None, :, :] + img[None, None, :]) - f1[:, None, :]
            cart = np.dot(v, self._matrix)
            d2 = np.sum(cart * cart, axis=-1)
            np.minimum(best, d2, out=best)
        return np.sqrt(best)

    def __eq__(self, other):
        return isinstance(other, Lattice) and np.allclose(self._matrix, other._matrix)


def to_displacements(positions: np.ndarray) -> np.ndarray:
    """Wrapped frame-to-frame fractional displacement, frame 0 = 0."""
    pos = np.asarray(positions, dtype=np.float64)
    d = np.subtract(pos, np.roll(pos, 1, axis=0))
    d[0] = 0.0
    return np.subtract(d, np.around(d))


def to_positions(displacements: np.ndarray, base_positions: np.ndarray) -> np.ndarray:
    """Inverse of ``to_displacements`` followed by GEMDAT's ``mod 1``."""
    pos = base_positions + np.cumsum(displacements, axis=0)
    return np.mod(pos, 1)
