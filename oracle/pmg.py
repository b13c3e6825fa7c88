"""Restatement of the pymatgen 2024.1.27 pieces the hot path reaches  [recalled].

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

* ``Lattice``: ``matrix``, ``lengths``, ``angles``, ``parameters``, ``metric_tensor``,
  ``volume``, ``get_cartesian_coords`` (= ``frac @ matrix``), ``get_all_distances``
  (minimum-image distance matrix, pymatgen ``pbc_shortest_vectors``: map both
  fractional sets into the LLL-reduced basis, wrap to [0,1) there, try the 27
  neighbouring images of that basis, keep the shortest), ``_calculate_lll``.
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

    # ---- LLL reduction (pymatgen Lattice._calculate_lll, delta = 0.75)  [recalled]
    def _calculate_lll(self, delta: float = 0.75):
        a = self._matrix.copy().T  # columns are the basis vectors
        b = np.zeros((3, 3))  # Gram-Schmidt vectors
        u = np.zeros((3, 3))  # Gram-Schmidt coefficients
        m = np.zeros(3)  # squared norms of the Gram-Schmidt vectors
        b[:, 0] = a[:, 0]
        m[0] = np.dot(b[:, 0], b[:, 0])
        for i in range(1, 3):
            u[i, 0:i] = np.dot(a[:, i].T, b[:, 0:i]) / m[0:i]
            b[:, i] = a[:, i] - np.dot(b[:, 0:i], u[i, 0:i].T)
            m[i] = np.dot(b[:, i], b[:, i])
        k = 2
        mapping = np.identity(3, dtype=np.double)
        while k <= 3:
            for i in range(k - 1, 0, -1):  # size reduction
                q = round(u[k - 1, i - 1])
                if q != 0:
                    a[:, k - 1] = a[:, k - 1] - q * a[:, i - 1]
                    mapping[:, k - 1] = mapping[:, k - 1] - q * mapping[:, i - 1]
                    uu = list(u[i - 1, 0:(i - 1)])
                    uu.append(1)
                    u[k - 1, 0:i] = u[k - 1, 0:i] - q * np.array(uu)
            # Lovasz condition
            if np.dot(b[:, k - 1], b[:, k - 1]) >= (delta - abs(u[k - 1, k - 2]) ** 2) * np.dot(
                    b[:, (k - 2)], b[:, (k - 2)]):
                k += 1
            else:
                v = a[:, k - 1].copy()
                a[:, k - 1] = a[:, k - 2].copy()
                a[:, k - 2] = v
                v_m = mapping[:, k - 1].copy()
                mapping[:, k - 1] = mapping[:, k - 2].copy()
                mapping[:, k - 2] = v_m
                for s in range(k - 1, k + 1):
                    u[s - 1, 0:(s - 1)] = np.dot(a[:, s - 1].T, b[:, 0:(s - 1)]) / m[0:(s - 1)]
                    b[:, s - 1] = a[:, s - 1] - np.dot(b[:, 0:(s - 1)], u[s - 1, 0:(s - 1)].T)
                    m[s - 1] = np.dot(b[:, s - 1], b[:, s - 1])
                if k > 2:
                    k -= 1
                else:
                    p = np.dot(a[:, k:3].T, b[:, (k - 2):k])
                    q = np.diag(m[(k - 2):k])
                    u[k:3, (k - 2):k] = np.linalg.lstsq(q.T, p.T, rcond=None)[0].T
        return a.T, mapping.T

    @property
    def lll_matrix(self) -> np.ndarray:
        return self._calculate_lll()[0]

    @property
    def lll_mapping(self) -> np.ndarray:
        return self._calculate_lll()[1]

    @property
    def lll_inverse(self) -> np.ndarray:
        return np.linalg.inv(self.lll_mapping)

    def get_lll_frac_coords(self, frac) -> np.ndarray:
        return np.dot(frac, self.lll_inverse)

    def get_all_distances(self, f1, f2) -> np.ndarray:
        """pymatgen ``pbc_shortest_vectors`` (coord_cython.pyx)  [recalled]: both sets go to the
        LLL-reduced basis and are wrapped to [0,1) there; ``cart = frac_lll @ lll_matrix`` per
        point (accumulated term by term from 0); ``pre = cart2 - cart1``; the answer is the
        smallest ``|pre + image @ lll_matrix|^2`` over the 27 images, summed
        ``(x^2 + y^2) + z^2``, then one square root."""
        f1 = np.atleast_2d(np.asarray(f1, dtype=np.float64))
        f2 = np.atleast_2d(np.asarray(f2, dtype=np.float64))
        lll_matrix, mapping = self._calculate_lll()
        # the mapping is an integer unimodular matrix: round the float inverse to its exact value
        inv = np.rint(np.linalg.inv(mapping)) + 0.0

        def dot3(a, m):  # ((0 + a0 m0k) + a1 m1k) + a2 m2k, no fused multiply-add
            out = np.zeros(a.shape[:-1] + (3,))
            for i in range(3):
                out = out + a[..., i:i + 1] * m[i][None, :]
            return out

        def to_cart(f):
            w = (f[:, 0:1] * inv[0][None, :] + f[:, 1:2] * inv[1][None, :]) + f[:, 2:3] * inv[2][None, :]
            w = w - np.floor(w)
            return dot3(w, lll_matrix)

        images = np.array(
            [[i, j, k] for i in (-1, 0, 1) for j in (-1, 0, 1) for k in (-1, 0, 1)], dtype=np.float64)
        cart_im = dot3(images, lll_matrix)
        pre = to_cart(f2)[None, :, :] - to_cart(f1)[:, None, :]
        best = np.full((len(f1), len(f2)), np.inf)
        for im in cart_im:
            d = pre + im[None, None, :]
            d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
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
