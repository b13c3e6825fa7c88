"""Light stand-ins for the pymatgen objects on the hot path's boundary.

GEMDAT hands the path a pymatgen ``Lattice`` (via ``Trajectory.get_lattice``) and a
pymatgen ``Structure`` of candidate sites (reference src/gemdat/io.py:61-96).  pymatgen is
not a dependency here; these classes expose the attributes the path reads
(``frac_coords``, ``labels``, ``is_ordered``, ``lattice``; ``matrix``, ``lengths``,
``angles``, ``parameters``, ``metric_tensor``, ``get_cartesian_coords``,
``get_all_distances``) and real pymatgen objects are accepted wherever these are
(duck typing, see ``as_lattice_matrix`` / ``site_arrays``).
"""

from __future__ import annotations

import ctypes as C
import re

import numpy as np

from . import _device as dv
from . import _lib


class FloatWithUnit(float):
    """``pymatgen.core.units.FloatWithUnit`` look-alike: a float that remembers its unit."""

    def __new__(cls, val, unit=None):
        obj = float.__new__(cls, val)
        obj.unit = unit
        return obj

    def __repr__(self):
        return f'{float(self)!r} {self.unit}'


def as_lattice_matrix(lattice) -> np.ndarray:
    m = getattr(lattice, 'matrix', lattice)
    return np.ascontiguousarray(np.asarray(m, dtype=np.float64).reshape(3, 3))


class Lattice:
    def __init__(self, matrix):
        self._matrix = as_lattice_matrix(matrix)

    @property
    def matrix(self) -> np.ndarray:
        return self._matrix

    @property
    def lengths(self) -> tuple[float, float, float]:
        return tuple(np.sqrt(np.sum(self._matrix**2, axis=1)).tolist())

    abc = lengths

    @property
    def angles(self) -> tuple[float, float, float]:
        m, lengths = self._matrix, self.lengths
        out = []
        for i in range(3):
            j, k = (i + 1) % 3, (i + 2) % 3
            c = float(np.dot(m[j], m[k]) / (lengths[j] * lengths[k]))
            out.append(float(np.degrees(np.arccos(min(1.0, max(-1.0, c))))))
        return tuple(out)

    @property
    def parameters(self):
        return (*self.lengths, *self.angles)

    @property
    def metric_tensor(self) -> np.ndarray:
        return self._matrix @ self._matrix.T

    @property
    def volume(self) -> float:
        return float(abs(np.linalg.det(self._matrix)))

    def get_cartesian_coords(self, frac) -> np.ndarray:
        return np.dot(np.asarray(frac, dtype=np.float64), self._matrix)

    def get_fractional_coords(self, cart) -> np.ndarray:
        return np.dot(np.asarray(cart, dtype=np.float64), np.linalg.inv(self._matrix))

    def get_all_distances(self, fcoords1, fcoords2) -> np.ndarray:
        """Minimum-image distance matrix (device kernel ``gd_pbc_distances``)."""
        f1 = np.atleast_2d(dv.host_f64(fcoords1))
        f2 = np.atleast_2d(dv.host_f64(fcoords2))
        d1, d2 = dv.to_device(f1), dv.to_device(f2)
        out = dv.empty((len(f1), len(f2)), dtype=d1.dtype)
        _lib.check(
            _lib.lib.gd_pbc_distances(
                self._matrix.ctypes.data, dv.ptr(d1), len(f1), dv.ptr(d2), len(f2), dv.ptr(out),
                dv.stream_ptr(),
            )
        )
        return dv.to_host(out)

    def __eq__(self, other):
        return hasattr(other, 'matrix') and np.allclose(self._matrix, as_lattice_matrix(other))

    def __repr__(self):
        return f'Lattice(abc={self.lengths}, angles={self.angles})'


class Site:
    def __init__(self, symbol: str, frac_coords, label: str, occupancy: float = 1.0):
        self.symbol = symbol
        self.frac_coords = np.asarray(frac_coords, dtype=np.float64)
        self.label = label
        self.occupancy = occupancy

    @property
    def is_ordered(self) -> bool:
        return self.occupancy == 1.0


class Sites:
    """Candidate sites: the subset of ``pymatgen.core.Structure`` the path reads."""

    def __init__(self, lattice, species, frac_coords, labels=None, occupancies=None):
        self.lattice = lattice if isinstance(lattice, Lattice) else Lattice(lattice)
        self._frac = np.ascontiguousarray(np.asarray(frac_coords, dtype=np.float64).reshape(-1, 3))
        species = [species] * len(self._frac) if isinstance(species, str) else list(species)
        self.symbols = [str(getattr(s, 'symbol', s)) for s in species]
        self.labels = list(labels) if labels is not None else list(self.symbols)
        self.occupancies = (
            np.ones(len(self._frac)) if occupancies is None else np.asarray(occupancies, dtype=float)
        )

    @property
    def frac_coords(self) -> np.ndarray:
        return self._frac

    @property
    def cart_coords(self) -> np.ndarray:
        return self.lattice.get_cartesian_coords(self._frac)

    @property
    def is_ordered(self) -> bool:
        return bool(np.all(self.occupancies == 1.0))

    def copy(self) -> 'Sites':
        return Sites(self.lattice, list(self.symbols), self._frac.copy(), list(self.labels),
                     self.occupancies.copy())

    def __len__(self):
        return len(self._frac)

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    def __getitem__(self, i):
        return Site(self.symbols[i], self._frac[i], self.labels[i], float(self.occupancies[i]))

    def make_supercell(self, scaling) -> 'Sites':
        """Diagonal supercell; labels are shared between copies (io.py:88-94)."""
        na, nb, nc = (int(s) for s in scaling)
        shifts = np.array([[i, j, k] for i in range(na) for j in range(nb) for k in range(nc)], float)
        frac = (self._frac[:, None, :] + shifts[None, :, :]).reshape(-1, 3) / np.array([na, nb, nc])
        reps = len(shifts)
        self.lattice = Lattice(self.lattice.matrix * np.array([[na], [nb], [nc]]))
        self._frac = np.ascontiguousarray(frac)
        self.symbols = [s for s in self.symbols for _ in range(reps)]
        self.labels = [re.sub(r'_\d+$', '', lab) for lab in self.labels for _ in range(reps)]
        self.occupancies = np.repeat(self.occupancies, reps)
        return self


def site_arrays(sites):
    """``(frac_coords float64 [S,3], labels list[str], is_ordered)`` of a Sites / pymatgen Structure."""
    frac = np.ascontiguousarray(np.asarray(sites.frac_coords, dtype=np.float64).reshape(-1, 3))
    labels = getattr(sites, 'labels', None)
    if labels is None:
        labels = [site.label for site in sites]
    return frac, [str(x) for x in labels], bool(getattr(sites, 'is_ordered', True))
