"""Oracle for the all-lag MSD: src/gemdat/trajectory.py:896-933.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

``msd_fft`` restates the reference's FFT formulation (S1 - 2*S2).  ``msd_direct`` is
the definition itself, summed per lag with ``math.fsum``-grade accuracy (numpy
longdouble accumulation), used as the accuracy yardstick: the FFT formulation
subtracts two O(<r^2>) numbers and is itself only good to ~1e-9..1e-10 relative at
small lags.
"""

from __future__ import annotations

import numpy as np

from .displ import cumulative_displacements
from .pmg import Lattice


def unwrapped_cartesian(positions, lattice_matrix) -> np.ndarray:
    """[A, T, 3] Cartesian cumulative displacement."""
    r = Lattice(lattice_matrix).get_cartesian_coords(cumulative_displacements(positions))
    return np.transpose(r, (1, 0, 2))


def msd_fft(positions, lattice_matrix) -> np.ndarray:
    pos = unwrapped_cartesian(positions, lattice_matrix)
    n = pos.shape[1]
    lagcount = n - np.arange(n)[None, :]
    power = np.abs(np.fft.fft(pos, n=2 * n, axis=-2)) ** 2
    s2 = np.fft.ifft(power, axis=-2)[:, :n, :].real.sum(axis=-1) / lagcount
    d = np.square(pos).sum(axis=-1)
    d = np.append(d, np.zeros((pos.shape[0], 1)), axis=-1)
    twice = 2 * np.sum(d, axis=-1)[:, None]
    running = np.cumsum(np.insert(d[:, 0:-1], 0, 0, axis=-1) + np.flip(d, axis=-1), axis=-1)
    s1 = (twice - running)[:, :-1] / lagcount
    return s1 - 2 * s2


def msd_direct(positions, lattice_matrix, lags=None) -> np.ndarray:
    """msd[a, k] = mean_t |r[a, t+k] - r[a, t]|^2, long-double accumulation."""
    pos = unwrapped_cartesian(positions, lattice_matrix)
    n_atoms, n, _ = pos.shape
    if lags is None:
        lags = np.arange(n)
    out = np.zeros((n_atoms, len(lags)))
    for j, k in enumerate(lags):
        if k == 0:
            continue
        diff = (pos[:, k:, :] - pos[:, : n - k, :]).astype(np.longdouble)
        out[:, j] = (np.sum(diff * diff, axis=(1, 2), dtype=np.longdouble) / (n - k)).astype(np.float64)
    return out
