"""Oracle for the displacement path: src/gemdat/trajectory.py:40-60, 182-193, 705-730
and src/gemdat/metrics.py:35-47, 76-97, 190-231.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""

from __future__ import annotations

import numpy as np

from .pmg import Lattice, to_displacements


def displacements(positions: np.ndarray) -> np.ndarray:
    """``Trajectory.displacements`` for a positions-mode trajectory: [T, A, 3] fractional."""
    return to_displacements(positions)


def cumulative_displacements(positions: np.ndarray) -> np.ndarray:
    """trajectory.py:705-713: sequential float64 cumsum over time."""
    return np.cumsum(displacements(positions), axis=0)


def lengths(vectors: np.ndarray, lattice_matrix) -> np.ndarray:
    """trajectory.py:40-60: sqrt(v G v^T) with the metric tensor G = M M^T."""
    g = Lattice(lattice_matrix).metric_tensor
    tmp = np.dot(vectors, g)
    return np.sqrt(np.einsum('ij,ij->i', tmp, vectors))


def distances_from_base_position(positions: np.ndarray, lattice_matrix) -> np.ndarray:
    """trajectory.py:715-730: [A, T] (note the transpose)."""
    cum = cumulative_displacements(positions)
    t, a, _ = cum.shape
    return lengths(cum.reshape(-1, 3), lattice_matrix).reshape(t, a).T


def speed(positions, lattice_matrix) -> np.ndarray:
    """metrics.py:35-47."""
    return np.diff(distances_from_base_position(positions, lattice_matrix), prepend=0)


def amplitudes(speed_arr: np.ndarray) -> np.ndarray:
    """metrics.py:209-231: sums of same-sign runs of ``speed``; the run before the first
    sign flip is merged with the one after it, and the last flip (a wrap-around
    artefact of np.roll) is ignored."""
    out = []
    for row in speed_arr:
        signs = np.sign(row)
        splits = np.where(signs != np.roll(signs, shift=-1))[0]
        for part in np.array_split(row, splits[1:-1] + 1):
            out.append(np.sum(part))
    return np.asarray(out)


def vibration_amplitude(positions, lattice_matrix) -> float:
    """metrics.py:190-207: std of the amplitudes."""
    return float(np.std(amplitudes(speed(positions, lattice_matrix))))


def tracer_diffusivity(positions, lattice_matrix, *, time_step: float, dimensions: int = 3) -> float:
    """metrics.py:76-97 (m^2/s)."""
    angstrom = 1e-10
    dist = distances_from_base_position(positions, lattice_matrix)
    msd = np.mean(dist[:, -1] ** 2)
    total_time = positions.shape[0] * time_step
    return float((msd * angstrom**2) / (2 * dimensions * total_time))
