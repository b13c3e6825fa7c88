"""Density volume of a trajectory (K5).

Mirror of ``trajectory_to_volume`` (reference src/gemdat/volume.py:449-509).  The
``Volume`` returned here is the plain container (data / lattice / label) -- peak
finding, watershed and free-energy methods of the reference's ``Volume`` operate on the
OUTPUT of this path and are out of scope.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .lattice import Lattice


@dataclass
class Volume:
    data: np.ndarray
    lattice: Lattice
    label: str = 'volume'

    @property
    def dims(self) -> tuple[int, int, int]:
        return self.data.shape

    @property
    def resolution(self) -> np.ndarray:
        return np.array(self.lattice.lengths) / np.array(self.dims)


def grid_edges(lattice: Lattice, resolution: float) -> np.ndarray:
    """Number of bin edges per axis, ``int(1 + length // resolution)`` (volume.py:482-484)."""
    return np.array([int(1 + length // resolution) for length in lattice.lengths], dtype=np.int32)


def density_device(pos: torch.Tensor, nedges: np.ndarray, grid: torch.Tensor | None = None):
    """Accumulate ``pos`` [.., 3] into a uint32 voxel grid on the device; returns (grid, flag)."""
    shape = tuple(int(n) - 1 for n in nedges)
    if grid is None:
        grid = dv.zeros(shape, torch.int32)  # reinterpreted as uint32 by the kernel
    flag = dv.zeros((1,), torch.int32)
    nedges = np.ascontiguousarray(nedges, dtype=np.int32)
    if pos.dim() == 3:  # [frames, atoms, 3]: run-length strip walker
        _lib.check(_lib.lib.gd_density_frames(dv.ptr(pos), int(pos.shape[0]), int(pos.shape[1]),
                                              nedges.ctypes.data, dv.ptr(grid), dv.ptr(flag), dv.stream_ptr()))
    else:
        _lib.check(_lib.lib.gd_density(dv.ptr(pos), pos.numel() // 3, nedges.ctypes.data, dv.ptr(grid),
                                       dv.ptr(flag), dv.stream_ptr()))
    return grid, flag


def trajectory_to_volume(trajectory, resolution: float = 0.2) -> Volume:
    lattice = trajectory.get_lattice()
    pos = trajectory._device_positions()
    nedges = grid_edges(lattice, resolution)
    grid, flag = density_device(pos, nedges)
    # the reference asserts 0 <= coords < 1 (volume.py:476-477)
    assert int(flag.item()) == 0
    data = dv.to_host(grid).view(np.uint32).astype(int)
    return Volume(data=data, lattice=lattice, label='trajectory')
