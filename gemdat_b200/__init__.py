"""gemdat_b200 -- B200-native implementation of GEMDAT's data-parallel hot path.

Site assignment -> transition events -> jumps, the displacement / MSD path and the
density histogram, behind GEMDAT's own Python API.  Everything O(frames x atoms) on
the path runs in hand-written sm_100a CUDA kernels (``csrc/``, C ABI in
``include/gemdat_b200.h``); library code appears only on secondary paths (CUB sort / scan
in the collective-jump search, cuFFT in ``attempt_frequency``).
There is no CPU fallback.
"""

from __future__ import annotations

from . import _lib  # fails loudly if libgemdat_b200.so is missing
from .collective import Collective
from .jumps import Jumps
from .lattice import FloatWithUnit, Lattice, Sites
from .metrics import TrajectoryMetrics
from .rdf import radial_distribution
from .trajectory import Trajectory
from .transitions import SiteRadius, Transitions
from .volume import Volume, trajectory_to_volume

__version__ = '0.1.0'
__all__ = [
    'Collective',
    'Jumps',
    'Lattice',
    'Sites',
    'SiteRadius',
    'Trajectory',
    'TrajectoryMetrics',
    'Transitions',
    'Volume',
    'radial_distribution',
    'trajectory_to_volume',
]
