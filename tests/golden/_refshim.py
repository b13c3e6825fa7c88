"""Execute the reference's OWN modules in the build container.

Used only by ``tests/golden/make_golden.py`` (never at test/bench time: the
reference does not exist on the GPU box).  GEMDAT is pure Python but its
third-party layer (pymatgen, MDAnalysis, uncertainties, scikit-image, plotting)
is not installed here, so that layer -- and only that layer -- is replaced by
stubs before ``/root/reference/src/gemdat/*.py`` is imported unmodified:

* ``pymatgen.core.trajectory.Trajectory`` / ``pymatgen.core.Lattice`` /
  ``Structure``  -> minimal restatements of pymatgen 2024.1.27 [recalled]
  (``oracle.pmg`` holds the arithmetic);
* ``MDAnalysis.lib.pkdtree.PeriodicKDTree`` -> ``oracle.mda.PeriodicKDTree``
  [recalled wrapper around the REAL scipy cKDTree];
* plotting / skimage / uncertainties -> inert placeholders.

Everything GEMDAT itself computes (``_calculate_atom_states`` control flow,
``_calculate_transition_events``, ``ffill/bfill/integer_remap``,
``_generic_transitions_to_jumps`` incl. its pandas ``iterrows`` state machine,
``Jumps.matrix/jump_diffusivity``, ``TrajectoryMetrics``, ``mean_squared_displacement``,
``trajectory_to_volume``) is therefore the reference's real code.
"""

from __future__ import annotations

import importlib
import sys
import types
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
REFERENCE_SRC = Path('/root/reference/src/gemdat')

if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

from oracle import mda as _mda  # noqa: E402
from oracle import pmg as _pmg  # noqa: E402


# ----------------------------------------------------------------------------- pymatgen stubs


class Element:
    def __init__(self, symbol: str):
        self.symbol = symbol
        self.name = symbol

    def __eq__(self, other):
        return getattr(other, 'symbol', None) == self.symbol

    def __hash__(self):
        return hash(self.symbol)

    def __repr__(self):
        return f'Element {self.symbol}'

    @property
    def atomic_mass(self):
        # pymatgen's periodic_table.json value [recalled]: the same table the product uses, so a
        # golden made through this stub pins the centre-of-mass ARITHMETIC, not the table
        from gemdat_b200.elements import ATOMIC_MASS

        return ATOMIC_MASS[self.symbol]


class Species(Element):
    pass


class Lattice(_pmg.Lattice):
    pass


class _Composition:
    """pymatgen Composition, reduced: one element, ``num_atoms`` = its amount  [recalled]."""

    def __init__(self, symbol, amount=1.0):
        self.elements = [Element(symbol)]
        self.num_atoms = float(amount)


class PeriodicSite:
    def __init__(self, symbol, frac_coords, label):
        amount = 1.0
        if isinstance(symbol, dict):  # Structure(species=[{name: occupancy}, ...]) (transitions.py:294-297)
            (symbol, amount), = symbol.items()
        else:
            symbol = str(symbol)
        self.specie = Element(symbol)
        self.species = _Composition(symbol, amount)
        self.frac_coords = np.asarray(frac_coords, dtype=float)
        self.label = label
        self.is_ordered = True


class Structure:
    def __init__(self, lattice, species, coords, labels=None, site_properties=None):
        self.lattice = lattice if isinstance(lattice, _pmg.Lattice) else Lattice(lattice)
        labels = list(labels) if labels is not None else [str(s) for s in species]
        self._sites = [PeriodicSite(s, c, lab) for s, c, lab in zip(species, coords, labels)]
        self.site_properties = site_properties or {}

    @property
    def frac_coords(self):
        return np.array([s.frac_coords for s in self._sites])

    @property
    def labels(self):
        return [s.label for s in self._sites]

    @property
    def species(self):
        return [s.specie for s in self._sites]

    @property
    def symbol_set(self):
        # pymatgen: sorted tuple of the element symbols present  [recalled]
        return tuple(sorted({s.specie.symbol for s in self._sites}))

    @property
    def is_ordered(self):
        return True

    def __len__(self):
        return len(self._sites)

    def __iter__(self):
        return iter(self._sites)

    def __getitem__(self, i):
        return self._sites[i]


class PymatgenTrajectory:
    """pymatgen.core.trajectory.Trajectory, reduced to what GEMDAT touches [recalled]."""

    def __init__(
        self,
        species,
        coords,
        charge=None,
        spin_multiplicity=None,
        lattice=None,
        *,
        site_properties=None,
        frame_properties=None,
        constant_lattice=True,
        time_step=None,
        coords_are_displacement=False,
        base_positions=None,
    ):
        coords = np.asarray(coords)
        if isinstance(lattice, _pmg.Lattice):
            lattice = lattice.matrix
        self.lattice = None if lattice is None else np.asarray(lattice)
        self.constant_lattice = constant_lattice
        if coords_are_displacement:
            if base_positions is None:
                base_positions = coords[0]
            self.base_positions = base_positions
        else:
            self.base_positions = coords[0]
        self.coords_are_displacement = coords_are_displacement
        self.species = [Species(s) if isinstance(s, str) else s for s in species]
        self.coords = coords
        self.time_step = time_step
        self.site_properties = site_properties
        self.frame_properties = frame_properties

    def __len__(self):
        return len(self.coords)

    def get_structure(self, idx):
        # pymatgen Trajectory.get_structure: positions of frame idx  [recalled]
        self.to_positions()
        return Structure(Lattice(self.lattice), [sp.symbol for sp in self.species], self.coords[idx])

    def to_positions(self):
        if not self.coords_are_displacement:
            return
        self.coords = self.base_positions + np.cumsum(self.coords, axis=0)
        self.coords_are_displacement = False

    def to_displacements(self):
        if self.coords_are_displacement:
            return
        self.coords = _pmg.to_displacements(self.coords)
        self.coords_are_displacement = True

    def __getitem__(self, frames):
        if isinstance(frames, slice):
            return self.__class__(
                species=self.species,
                coords=self.coords[frames],
                lattice=self.lattice,
                constant_lattice=self.constant_lattice,
                time_step=self.time_step,
                coords_are_displacement=False,
                base_positions=self.base_positions,
            )
        raise NotImplementedError


class FloatWithUnit(float):
    def __new__(cls, val, unit=None):
        obj = float.__new__(cls, val)
        obj.unit = unit
        return obj


def _module(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


def _identity_decorator(func):
    return func


def install_stubs():
    core = _module(
        'pymatgen.core',
        Element=Element, Species=Species, Lattice=Lattice, Structure=Structure,
        PeriodicSite=PeriodicSite,
    )
    core.__path__ = []
    pm = _module('pymatgen', core=core)
    pm.__path__ = []
    _module('pymatgen.core.trajectory', Trajectory=PymatgenTrajectory)
    _module('pymatgen.core.units', FloatWithUnit=FloatWithUnit, Unit=str)
    vasp = _module('pymatgen.io.vasp', VolumetricData=object, Vasprun=object)
    io = _module('pymatgen.io', vasp=vasp)
    io.__path__ = []
    pkd = _module('MDAnalysis.lib.pkdtree', PeriodicKDTree=_mda.PeriodicKDTree)
    lib = _module('MDAnalysis.lib', pkdtree=pkd)
    lib.__path__ = []
    m = _module('MDAnalysis', lib=lib)
    m.__path__ = []
    _module('uncertainties', ufloat=lambda n, s: (n, s), UFloat=object)
    sk = _module('skimage')
    sk.__path__ = []
    _module('skimage.feature', blob_dog=None)
    _module('skimage.measure', regionprops=None)
    _module('skimage.segmentation', watershed=None)
    pkg = _module('gemdat')
    pkg.__path__ = [str(REFERENCE_SRC)]
    _module('gemdat._plot_backend', plot_backend=_identity_decorator)


_loaded = {}


def load_reference():
    """Import the reference modules (unmodified source) on top of the stubs."""
    if _loaded:
        return _loaded
    if not REFERENCE_SRC.exists():
        raise RuntimeError('/root/reference is only available in the build container')
    install_stubs()
    import importlib.resources as _res

    _orig_files = _res.files
    _res.files = lambda anchor: REFERENCE_SRC if anchor == 'gemdat' else _orig_files(anchor)
    try:
        _import_all()
    finally:
        _res.files = _orig_files
    return _loaded


def _import_all():
    for name in ('utils', 'caching', 'metrics', 'transitions', 'collective', 'jumps',
                 'trajectory', 'volume', 'rdf'):
        _loaded[name] = importlib.import_module(f'gemdat.{name}')
