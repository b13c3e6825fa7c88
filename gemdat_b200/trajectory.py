"""Trajectory: the input contract of the hot path and its displacement / MSD / density methods.

Mirrors the parts of ``gemdat.Trajectory`` (reference src/gemdat/trajectory.py) the
path touches: construction kwargs, ``positions`` / ``displacements`` (:170-193),
``filter`` (:833-864), ``get_lattice`` (:686-703), ``split`` (:866-894),
``cumulative_displacements`` (:705-713), ``distances_from_base_position`` (:715-730),
``mean_squared_displacement`` (:896-933), ``to_volume`` (:129-150),
``transitions_between_sites`` (:1066-1110), ``metrics`` (:1112-1116).

File readers, drift correction, plotting and the kinisi bridge are out of scope
(DESIGN.md).  ``coords`` may be a numpy array (copied to the GPU per call) or a
CUDA ``torch.Tensor`` (device resident: the 24 GB C3 trajectory never visits the
host).  All arithmetic on the path runs in ``libgemdat_b200.so``.
"""

from __future__ import annotations

from itertools import pairwise
from typing import Collection

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .lattice import Lattice, as_lattice_matrix


def _symbol(sp) -> str:
    """Element symbol of a species given as 'Li', 'Li+', pymatgen Species/Element, ..."""
    s = getattr(sp, 'symbol', None)
    if s is not None:
        return str(s)
    s = str(sp)
    out = ''
    for ch in s:
        if ch.isalpha():
            out += ch
        else:
            break
    return out or s


def _wrap_unit_device(t: torch.Tensor, *, in_place: bool) -> torch.Tensor | None:
    """``np.mod(t, 1)`` on the device if some element lies outside [0, 1); None when all are
    inside (one read-only pass, ``gd_wrap_unit``).  ``in_place=False`` leaves ``t`` untouched."""
    flag = dv.zeros((1,), torch.int32)
    lib, s = _lib.lib, dv.stream_ptr()
    _lib.check(lib.gd_wrap_unit(dv.ptr(t), None, t.numel(), dv.ptr(flag), s))
    if int(flag.item()) == 0:
        return None
    out = t if in_place else torch.empty_like(t)
    _lib.check(lib.gd_wrap_unit(dv.ptr(t), dv.ptr(out), t.numel(), dv.ptr(flag), s))
    return out


class Trajectory:
    def __init__(
        self,
        *,
        species,
        coords,
        lattice,
        metadata: dict | None = None,
        time_step: float | None = None,
        constant_lattice: bool = True,
        coords_are_displacement: bool = False,
        base_positions=None,
        **_ignored,
    ):
        self.species = list(species)
        self.coords = coords if dv.is_device(coords) else np.asarray(coords, dtype=np.float64)
        if self.coords.ndim != 3 or self.coords.shape[2] != 3:
            raise ValueError('coords must have shape [n_frames, n_atoms, 3]')
        if self.coords.shape[1] != len(self.species):
            raise ValueError('len(species) must equal coords.shape[1]')
        self.constant_lattice = bool(constant_lattice)
        if self.constant_lattice:
            self.lattice = as_lattice_matrix(lattice)
        else:
            # one cell per frame.  Only `radial_distribution_between_species` walks them
            # (rdf.py:232-261); everything else on the path asks for THE lattice, like the
            # reference, whose get_lattice() without an index is meaningless here
            # (trajectory.py:686-703, transitions.py:769)
            self.lattice = np.ascontiguousarray(np.asarray(lattice, dtype=np.float64).reshape(-1, 3, 3))
            if len(self.lattice) != self.coords.shape[0]:
                raise ValueError('constant_lattice=False needs one 3x3 lattice per frame')
        self.metadata = metadata if metadata else {}
        self.time_step = time_step
        self.coords_are_displacement = coords_are_displacement
        if coords_are_displacement:
            self.base_positions = self.coords[0] if base_positions is None else base_positions
        else:
            self.base_positions = self.coords[0]
        # positions known to lie in [0, 1)?  None = not checked yet (see `positions`)
        self._unit_wrapped: bool | None = None

    # ------------------------------------------------------------------ container protocol
    def __len__(self) -> int:
        return int(self.coords.shape[0])

    def __getitem__(self, frames):
        if not isinstance(frames, slice):
            raise TypeError('only slicing along the frame axis is supported')
        self.to_positions()
        return self.__class__(
            species=self.species,
            coords=self.coords[frames],
            lattice=self.lattice if self.constant_lattice else self.lattice[frames],
            constant_lattice=self.constant_lattice,
            metadata=self.metadata,
            time_step=self.time_step,
        )

    def __repr__(self):
        return (
            f'Trajectory({len(self.species)} sites, {len(self)} time steps, '
            f'abc={Lattice(self.lattice if self.constant_lattice else self.lattice[0]).lengths}, '
            f'constant_lattice={self.constant_lattice}, device_resident={dv.is_device(self.coords)})'
        )

    # ------------------------------------------------------------------ pickling / cache files
    def __getstate__(self):
        """Host data only: the cached device copy is dropped, device-resident coords are
        downloaded (trajectory.py:110-119 drops its runtime-only caches the same way)."""
        state = self.__dict__.copy()
        state.pop('_dev_holder', None)
        for key in ('coords', 'base_positions'):
            if dv.is_device(state.get(key)):
                state[key] = dv.to_host(state[key])
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)

    def to_cache(self, cache) -> None:
        """trajectory.py:208-217: pickle the trajectory to ``cache``."""
        import pickle

        with open(cache, 'wb') as f:
            pickle.dump(self, f)

    @classmethod
    def from_cache(cls, cache) -> 'Trajectory':
        """trajectory.py:195-206: load a trajectory pickled by ``to_cache``."""
        import pickle

        with open(cache, 'rb') as f:
            return pickle.load(f)

    # ------------------------------------------------------------------ representations
    def to_positions(self):
        """trajectory.py:121-127: positions representation, then ``np.mod(coords, 1)``.

        The reference applies the modulo on every call; here it is applied once per coords
        array (it is idempotent): coordinates outside [0, 1) are detected -- on the device for
        CUDA tensors (``gd_wrap_unit``), with one min/max pass for numpy arrays -- and only then
        a wrapped copy replaces ``coords``."""
        if self.coords_are_displacement:
            coords = self._host_coords()
            base = np.asarray(dv.to_host(self.base_positions) if dv.is_device(self.base_positions)
                              else self.base_positions)
            self.coords = np.mod(base + np.cumsum(coords, axis=0), 1)
            self.coords_are_displacement = False
            self._unit_wrapped = True
            return
        if self._unit_wrapped:
            return
        coords = self.coords
        if dv.is_device(coords):
            wrapped = _wrap_unit_device(coords, in_place=False)
            if wrapped is not None:
                self.coords = wrapped
        elif coords.size and (coords.min() < 0.0 or coords.max() >= 1.0):
            self._rebind_coords(np.mod(coords, 1))
        self._unit_wrapped = True

    def _rebind_coords(self, new_coords: np.ndarray) -> None:
        """Replace the host array, keeping the cached device copy (which holds the wrapped
        values already) attached to the new object."""
        holder = self.__dict__.get('_dev_holder') or {}
        cached = holder.get('pos')
        if cached is not None and cached[0] is self.coords:
            holder['pos'] = (new_coords, cached[1])
        self.coords = new_coords

    def to_displacements(self):
        """pymatgen ``to_displacements``: wrapped frame-to-frame steps, frame 0 = 0."""
        if self.coords_are_displacement:
            return
        coords = self._host_coords()
        d = np.subtract(coords, np.roll(coords, 1, axis=0))
        d[0] = 0.0
        self.coords = np.subtract(d, np.around(d))
        self.coords_are_displacement = True

    def _host_coords(self) -> np.ndarray:
        return dv.to_host(self.coords) if dv.is_device(self.coords) else self.coords

    @property
    def positions(self):
        """Fractional positions [n_frames, n_atoms, 3] (numpy, or the CUDA tensor if device resident)."""
        self.to_positions()
        return self.coords

    @property
    def displacements(self) -> np.ndarray:
        self.to_displacements()
        return self.coords

    def _device_positions(self) -> torch.Tensor:
        """float64 CUDA tensor [T, A, 3] of wrapped positions.

        Host arrays are uploaded once per coords array: the device copy is cached on the
        object (and shared with ``filter`` results that keep every atom), so
        ``transitions_between_sites`` + ``metrics`` + ``to_volume`` on one trajectory pay for
        one host->device copy, not three.
        """
        if self.coords_are_displacement or dv.is_device(self.coords):
            self.to_positions()
        if dv.is_device(self.coords):
            return dv.to_device(self.coords, torch.float64)
        holder = self.__dict__.setdefault('_dev_holder', {})
        cached = holder.get('pos')
        if cached is not None and cached[0] is self.coords:
            return cached[1]
        t = dv.to_device(self.coords, torch.float64)
        # np.mod(coords, 1) of to_positions, on the device copy (ours: wrapped in place); the
        # host array is only rewritten if someone reads `positions` from it
        if not self._unit_wrapped and _wrap_unit_device(t, in_place=True) is None:
            self._unit_wrapped = True
        holder['pos'] = (self.coords, t)
        return t

    # ------------------------------------------------------------------ scalars
    @property
    def time_step_ps(self) -> float:
        assert self.time_step
        return self.time_step * 1e12

    @property
    def total_time(self) -> float:
        assert self.time_step
        return len(self) * self.time_step

    @property
    def sampling_frequency(self) -> float:
        assert self.time_step
        return 1 / self.time_step

    def get_lattice(self, idx: int | None = None) -> Lattice:
        """trajectory.py:686-703: the lattice, or the one of frame ``idx`` when it is not constant."""
        if self.constant_lattice:
            return Lattice(self.lattice)
        if idx is None:
            raise NotImplementedError(
                'this trajectory has one lattice per frame: only radial_distribution_between_species '
                'supports that (the rest of the path uses a single lattice, like the reference)')
        return Lattice(self.lattice[idx])

    # ------------------------------------------------------------------ selection
    def filter(self, species: str | Collection[str]) -> 'Trajectory':
        """trajectory.py:833-864: keep the atoms whose element symbol is in ``species``."""
        if isinstance(species, str):
            species = [species]
        wanted = {_symbol(s) for s in species}
        idx = [_symbol(sp) in wanted for sp in self.species]
        # (positions are wrapped lazily: by `positions` on the host, by `_device_positions` on
        # the device copy -- filtering commutes with the elementwise modulo)
        pos = self.positions if self.coords_are_displacement else self.coords
        keep_all = all(idx)
        if keep_all:
            # the reference copies (trajectory.py:854); coords are never modified in place here
            # (representation changes rebind the attribute), so the array is shared
            new_coords = pos
        elif dv.is_device(pos):
            sel = torch.nonzero(torch.tensor(idx, device=pos.device)).flatten()
            new_coords = pos.index_select(1, sel)
        else:
            new_coords = np.take(pos, np.nonzero(idx)[0], axis=1)
        out = self.__class__(
            species=[sp for sp, keep in zip(self.species, idx) if keep],
            coords=new_coords,
            lattice=self.lattice,
            constant_lattice=self.constant_lattice,
            metadata=self.metadata,
            time_step=self.time_step,
        )
        out._unit_wrapped = self._unit_wrapped
        if keep_all:  # same coords object: parent and child share one device copy
            out._dev_holder = self.__dict__.setdefault('_dev_holder', {})
        return out

    def split(self, n_parts: int = 10, equal_parts: bool = False) -> list['Trajectory']:
        """trajectory.py:866-894."""
        interval = np.linspace(0, len(self) - 1, n_parts + 1, dtype=int)
        parts = [self[int(a):int(b)] for a, b in pairwise(interval)]
        if equal_parts:
            minsize = min(int(b - a) for a, b in pairwise(interval))
            parts = [p[0:minsize] for p in parts]
        return parts

    # ------------------------------------------------------------------ K0: displacement path
    def _unwrap(self, *, want_frac=False, want_cart=False, want_dist=False, want_speed=False,
                keep_on_device=False):
        """K0.  ``speed`` comes out TIME-major [n_frames, n_atoms] (the layout the amplitude
        kernels stream); ``dist`` atom-major like the reference's array."""
        lattice = self.get_lattice().matrix  # (raises for one lattice per frame)
        pos = self._device_positions()
        n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
        n_chunks = int(_lib.lib.gd_unwrap_chunks(n_frames, n_atoms))
        work = dv.empty((n_chunks * n_atoms * 3,), torch.int32)
        frac = dv.empty((n_frames, n_atoms, 3), torch.float64) if want_frac else None
        cart = dv.empty((n_frames, n_atoms, 3), torch.float64) if want_cart else None
        dist = dv.empty((n_atoms, n_frames), torch.float64) if want_dist else None
        speed = dv.empty((n_frames, n_atoms), torch.float64) if want_speed else None
        wind = dv.empty((n_atoms, 3), torch.int32)
        _lib.check(
            _lib.lib.gd_unwrap(
                dv.ptr(pos), n_frames, n_atoms, lattice.ctypes.data, None, None, None,
                dv.ptr(work), dv.ptr(frac), dv.ptr(cart), dv.ptr(dist), dv.ptr(speed), dv.ptr(wind),
                dv.stream_ptr(),
            )
        )
        out = {'frac': frac, 'cart': cart, 'dist': dist, 'speed': speed, 'winding': wind, 'pos': pos}
        if not keep_on_device:
            out = {k: (dv.to_host(v) if v is not None and k != 'pos' else v) for k, v in out.items()}
        return out

    @property
    def cumulative_displacements(self) -> np.ndarray:
        """trajectory.py:705-713, [n_frames, n_atoms, 3] fractional."""
        return self._unwrap(want_frac=True)['frac']

    def distances_from_base_position(self) -> np.ndarray:
        """trajectory.py:715-730, [n_atoms, n_frames] in Angstrom."""
        return self._unwrap(want_dist=True)['dist']

    def final_displacement_sq(self) -> np.ndarray:
        """|r(T-1) - r(0)|^2 per atom (Angstrom^2) from the winding counts alone."""
        res = self._unwrap(keep_on_device=True)
        pos = res['pos']
        first, last = dv.to_host(pos[0]), dv.to_host(pos[-1])
        frac = (last - first) - dv.to_host(res['winding']).astype(np.float64)
        g = self.get_lattice().metric_tensor
        return np.einsum('ij,ij->i', frac @ g, frac)

    # ------------------------------------------------------------------ per-frame averages, drift
    def _selected_atoms(self, fixed_species, floating_species) -> np.ndarray:
        """Atom numbers the drift is averaged over (trajectory.py:778-789: the fixed species, or
        every species that is not floating, or all atoms)."""
        symbols = [_symbol(sp) for sp in self.species]
        if fixed_species:
            wanted = [fixed_species] if isinstance(fixed_species, str) else fixed_species
            wanted = {_symbol(s) for s in wanted}
            keep = [s in wanted for s in symbols]
        elif floating_species:
            keep = [s not in floating_species for s in symbols]  # (`in` on a str: like the reference)
        else:
            keep = [True] * len(symbols)
        return np.nonzero(keep)[0].astype(np.int32)

    def _drift_device(self, fixed_species, floating_species) -> torch.Tensor | None:
        """float64 [n_frames, 3] on the device, None when no atom is selected."""
        sel = self._selected_atoms(fixed_species, floating_species)
        if sel.size == 0:
            return None
        pos = self._device_positions()
        n_frames, n_atoms = int(pos.shape[0]), int(pos.shape[1])
        out = dv.empty((n_frames, 3), torch.float64)
        d_sel = None if sel.size == n_atoms else dv.to_device(sel)
        _lib.check(_lib.lib.gd_drift(dv.ptr(pos), n_frames, n_atoms, dv.ptr(d_sel), int(sel.size), None,
                                     dv.ptr(out), dv.stream_ptr()))
        return out

    def drift(self, *, fixed_species=None, floating_species=None) -> np.ndarray:
        """trajectory.py:751-791: mean wrapped step of the fixed atoms per frame, [n_frames, 1, 3]
        (``gd_drift``: one CTA per frame)."""
        out = self._drift_device(fixed_species, floating_species)
        if out is None:
            import warnings

            warnings.warn('Mean of empty slice.', RuntimeWarning, stacklevel=2)  # numpy's, over 0 atoms
            return np.full((len(self), 1, 3), np.nan)
        return dv.to_host(out)[:, None, :]

    def apply_drift_correction(self, *, fixed_species=None, floating_species=None) -> 'Trajectory':
        """trajectory.py:793-830: subtract the drift from every atom's steps.

        The reference returns the corrected steps in displacement form and sums them along the
        frames when positions are asked for; here the corrected, wrapped POSITIONS come back
        directly (K0's cumulative displacements minus the summed drift, ``gd_drift_correct``),
        on the device when the input was: the same trajectory up to rounding of the running sum.
        """
        drift = self._drift_device(fixed_species, floating_species)
        n_frames, n_atoms = len(self), len(self.species)
        if drift is None:
            raise ValueError('no atoms selected to compute the drift from')
        lib, s = _lib.lib, dv.stream_ptr()
        summed = dv.empty((n_frames, 3), torch.float64)
        scan_work = dv.empty((int(lib.gd_cumsum_frames3_work_bytes(n_frames)),), torch.uint8)
        _lib.check(lib.gd_cumsum_frames3(dv.ptr(drift), n_frames, dv.ptr(scan_work), dv.ptr(summed), s))
        was_device = dv.is_device(self.coords)
        base = dv.to_device(np.ascontiguousarray(self._host_base_positions()), torch.float64)
        cum = self._unwrap(want_frac=True, keep_on_device=True)['frac']
        _lib.check(lib.gd_drift_correct(dv.ptr(cum), dv.ptr(base), dv.ptr(summed), n_frames, n_atoms,
                                        dv.ptr(cum), s))
        out = self.__class__(species=self.species, coords=cum if was_device else dv.to_host(cum),
                             lattice=self.lattice, metadata=self.metadata, time_step=self.time_step)
        out._unit_wrapped = True
        return out

    def _host_base_positions(self) -> np.ndarray:
        base = self.base_positions
        return dv.to_host(base) if dv.is_device(base) else np.asarray(base, dtype=np.float64)

    def center_of_mass(self) -> 'Trajectory':
        """trajectory.py:732-749: one pseudo atom 'X' at the mass-weighted mean of the unwrapped
        positions (base positions + K0's cumulative displacements, ``gd_frame_mean``)."""
        from .elements import atomic_mass

        weights = np.array([atomic_mass(sp) for sp in self.species], dtype=np.float64)
        cum = self._unwrap(want_frac=True, keep_on_device=True)['frac']
        n_frames, n_atoms = int(cum.shape[0]), int(cum.shape[1])
        base = dv.to_device(np.ascontiguousarray(self._host_base_positions()), torch.float64)
        d_weights = dv.to_device(weights)
        out = dv.empty((n_frames, 3), torch.float64)
        _lib.check(_lib.lib.gd_frame_mean(dv.ptr(cum), dv.ptr(base), dv.ptr(d_weights), n_frames, n_atoms,
                                          dv.ptr(out), dv.stream_ptr()))
        return self.__class__(species=['X'], coords=dv.to_host(out)[:, None, :], lattice=self.lattice,
                              metadata=self.metadata, time_step=self.time_step)

    # ------------------------------------------------------------------ K4: MSD
    def mean_squared_displacement(self, lags: tuple[int, int] | None = None) -> np.ndarray:
        """trajectory.py:896-933: per-atom MSD for every lag, [n_atoms, n_frames] Angstrom^2.

        ``lags=(first, count)`` restricts the computation to a contiguous lag range
        (extension; the reference always returns all lags).
        """
        res = self._unwrap(want_cart=True, keep_on_device=True)
        cart = res['cart']
        n_frames, n_atoms = int(cart.shape[0]), int(cart.shape[1])
        soa = dv.empty((n_atoms, 3, n_frames), torch.float64)
        _lib.check(_lib.lib.gd_transpose_soa(dv.ptr(cart), n_frames, n_atoms, dv.ptr(soa), dv.stream_ptr()))
        del cart, res
        lag0, n_lags = (0, n_frames) if lags is None else (int(lags[0]), int(lags[1]))
        msd = dv.empty((n_atoms, n_lags), torch.float64)
        _lib.check(
            _lib.lib.gd_msd_direct(dv.ptr(soa), n_frames, n_atoms, lag0, n_lags, dv.ptr(msd), dv.stream_ptr())
        )
        return dv.to_host(msd)

    # ------------------------------------------------------------------ K5: density
    def to_volume(self, resolution: float = 0.2):
        from .volume import trajectory_to_volume

        return trajectory_to_volume(self, resolution=resolution)

    # ------------------------------------------------------------------ K1-K3
    def transitions_between_sites(
        self,
        sites,
        floating_specie: str,
        site_radius: float | dict[str, float] | None = None,
        site_inner_fraction: float | dict[str, float] = 1.0,
        remove_part_occup_from_structure: bool = False,
        fraction_of_overlap: float = 0.0,
    ):
        """trajectory.py:1066-1110."""
        from .transitions import Transitions

        return Transitions.from_trajectory(
            trajectory=self,
            sites=sites,
            floating_specie=floating_specie,
            site_radius=site_radius,
            site_inner_fraction=site_inner_fraction,
            remove_part_occup_from_structure=remove_part_occup_from_structure,
            fraction_of_overlap=fraction_of_overlap,
        )

    def metrics(self):
        from .metrics import TrajectoryMetrics

        return TrajectoryMetrics(trajectory=self)

    def radial_distribution_between_species(self, **kwargs):
        """trajectory.py:1118-1124."""
        from . import rdf

        return rdf.radial_distribution_between_species(trajectory=self, **kwargs)
