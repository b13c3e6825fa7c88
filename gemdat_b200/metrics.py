"""TrajectoryMetrics: the displacement-derived scalars on the hot path.

Mirror of reference src/gemdat/metrics.py:21-231 for ``speed`` (:35-47),
``tracer_diffusivity`` (:76-97), ``amplitudes`` (:209-231), ``vibration_amplitude``
(:190-207) and ``particle_density`` (:49-62).  Everything O(frames x atoms) stays in HBM:
K0 (``gd_unwrap``) writes the speed array time-major, ``gd_amplitudes_count/fill`` cut it at the
sign flips and sum the pieces, ``gd_mean_std`` reduces them -- ``vibration_amplitude()`` (the
``site_radius=None`` route of ``transitions_between_sites``) never downloads an [A, T] array.
``attempt_frequency`` (:167-188) runs on the same device array: the per-atom
spectrum comes from cuFFT through ``torch.fft`` (a library transform, not one of this
package's kernels), the periodogram weighting and the mean-frequency reductions of
``utils.py:128-160`` are one pass of ``gd_meanfreq`` over it.  ``mol_per_liter`` (:64-74), ``tracer_diffusivity_center_of_mass``
(:99-120), ``haven_ratio`` (:122-137) and ``tracer_conductivity`` (:140-165) are scalars on top of
these; the multi-trajectory statistics (``TrajectoryMetricsStd``) and the Arrhenius helpers are out
of scope.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .caching import weak_lru_cache
from .lattice import FloatWithUnit

ANGSTROM = 1e-10


class TrajectoryMetrics:
    def __init__(self, trajectory):
        self.trajectory = trajectory

    # ---- device-resident intermediates (cached per object like the reference's results)
    @weak_lru_cache()
    def _speed_tm(self) -> torch.Tensor:
        """np.diff(distance, prepend=0) as a TIME-major [n_frames, n_atoms] device array."""
        dv.require_cuda()
        return self.trajectory._unwrap(want_speed=True, keep_on_device=True)['speed']

    @weak_lru_cache()
    def _amplitudes_device(self) -> torch.Tensor:
        speed = self._speed_tm()
        n_frames, n_atoms = int(speed.shape[0]), int(speed.shape[1])
        lib, s = _lib.lib, dv.stream_ptr()
        work = dv.empty((int(lib.gd_amplitudes_work_bytes(n_frames, n_atoms)),), torch.uint8)
        offsets = dv.empty((n_atoms + 1,), torch.int64)
        n_amp = C.c_int64(0)
        _lib.check(lib.gd_amplitudes_count(dv.ptr(speed), n_frames, n_atoms, dv.ptr(work),
                                           dv.ptr(offsets), C.byref(n_amp), s))
        amp = dv.empty((n_amp.value,), torch.float64)
        _lib.check(lib.gd_amplitudes_fill(dv.ptr(speed), n_frames, n_atoms, dv.ptr(work),
                                          dv.ptr(offsets), dv.ptr(amp), s))
        return amp

    @weak_lru_cache()
    def speed(self) -> np.ndarray:
        """metrics.py:35-47: change of the distance from the base position per step, [A, T]."""
        tm = self._speed_tm()
        out = dv.empty((int(tm.shape[1]), int(tm.shape[0])), torch.float64)
        _lib.check(_lib.lib.gd_transpose_f64(dv.ptr(tm), int(tm.shape[0]), int(tm.shape[1]), dv.ptr(out),
                                             dv.stream_ptr()))
        return dv.to_host(out)

    @weak_lru_cache()
    def _meanfreq(self) -> np.ndarray:
        """utils.py:128-160 on the device: mean frequency of every atom's speed, [A, 1].

        ``scipy.signal.periodogram`` defaults restated: constant detrend, boxcar window, one-sided
        spectrum with every bin but DC (and Nyquist for even N) doubled; ``mnfreq = sum(P f) /
        sum(P)``.  The transform is cuFFT (``torch.fft.rfft`` of one row per atom, a library
        call); everything after it -- weighting, both sums, the ratio -- is one
        pass of ``gd_meanfreq`` over the spectrum.  A constant offset only lands in the DC bin,
        which the kernel skips, so no detrended copy of the array is made.
        """
        tm = self._speed_tm()  # [T, A]
        n, n_atoms = int(tm.shape[0]), int(tm.shape[1])
        lib, s = _lib.lib, dv.stream_ptr()
        rows = dv.empty((n_atoms, n), torch.float64)  # one row per atom: what the batched FFT wants
        _lib.check(lib.gd_transpose_f64(dv.ptr(tm), n, n_atoms, dv.ptr(rows), s))
        spec = torch.fft.rfft(rows, dim=1)  # [A, n // 2 + 1] complex128
        del rows
        out = dv.empty((n_atoms,), torch.float64)
        _lib.check(lib.gd_meanfreq(dv.ptr(spec), n, n_atoms, float(self.trajectory.sampling_frequency),
                                   dv.ptr(out), s))
        return dv.to_host(out).reshape(-1, 1)

    @weak_lru_cache()
    def attempt_frequency(self) -> tuple[FloatWithUnit, FloatWithUnit]:
        """metrics.py:167-188: mean and standard deviation over atoms of the mean frequency, Hz."""
        freq_mean = self._meanfreq()
        return FloatWithUnit(np.mean(freq_mean), 'hz'), FloatWithUnit(np.std(freq_mean), 'hz')

    @weak_lru_cache()
    def particle_density(self) -> FloatWithUnit:
        volume_m3 = self.trajectory.get_lattice().volume * ANGSTROM**3
        return FloatWithUnit(len(self.trajectory.species) / volume_m3, 'm^-3')

    @weak_lru_cache()
    def mol_per_liter(self) -> FloatWithUnit:
        """metrics.py:64-74."""
        from scipy.constants import Avogadro

        return FloatWithUnit((self.particle_density() * 1e-3) / Avogadro, 'mol l^-1')

    @weak_lru_cache()
    def tracer_diffusivity_center_of_mass(self, *, dimensions: int = 3) -> FloatWithUnit:
        """metrics.py:99-120: tracer diffusivity of the centre of mass (one pseudo atom)."""
        return TrajectoryMetrics(self.trajectory.center_of_mass()).tracer_diffusivity(dimensions=dimensions)

    @weak_lru_cache()
    def haven_ratio(self, *, dimensions: int = 3) -> float:
        """metrics.py:122-137."""
        return self.tracer_diffusivity(dimensions=dimensions) / self.tracer_diffusivity_center_of_mass(
            dimensions=dimensions)

    @weak_lru_cache()
    def tracer_conductivity(self, *, z_ion: int, dimensions: int = 3) -> FloatWithUnit:
        """metrics.py:140-165: e^2 z^2 D n / (k_B T), S/m."""
        from scipy.constants import Boltzmann, elementary_charge

        temperature = self.trajectory.metadata['temperature']
        tracer_diff = self.tracer_diffusivity(dimensions=dimensions)
        conductivity = ((elementary_charge**2) * (z_ion**2) * tracer_diff * self.particle_density()) / (
            Boltzmann * temperature)
        return FloatWithUnit(conductivity, 'S m^-1')

    @weak_lru_cache()
    def tracer_diffusivity(self, *, dimensions: int = 3) -> FloatWithUnit:
        """metrics.py:76-97: mean |r(T-1)-r(0)|^2 / (2 d total_time), m^2/s.

        Only the final frame matters, so this reads the winding totals of K0 and never
        materialises the [A, T] distance array the reference builds.
        """
        msd = float(np.mean(self.trajectory.final_displacement_sq()))
        tracer = (msd * ANGSTROM**2) / (2 * dimensions * self.trajectory.total_time)
        return FloatWithUnit(tracer, 'm^2 s^-1')

    @weak_lru_cache()
    def amplitudes(self) -> np.ndarray:
        """metrics.py:209-231: every atom's speed row cut after each sign flip except the first
        and the last one (``splits[1:-1]``), each piece summed; atoms in order, pieces in time
        order (``gd_amplitudes_count`` / ``gd_amplitudes_fill``)."""
        return dv.to_host(self._amplitudes_device())

    @weak_lru_cache()
    def vibration_amplitude(self) -> FloatWithUnit:
        """metrics.py:190-207: standard deviation of the amplitudes, Angstrom (``gd_mean_std``
        on the device array: nothing O(frames) is downloaded)."""
        amp = self._amplitudes_device()
        lib = _lib.lib
        work = dv.empty((int(lib.gd_mean_std_work_bytes()),), torch.uint8)
        mean, std = C.c_double(0.0), C.c_double(0.0)
        _lib.check(lib.gd_mean_std(dv.ptr(amp), amp.numel(), dv.ptr(work), C.byref(mean), C.byref(std),
                                   dv.stream_ptr()))
        return FloatWithUnit(std.value, 'ang')
