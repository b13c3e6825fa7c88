"""TrajectoryMetrics: the displacement-derived scalars on the hot path.

Mirror of reference src/gemdat/metrics.py:21-231 for ``speed`` (:35-47),
``tracer_diffusivity`` (:76-97), ``amplitudes`` (:209-231), ``vibration_amplitude``
(:190-207) and ``particle_density`` (:49-62).  The per-frame distances come from the K0
kernel (``gd_unwrap``) instead of the reference's Python loop over frames; the
amplitude run-length bookkeeping is O(sign flips) host work on the returned array.
``attempt_frequency`` (:167-188) runs on the device array too: the
per-atom spectrum comes from cuFFT through ``torch.fft`` (a library transform, not one of
this package's kernels), the periodogram weighting and the mean-frequency reductions of
``utils.py:128-160`` follow it on the device.  Conductivity, Haven ratio and the Arrhenius
helpers are out of scope (DESIGN.md "next rows").
"""

from __future__ import annotations

import numpy as np
import torch

from . import _device as dv
from .caching import weak_lru_cache
from .lattice import FloatWithUnit

ANGSTROM = 1e-10


class TrajectoryMetrics:
    def __init__(self, trajectory):
        self.trajectory = trajectory

    @weak_lru_cache()
    def _distances(self) -> np.ndarray:
        return self.trajectory.distances_from_base_position()

    @weak_lru_cache()
    def speed(self) -> np.ndarray:
        """metrics.py:35-47: change of the distance from the base position per step, [A, T]."""
        return np.diff(self._distances(), prepend=0)

    @weak_lru_cache()
    def _meanfreq(self) -> np.ndarray:
        """utils.py:128-160 on the device: mean frequency of every atom's speed, [A, 1].

        ``scipy.signal.periodogram`` defaults restated: constant detrend, boxcar window,
        density scaling 1/(fs N), one-sided spectrum with every bin but DC (and Nyquist for
        even N) doubled; ``mnfreq = sum(P f) / sum(P)`` with ``P = Pxx (f[1] - f[0])``.
        """
        dv.require_cuda()
        dist = self.trajectory._unwrap(want_dist=True, keep_on_device=True)['dist']  # [A, T] device
        n = int(dist.shape[1])
        fs = float(self.trajectory.sampling_frequency)
        speed = torch.cat([dist[:, :1], dist[:, 1:] - dist[:, :-1]], dim=1)  # np.diff(prepend=0)
        del dist
        x = speed - speed.mean(dim=1, keepdim=True)
        del speed
        spec = torch.fft.rfft(x, dim=1)
        del x
        pxx = (spec.real * spec.real + spec.imag * spec.imag) * (1.0 / (fs * n))
        del spec
        if n % 2 == 0:
            pxx[:, 1:-1] *= 2.0
        else:
            pxx[:, 1:] *= 2.0
        f = torch.fft.rfftfreq(n, d=1.0 / fs, dtype=torch.float64, device=pxx.device)
        width = (f[1] - f[0]) if n > 1 else f.new_tensor(float('nan'))
        p = pxx * width
        mnfreq = (p @ f.reshape(-1, 1)) / p.sum(dim=1, keepdim=True)
        return dv.to_host(mnfreq)

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
        """metrics.py:209-231: sums of ``speed`` over runs of equal sign.

        The reference splits each atom's speed at every sign flip except the first and
        the last one (``splits[1:-1]``) and sums each piece.
        """
        out = []
        for row in self.speed():
            signs = np.sign(row)
            flips = np.nonzero(signs != np.roll(signs, shift=-1))[0]
            cuts = flips[1:-1] + 1
            starts = np.concatenate([[0], cuts])
            if len(row):
                out.append(np.add.reduceat(row, starts))
        return np.concatenate(out) if out else np.zeros(0)

    @weak_lru_cache()
    def vibration_amplitude(self) -> FloatWithUnit:
        """metrics.py:190-207: standard deviation of the amplitudes, Angstrom."""
        return FloatWithUnit(np.std(self.amplitudes()), 'ang')
