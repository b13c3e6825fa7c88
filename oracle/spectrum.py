"""Oracle for the attempt frequency: src/gemdat/metrics.py:167-188 and src/gemdat/utils.py:128-160.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

``meanfreq`` calls ``scipy.signal.periodogram(x, fs, axis=-1)`` with its defaults (scipy 1.11.2
pinned, ``requirements_full.txt:153``): boxcar window, constant detrend, density scaling
``1 / (fs N)``, one-sided spectrum with every bin except DC (and Nyquist for even N) doubled.
Restated with numpy's FFT; pinned against goldens produced by executing the reference's own
``meanfreq`` on real scipy (``tests/golden/make_golden_attempt.py``).
"""

from __future__ import annotations

import numpy as np


def meanfreq(x: np.ndarray, fs: float = 1.0) -> np.ndarray:
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    n = x.shape[1]
    spec = np.fft.rfft(x - x.mean(axis=1, keepdims=True), axis=1)
    pxx = (spec.real**2 + spec.imag**2) / (fs * n)
    if n % 2 == 0:
        pxx[:, 1:-1] *= 2.0
    else:
        pxx[:, 1:] *= 2.0
    f = np.fft.rfftfreq(n, 1.0 / fs)
    p = pxx * (f[1] - f[0])
    with np.errstate(invalid='ignore', divide='ignore'):
        return (p @ f.reshape(-1, 1)) / p.sum(axis=1, keepdims=True)


def attempt_frequency(speed: np.ndarray, fs: float) -> tuple[float, float]:
    """metrics.py:177-186: mean and (population) standard deviation over atoms, Hz."""
    fm = meanfreq(speed, fs)
    return float(np.mean(fm)), float(np.std(fm))
