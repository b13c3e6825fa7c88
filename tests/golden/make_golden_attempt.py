"""Golden values for TrajectoryMetrics.attempt_frequency by EXECUTING the
reference's own metrics.py:167-188 and utils.py:128-160 (real scipy.signal.periodogram) on the
trajectories already stored in displacement_metrics.npz.  Build container only:

    python tests/golden/make_golden_attempt.py   ->  tests/golden/attempt_frequency.npz
"""

from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parents[1]))

import _refshim  # noqa: E402

ref = _refshim.load_reference()
RefTrajectory = ref['trajectory'].Trajectory
RefMetrics = ref['metrics'].TrajectoryMetrics


def main():
    src = np.load(HERE / 'displacement_metrics.npz')
    out = {}
    for tag in ('unit', 'walk', 'c2'):
        p, lattice, ts = src[f'{tag}_positions'], src[f'{tag}_lattice'], float(src[f'{tag}_time_step'])
        t = RefTrajectory(species=['Li'] * p.shape[1], coords=p.copy(), lattice=np.array(lattice),
                          metadata={'temperature': 650.0}, time_step=ts)
        m = RefMetrics(t)
        freq, std = m.attempt_frequency()
        out[f'{tag}_attempt_frequency'] = np.array(float(freq))
        out[f'{tag}_attempt_frequency_std'] = np.array(float(std))
        out[f'{tag}_meanfreq'] = ref['utils'].meanfreq(m.speed(), fs=t.sampling_frequency)
    np.savez_compressed(HERE / 'attempt_frequency.npz', **out)
    print({k: (v.shape, float(np.ravel(v)[0])) for k, v in out.items()})


if __name__ == '__main__':
    main()
