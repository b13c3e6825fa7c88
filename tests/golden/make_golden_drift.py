"""Generate tests/golden/drift.npz by EXECUTING THE REFERENCE's own code.

    python tests/golden/make_golden_drift.py        (build container only: needs /root/reference)

Runs ``Trajectory.drift`` / ``apply_drift_correction`` / ``center_of_mass`` of
/root/reference/src/gemdat/trajectory.py:732-830 and ``TrajectoryMetrics.mol_per_liter`` /
``tracer_diffusivity_center_of_mass`` / ``haven_ratio`` / ``tracer_conductivity`` of
metrics.py:64-165, unmodified, on a seeded mixed-species random walk with a common drift
(``_refshim.py`` says what is stubbed; atomic masses come from the product's own table).

``floating_species=...`` comes out as NaN: trajectory.py:781-787 collects Species OBJECTS and
``filter`` then tests ``sp.symbol in species`` (a str against a set of objects), which never
matches, so the mean runs over zero atoms.  The NaN arrays are kept in the file as the record of
that; the product follows the docstring instead (tests compare with the equivalent
``fixed_species`` selection).
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

SPECIES = ['Li'] * 7 + ['P'] * 3 + ['S'] * 9 + ['Br'] * 2
LATTICE = np.array([[9.9, 0.0, 0.0], [0.3, 10.4, 0.0], [-0.2, 0.5, 11.1]])


def walk(seed=7, n_frames=400):
    rng = np.random.default_rng(seed)
    n = len(SPECIES)
    start = rng.random((n, 3))
    steps = rng.normal(scale=0.004, size=(n_frames, n, 3))
    steps[:, :7] *= 6.0  # the floating species moves more
    common = rng.normal(scale=0.002, size=(n_frames, 1, 3)) + np.array([0.0015, -0.001, 0.0005])
    steps = steps + common
    steps[0] = 0.0
    return np.mod(start + np.cumsum(steps, axis=0), 1.0)


def main():
    coords = walk()
    out = {'coords': coords, 'lattice': LATTICE, 'species': np.array(SPECIES), 'time_step': np.array(2e-15),
           'temperature': np.array(700.0)}

    def traj():
        return RefTrajectory(species=SPECIES, coords=coords.copy(), lattice=LATTICE.copy(),
                             metadata={'temperature': 700.0}, time_step=2e-15)

    selections = {
        'all': {},
        'fixed_P_S': {'fixed_species': ('P', 'S')},
        'fixed_Br': {'fixed_species': 'Br'},
        'fixed_P_S_Br': {'fixed_species': ['P', 'S', 'Br']},
        'floating_Li': {'floating_species': 'Li'},
        'floating_Li_Br': {'floating_species': ('Li', 'Br')},
    }
    for tag, kw in selections.items():
        out[f'drift/{tag}'] = traj().drift(**kw)
        corrected = traj().apply_drift_correction(**kw)
        out[f'corrected/{tag}/positions'] = np.array(corrected.positions)
        out[f'corrected/{tag}/displacements'] = np.array(traj().apply_drift_correction(**kw).displacements)
        m = RefMetrics(corrected.filter('Li'))
        out[f'corrected/{tag}/tracer_diffusivity_Li'] = np.array(float(m.tracer_diffusivity(dimensions=3)))
        print(tag, out[f'drift/{tag}'].shape, float(m.tracer_diffusivity(dimensions=3)))
    com = traj().center_of_mass()
    out['com/coords'] = np.array(com.coords)
    out['com/positions'] = np.array(traj().center_of_mass().positions)
    li = traj().filter('Li')
    out['com_Li/coords'] = np.array(li.center_of_mass().coords)
    m = RefMetrics(traj().filter('Li'))
    out['metrics/mol_per_liter'] = np.array(float(m.mol_per_liter()))
    out['metrics/tracer_diffusivity'] = np.array(float(m.tracer_diffusivity(dimensions=3)))
    out['metrics/tracer_diffusivity_com'] = np.array(float(m.tracer_diffusivity_center_of_mass(dimensions=3)))
    out['metrics/haven_ratio'] = np.array(float(m.haven_ratio(dimensions=3)))
    out['metrics/tracer_conductivity'] = np.array(float(m.tracer_conductivity(z_ion=1, dimensions=3)))
    out['metrics/tracer_conductivity_2d'] = np.array(float(m.tracer_conductivity(z_ion=2, dimensions=2)))
    print({k: float(v) for k, v in out.items() if k.startswith('metrics/')})
    np.savez_compressed(HERE / 'drift.npz', **out)


if __name__ == '__main__':
    main()
