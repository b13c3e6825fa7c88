"""Golden values for the species RDF of a trajectory whose lattice changes from frame to frame
(``constant_lattice=False``), by EXECUTING the reference's own rdf.py:193-268 (per-frame
``Lattice(lat).get_all_distances`` + ``np.histogram``, every frame normalised with its own volume).
Build container only:

    python tests/golden/make_golden_rdf_frames.py   ->  tests/golden/rdf_frames.npz

Positions: the first 60 frames of the stored mixed-species pipeline fixture; lattices: its box
under a seeded per-frame strain (stretch up to 3 %, shear up to 4 %, so the reduced basis and the
general 27-image search are exercised, not only the axis-aligned case), plus one strongly skewed cell.
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
rdf = ref['rdf']
RefTrajectory = ref['trajectory'].Trajectory

N_FRAMES = 60


def lattices_for(base: np.ndarray, seed: int, shear: float) -> np.ndarray:
    rng = np.random.default_rng(seed)
    out = np.empty((N_FRAMES, 3, 3))
    for t in range(N_FRAMES):
        strain = np.eye(3) + np.diag(rng.uniform(-0.03, 0.03, 3))
        strain[0, 1] = rng.uniform(-shear, shear)
        strain[1, 2] = rng.uniform(-shear, shear)
        strain[2, 0] = rng.uniform(-shear, shear)
        out[t] = base @ strain
    return out


def main():
    c = np.load(HERE / 'pipeline_box_mixed.npz')
    species = [str(s) for s in c['species']]
    positions = c['positions'][:N_FRAMES].copy()
    base = np.array(c['lattice'], dtype=float)
    out = {'species': np.array(species), 'positions': positions}
    cases = {
        'diag': lattices_for(base, 1, 0.0),       # axis-aligned cells of changing size
        'sheared': lattices_for(base, 2, 0.04),   # general cells
        'skewed': lattices_for(base @ np.array([[1.0, 0.0, 0.0], [0.9, 1.0, 0.0], [0.0, 0.7, 1.0]]), 3, 0.02),
    }
    for name, lattices in cases.items():
        out[f'{name}_lattices'] = lattices
        for n, (s1, s2, max_dist, resolution) in enumerate([('Li', 'S', 5.0, 0.1), ('Li', 'Li', 4.0, 0.2),
                                                           (['Li', 'S'], 'S', 3.0, 0.25)]):
            traj = RefTrajectory(species=species, coords=positions.copy(), lattice=lattices.copy(),
                                 constant_lattice=False, metadata={'temperature': 650.0}, time_step=2e-15)
            d = rdf.radial_distribution_between_species(trajectory=traj, specie_1=s1, specie_2=s2,
                                                        max_dist=max_dist, resolution=resolution)
            t = f'{name}_between{n}'
            out[f'{t}_specie_1'], out[f'{t}_specie_2'] = np.array(s1), np.array(s2)
            out[f'{t}_max_dist'], out[f'{t}_resolution'] = np.array(max_dist), np.array(resolution)
            out[f'{t}_x'], out[f'{t}_y'], out[f'{t}_label'] = d.x, d.y, np.array(d.label)
            print(t, d.label, d.y[:4], float(d.y.sum()))
    np.savez_compressed(HERE / 'rdf_frames.npz', **out)


if __name__ == '__main__':
    main()
