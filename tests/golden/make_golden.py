"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE's own modules.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

See ``_refshim.py`` for what is real reference code and what is stubbed: GEMDAT's own
Python (state assignment control flow, event extraction, the pandas jump state
machine, matrices, diffusivities, metrics, FFT MSD, density histogram) runs
unmodified; pymatgen / MDAnalysis are replaced by restatements [recalled] (the ball
query underneath is real scipy).  The fixtures hold inputs AND outputs, so the tests
need neither the reference nor this script.
"""

from __future__ import annotations

import sys
import warnings
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parents[1]))

import _refshim  # noqa: E402

ref = _refshim.load_reference()
RefTrajectory = ref['trajectory'].Trajectory
RefTransitions = ref['transitions'].Transitions
RefJumps = ref['jumps'].Jumps
RefMetrics = ref['metrics'].TrajectoryMetrics


def ref_structure(lattice, frac, labels):
    return _refshim.Structure(_refshim.Lattice(lattice), ['Li'] * len(frac), frac, labels=labels)


def ref_trajectory(positions, lattice, species=None, time_step=2e-15):
    species = species or ['Li'] * positions.shape[1]
    return RefTrajectory(species=species, coords=positions.copy(), lattice=np.array(lattice),
                         metadata={'temperature': 650.0}, time_step=time_step)


def run_pipeline(name, *, positions, lattice, site_frac, site_labels, site_radius, inner_fraction,
                 fraction_of_overlap=0.0, residences=(0, 3, 100), species=None, floating='Li'):
    traj = ref_trajectory(positions, lattice, species)
    sites = ref_structure(lattice, site_frac, site_labels)
    out = {
        'positions': positions, 'lattice': np.array(lattice, dtype=float), 'site_frac': site_frac,
        'site_labels': np.array(site_labels), 'fraction_of_overlap': fraction_of_overlap,
    }
    if species is not None:
        out['species'] = np.array(species)
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter('always')
        tr = traj.transitions_between_sites(
            sites, floating, site_radius=site_radius, site_inner_fraction=inner_fraction,
            fraction_of_overlap=fraction_of_overlap,
        )
    out['warnings'] = np.array([str(w.message) for w in caught])
    out['radius_keys'] = np.array(list(tr.site_radius.radius.keys()))
    out['radius_values'] = np.array(list(tr.site_radius.radius.values()), dtype=float)
    out['inner_values'] = np.array([tr.site_radius.inner_fraction[k] for k in tr.site_radius.radius],
                                   dtype=float)
    out['states'] = tr.states
    out['inner_states'] = tr.inner_states
    out['events'] = tr.events.to_numpy()
    out['event_columns'] = np.array(list(tr.events.columns))
    out['matrix'] = tr.matrix()
    out['states_next'] = tr.states_next()
    out['states_prev'] = tr.states_prev()
    for mr in residences:
        try:
            j = RefJumps(tr, minimal_residence=mr)
        except ValueError as exc:
            out[f'jumps_{mr}_error'] = np.array(str(exc))
            continue
        out[f'jumps_{mr}'] = j.data.to_numpy()
        out['jump_columns'] = np.array(list(j.data.columns))
        out[f'jumps_{mr}_matrix'] = j.matrix()
        out[f'jumps_{mr}_diffusivity'] = np.array(float(j.jump_diffusivity(3)))
    np.savez_compressed(HERE / f'{name}.npz', **out)
    print(f'{name}: states {tr.states.shape}, events {len(tr.events)}, '
          f'jumps { {mr: len(out[f"jumps_{mr}"]) for mr in residences if f"jumps_{mr}" in out} }, '
          f'radius {dict(tr.site_radius.radius)}, warnings {len(caught)}')


def main():
    from gemdat_b200 import synth  # recipe + host replay only (pure numpy)

    # ---- 1/2: argyrodite 2x1x1, given radius and default (vibration amplitude) radius
    rec = synth.config_c2(n_frames=1200, seed=11)
    pos = rec.host_positions()
    frac, labels = rec.sites.frac_coords, rec.sites.labels
    run_pipeline('pipeline_c2_r09_inner05', positions=pos, lattice=rec.lattice, site_frac=frac,
                 site_labels=labels, site_radius=0.9, inner_fraction=0.5)
    run_pipeline('pipeline_c2_r09_inner10', positions=pos, lattice=rec.lattice, site_frac=frac,
                 site_labels=labels, site_radius=0.9, inner_fraction=1.0)
    run_pipeline('pipeline_c2_default_radius', positions=pos[:600], lattice=rec.lattice,
                 site_frac=frac, site_labels=labels, site_radius=None, inner_fraction=1.0)

    # ---- 3: per-label radii (reference tests/integration/transitions_test.py:32-47 pattern);
    #         short trajectory so that some sites are never visited -> integer_remap quirk
    lab3 = list(labels)
    for i in range(len(lab3)):
        lab3[i] = 'ABC'[i % 3]
    run_pipeline('pipeline_c2_dict_radius', positions=pos[:400], lattice=rec.lattice, site_frac=frac,
                 site_labels=lab3, site_radius={'A': 0.5, 'B': 0.6, 'C': 0.7},
                 inner_fraction={'A': 0.8, 'B': 1.0, 'C': 0.5})

    # ---- 4: overlapping spheres allowed (tie rule: last write wins)
    run_pipeline('pipeline_c2_overlap', positions=pos[:400], lattice=rec.lattice, site_frac=frac,
                 site_labels=lab3, site_radius={'C': 1.3, 'A': 1.2, 'B': 1.25},
                 inner_fraction={'C': 0.9, 'A': 0.6, 'B': 1.0}, fraction_of_overlap=1.0)

    # ---- 5: small orthorhombic box, uniform random atoms (faces, edges, corners, exact 0)
    rng = np.random.default_rng(5)
    lat = np.diag([7.31, 5.07, 6.43])
    fixed = [[0.0, 0.0, 0.0], [0.999999, 0.5, 1e-9],
             [1.25, -0.3, 2.0000001]]  # last one outside [0,1): wrapped by apply_PBC
    sfrac = [np.array(f) for f in fixed]
    while len(sfrac) < 24:  # spheres may overlap by up to half (fraction_of_overlap=1)
        cand = rng.random(3)
        d = np.array(sfrac) - cand
        d -= np.round(d)
        if np.min(np.linalg.norm(d @ lat, axis=1)) >= 1.15:
            sfrac.append(cand)
    sfrac = np.array(sfrac)
    p5 = rng.random((200, 30, 3))
    p5[0, 0] = [0.0, 0.0, 0.0]
    p5[0, 1] = [1.0 - 2**-53, 0.5, 0.5]
    p5[0, 2] = [0.9999999, 0.9999999, 0.9999999]
    p5[0, 3] = [1e-12, 1e-12, 1e-12]
    p5[1:50, 4] = sfrac[0] + rng.normal(scale=0.02, size=(49, 3))
    p5 = np.mod(p5, 1)
    slab = ['P' if i % 2 else 'Q' for i in range(len(sfrac))]
    run_pipeline('pipeline_box_mixed', positions=p5, lattice=lat, site_frac=sfrac, site_labels=slab,
                 site_radius={'Q': 1.1, 'P': 0.8}, inner_fraction={'Q': 0.7, 'P': 1.0},
                 fraction_of_overlap=1.0, residences=(0, 2),
                 species=['Li'] * 20 + ['S'] * 5 + ['Li'] * 5)

    # ---- 6: displacement / metrics / MSD / volume on the reference's own unit fixture and a walk
    coords = np.array(
        [[[0.2, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5]],
         [[0.4, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5]],
         [[0.6, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5]],
         [[0.8, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5]],
         [[0.1, 0.0, 0.0], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5], [0.0, 0.0, 0.5]]])
    out = {}
    for tag, p, lattice, ts in (
        ('unit', coords, np.eye(3), 1.0),
        ('walk', np.mod(0.3 + np.cumsum(rng.normal(scale=0.03, size=(700, 6, 3)), axis=0), 1),
         np.diag([9.0, 10.5, 12.25]), 2e-15),
        ('c2', pos[:500], rec.lattice, 2e-15),
    ):
        t = ref_trajectory(p, lattice, time_step=ts)
        out[f'{tag}_positions'] = p
        out[f'{tag}_lattice'] = np.array(lattice, dtype=float)
        out[f'{tag}_time_step'] = np.array(ts)
        out[f'{tag}_cumulative'] = ref_trajectory(p, lattice, time_step=ts).cumulative_displacements
        out[f'{tag}_distances'] = ref_trajectory(p, lattice, time_step=ts).distances_from_base_position()
        out[f'{tag}_msd'] = ref_trajectory(p, lattice, time_step=ts).mean_squared_displacement()
        m = RefMetrics(t)
        out[f'{tag}_speed'] = m.speed()
        out[f'{tag}_amplitudes'] = m.amplitudes()
        out[f'{tag}_vibration_amplitude'] = np.array(float(m.vibration_amplitude()))
        out[f'{tag}_tracer_diffusivity'] = np.array(float(m.tracer_diffusivity(dimensions=3)))
        for res in (0.2, 0.5) if tag != 'unit' else (0.2,):
            vol = ref['volume'].trajectory_to_volume(ref_trajectory(p, lattice, time_step=ts), resolution=res)
            out[f'{tag}_volume_{res}'] = vol.data
    np.savez_compressed(HERE / 'displacement_metrics.npz', **out)
    print('displacement_metrics:', {k: v.shape for k, v in out.items() if k.endswith(('msd', 'amplitudes'))})

    # ---- 7: ffill / bfill / integer_remap / transitions matrix on the reference's unit-test inputs
    u = ref['utils']
    arr = np.array([[5, -1, -1, 7, 2], [3, -1, 1, 8, -1], [4, 9, 6, -1, -1]])
    rnd = rng.integers(-1, 6, size=(64, 37))
    rnd[rng.random(rnd.shape) < 0.5] = -1
    import pandas as pd

    ev = pd.DataFrame(
        data=np.array([[0, 0, 1, 10], [0, 1, 0, 20], [1, 2, 1, 5], [1, 1, 0, 6], [2, 2, 0, 30]]),
        columns=['atom index', 'start site', 'destination site', 'time'])
    np.savez_compressed(
        HERE / 'utils.npz',
        arr=arr, ffill=u.ffill(arr), bfill=u.bfill(arr), ffill0=u.ffill(arr, axis=0),
        bfill0=u.bfill(arr, axis=0), rnd=rnd, rnd_ffill0=u.ffill(rnd, axis=0),
        rnd_bfill0=u.bfill(rnd, axis=0), rnd_ffill1=u.ffill(rnd), rnd_bfill1=u.bfill(rnd),
        remap=u.integer_remap(np.array([4, 2, 1, 3]), key=np.array([10, 20, 30, 40])),
        matrix_events=ev.to_numpy(),
        matrix=ref['transitions']._calculate_transitions_matrix(ev, n_sites=3),
    )
    print('utils: done')


if __name__ == '__main__':
    main()
