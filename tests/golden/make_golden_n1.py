"""Generate tests/golden/n1.npz by EXECUTING THE REFERENCE's own N1 code.

    python tests/golden/make_golden_n1.py          (build container only: needs /root/reference)

Runs ``Transitions.occupancy`` / ``occupancy_by_site_type`` / ``atom_locations`` /
``residence_time`` / ``split`` (+ ``_split_transitions_events``) and ``Jumps.split`` of
/root/reference/src/gemdat/transitions.py:280-410, 837-879 and jumps.py:412-433, unmodified,
on the inputs of the committed pipeline fixtures (``_refshim.py`` says what is stubbed).
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
RefJumps = ref['jumps'].Jumps

CASES = {
    # fixture name -> (split sizes, minimal_residence of the Jumps.split check)
    'pipeline_c2_r09_inner05': ((3, 10), 3),
    'pipeline_c2_dict_radius': ((2, 7), 0),
    'pipeline_box_mixed': ((4,), 0),
}


def load(name):
    with np.load(HERE / f'{name}.npz', allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def main():
    out = {}
    for name, (splits, mr) in CASES.items():
        g = load(name)
        keys = [str(k) for k in g['radius_keys']]
        radius = {k: float(v) for k, v in zip(keys, g['radius_values'])}
        inner = {k: float(v) for k, v in zip(keys, g['inner_values'])}
        species = [str(s) for s in g['species']] if 'species' in g else ['Li'] * g['positions'].shape[1]
        traj = RefTrajectory(species=species, coords=g['positions'].copy(), lattice=np.array(g['lattice']),
                             metadata={'temperature': 650.0}, time_step=2e-15)
        labels = [str(x) for x in g['site_labels']]
        sites = _refshim.Structure(_refshim.Lattice(g['lattice']), ['Li'] * len(labels), g['site_frac'],
                                   labels=labels)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            tr = traj.transitions_between_sites(sites, 'Li', site_radius=radius, site_inner_fraction=inner,
                                                fraction_of_overlap=float(g['fraction_of_overlap']))
        assert np.array_equal(tr.states, g['states'])
        occ = tr.occupancy()
        out[f'{name}/occupancy'] = np.array([site.species.num_atoms for site in occ], dtype=float)
        by_type = tr.occupancy_by_site_type()
        out[f'{name}/by_type_keys'] = np.array(list(by_type.keys()))
        out[f'{name}/by_type_values'] = np.array(list(by_type.values()), dtype=float)
        loc = tr.atom_locations()
        out[f'{name}/locations_keys'] = np.array(list(loc.keys()))
        out[f'{name}/locations_values'] = np.array(list(loc.values()), dtype=float)
        rt = tr.residence_time()
        out[f'{name}/residence'] = rt[['atom index', 'site', 'time']].to_numpy(dtype=np.int64).reshape(-1, 3)
        out[f'{name}/residence_labels'] = np.array([str(x) for x in rt['label']])
        for n in splits:
            parts = tr.split(n)
            for i, part in enumerate(parts):
                tag = f'{name}/split{n}/{i}'
                out[f'{tag}/states'] = part.states
                out[f'{tag}/inner_states'] = part.inner_states
                out[f'{tag}/events'] = part.events.to_numpy().reshape(-1, 6)
                out[f'{tag}/events_index'] = part.events.index.to_numpy()
                out[f'{tag}/n_frames_trajectory'] = np.array(len(part.trajectory))
                out[f'{tag}/occupancy'] = np.array([s.species.num_atoms for s in part.occupancy()], dtype=float)
                prt = part.residence_time()
                out[f'{tag}/residence'] = prt[['atom index', 'site', 'time']].to_numpy(dtype=np.int64).reshape(-1, 3)
                try:
                    j = RefJumps(part, minimal_residence=mr)
                    out[f'{tag}/jumps'] = j.data.to_numpy().reshape(-1, 6)
                except ValueError as exc:
                    out[f'{tag}/jumps_error'] = np.array(str(exc))
            print(name, 'split', n, [len(p.events) for p in parts],
                  [len(out.get(f'{name}/split{n}/{i}/jumps', [])) for i in range(n)])
        out[f'{name}/minimal_residence'] = np.array(mr)
        out[f'{name}/splits'] = np.array(splits)
        # the error path of _split_transitions_events (transitions.py:863-864)
        try:
            tr.split(len(tr.events) + 1)
        except ValueError as exc:
            out[f'{name}/split_error'] = np.array(str(exc))
        print(name, 'occupancy sum', out[f'{name}/occupancy'].sum(), 'residence rows', len(rt), by_type)
    np.savez_compressed(HERE / 'n1.npz', **out)


if __name__ == '__main__':
    main()
