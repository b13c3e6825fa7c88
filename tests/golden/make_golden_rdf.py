"""Golden values for the state-resolved RDF by EXECUTING the reference's own
rdf.py:111-268 on the stored mixed-species pipeline fixture.  Build container only:

    python tests/golden/make_golden_rdf.py   ->  tests/golden/rdf.npz

rdf.py builds its label list with ``list(set(labels))``, whose order depends on the process'
string hashing; the order used in this run is stored (``unique_labels``) because the name of the
"~>" states depends on it (rdf.py:40-45 indexes that list with -1).
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
rdf = ref['rdf']
RefTrajectory = ref['trajectory'].Trajectory


def main():
    c = np.load(HERE / 'pipeline_box_mixed.npz')
    species = [str(s) for s in c['species']]
    labels = [str(x) for x in c['site_labels']]
    traj = RefTrajectory(species=species, coords=c['positions'].copy(), lattice=np.array(c['lattice']),
                         metadata={'temperature': 650.0}, time_step=2e-15)
    sites = _refshim.Structure(_refshim.Lattice(c['lattice']), ['Li'] * len(labels), c['site_frac'], labels=labels)
    keys = [str(k) for k in c['radius_keys']]
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(
            sites, 'Li', site_radius={k: float(v) for k, v in zip(keys, c['radius_values'])},
            site_inner_fraction={k: float(v) for k, v in zip(keys, c['inner_values'])},
            fraction_of_overlap=float(c['fraction_of_overlap']))
    assert np.array_equal(tr.states, c['states'])
    out = {'fixture': np.array('pipeline_box_mixed'), 'unique_labels': np.array(list(set(labels)))}
    out['uniqify_states'] = rdf._uniqify_labels(tr.states, labels)
    out['states_array'] = rdf._get_states_array(tr, labels)
    s2s = rdf._get_states(labels)
    out['states2str_codes'] = np.array(list(s2s.keys()), dtype=np.int64)
    out['states2str_names'] = np.array(list(s2s.values()))
    for n, (max_dist, resolution) in enumerate([(5.0, 0.1), (3.3, 0.25)]):
        res = rdf.radial_distribution(transitions=tr, floating_specie='Li', max_dist=max_dist, resolution=resolution)
        t = f'rdf{n}'
        out[f'{t}_max_dist'], out[f'{t}_resolution'] = np.array(max_dist), np.array(resolution)
        flat = [(state, d.label, d.x, d.y) for state, coll in res.items() for d in coll]
        out[f'{t}_states'] = np.array([f[0] for f in flat])
        out[f'{t}_symbols'] = np.array([f[1] for f in flat])
        out[f'{t}_x'] = flat[0][2]
        out[f'{t}_y'] = np.stack([f[3] for f in flat]).astype(np.int64)
        print(t, 'curves', len(flat), 'states', list(res.keys()), 'total', int(out[f'{t}_y'].sum()))
    for n, (s1, s2, max_dist, resolution) in enumerate([('Li', 'S', 5.0, 0.1), ('Li', 'Li', 4.0, 0.2),
                                                       (['Li', 'S'], 'S', 5.0, 0.1)]):
        d = rdf.radial_distribution_between_species(trajectory=traj, specie_1=s1, specie_2=s2, max_dist=max_dist,
                                                    resolution=resolution)
        t = f'between{n}'
        out[f'{t}_specie_1'] = np.array(s1)
        out[f'{t}_specie_2'] = np.array(s2)
        out[f'{t}_max_dist'], out[f'{t}_resolution'] = np.array(max_dist), np.array(resolution)
        out[f'{t}_x'], out[f'{t}_y'], out[f'{t}_label'] = d.x, d.y, np.array(d.label)
        print(t, d.label, d.y[:4], float(d.y.sum()))
    np.savez_compressed(HERE / 'rdf.npz', **out)


if __name__ == '__main__':
    main()
