"""Generate tests/golden/rates.npz by EXECUTING THE REFERENCE's own code.

    python tests/golden/make_golden_rates.py        (build container only: needs /root/reference)

Runs ``Jumps.rates`` / ``activation_energies`` / ``to_graph`` / ``activation_energy_between_sites``
of /root/reference/src/gemdat/jumps.py:253-308, 340-462, unmodified (with everything under them:
``Transitions.split`` / ``atom_locations`` / ``occupancy``, ``Jumps.split`` / ``counter``,
``TrajectoryMetrics.attempt_frequency`` with the real scipy periodogram), on the inputs of the
committed pipeline fixtures (``_refshim.py`` says what is stubbed).
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
    # fixture name -> (minimal_residence, part counts to try)
    'pipeline_c2_r09_inner05': (3, (2, 3, 10)),
    'pipeline_c2_dict_radius': (0, (2, 4)),
    'pipeline_box_mixed': (0, (2, 3)),
}
TEMPERATURE = 650.0


def load(name):
    with np.load(HERE / f'{name}.npz', allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def table(df):
    keys = np.array([list(k) for k in df.index])
    return keys, df.to_numpy(dtype=float)


def main():
    out = {'temperature': np.array(TEMPERATURE)}
    for name, (mr, part_counts) in CASES.items():
        g = load(name)
        keys = [str(k) for k in g['radius_keys']]
        radius = {k: float(v) for k, v in zip(keys, g['radius_values'])}
        inner = {k: float(v) for k, v in zip(keys, g['inner_values'])}
        species = [str(s) for s in g['species']] if 'species' in g else ['Li'] * g['positions'].shape[1]
        traj = RefTrajectory(species=species, coords=g['positions'].copy(), lattice=np.array(g['lattice']),
                             metadata={'temperature': TEMPERATURE}, time_step=2e-15)
        labels = [str(x) for x in g['site_labels']]
        sites = _refshim.Structure(_refshim.Lattice(g['lattice']), ['Li'] * len(labels), g['site_frac'],
                                   labels=labels)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            tr = traj.transitions_between_sites(sites, 'Li', site_radius=radius, site_inner_fraction=inner,
                                                fraction_of_overlap=float(g['fraction_of_overlap']))
            jumps = RefJumps(tr, minimal_residence=mr)
            out[f'{name}/minimal_residence'] = np.array(mr)
            out[f'{name}/part_counts'] = np.array(part_counts)
            freq, freq_std = ref['metrics'].TrajectoryMetrics(traj.filter('Li')).attempt_frequency()
            out[f'{name}/attempt_frequency'] = np.array([float(freq), float(freq_std)])
            for n in part_counts:
                try:
                    k, v = table(jumps.rates(n_parts=n))
                    out[f'{name}/rates{n}/keys'], out[f'{name}/rates{n}/values'] = k, v
                    k, v = table(jumps.activation_energies(n_parts=n))
                    out[f'{name}/eact{n}/keys'], out[f'{name}/eact{n}/values'] = k, v
                    print(name, n, 'rates', v.shape)
                except ValueError as exc:
                    out[f'{name}/parts{n}/error'] = np.array(str(exc))
                    print(name, n, 'error', exc)
            graph = jumps.to_graph()
            edges = np.array([(a, b, d['e_act']) for a, b, d in graph.edges(data=True)], dtype=float).reshape(-1, 3)
            out[f'{name}/graph_edges'] = edges
            out[f'{name}/graph_nodes'] = np.array([d['label'] for _, d in graph.nodes(data=True)])
            lo, hi = np.quantile(edges[:, 2], [0.25, 0.75])
            cut = jumps.to_graph(min_e_act=float(lo), max_e_act=float(hi))
            out[f'{name}/graph_cut'] = np.array([lo, hi])
            out[f'{name}/graph_cut_edges'] = np.array(
                [(a, b, d['e_act']) for a, b, d in cut.edges(data=True)], dtype=float).reshape(-1, 3)
            a, b = int(edges[0, 0]), int(edges[0, 1])
            out[f'{name}/between'] = np.array([a, b, jumps.activation_energy_between_sites(a, b)], dtype=float)
            print(name, 'edges', len(edges), 'cut', len(out[f'{name}/graph_cut_edges']))
    np.savez_compressed(HERE / 'rates.npz', **out)


if __name__ == '__main__':
    main()
