"""Golden values for collective jumps by EXECUTING the reference's own
collective.py:61-106 (pandas double loop) on jump tables stored in the pipeline fixtures.
Build container only:

    python tests/golden/make_golden_collective.py   ->  tests/golden/collective.npz
"""

from __future__ import annotations

import sys
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import pandas as pd

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parents[1]))

import _refshim  # noqa: E402

ref = _refshim.load_reference()
RefCollective = ref['collective'].Collective

CASES = [  # (fixture, jumps key, max_steps, max_dist)
    ('pipeline_c2_r09_inner05', 'jumps_0', 30, 1.0),
    ('pipeline_c2_r09_inner05', 'jumps_0', 400, 4.5),
    ('pipeline_c2_r09_inner10', 'jumps_3', 150, 3.0),
    ('pipeline_box_mixed', 'jumps_0', 60, 2.5),
]


def main():
    out = {}
    for n, (fixture, key, max_steps, max_dist) in enumerate(CASES):
        c = np.load(HERE / f'{fixture}.npz')
        data = pd.DataFrame(c[key], columns=[str(x) for x in c['jump_columns']])
        labels = [str(x) for x in c['site_labels']]
        sites = _refshim.Structure(_refshim.Lattice(c['lattice']), ['Li'] * len(labels), c['site_frac'],
                                   labels=labels)
        coll = RefCollective(jumps=SimpleNamespace(data=data), sites=sites, lattice=_refshim.Lattice(c['lattice']),
                             max_steps=max_steps, max_dist=max_dist)
        ordered = data.sort_values(['stop time', 'start time'], ignore_index=True)
        # index pairs (i, j) into the sorted table, in the reference's emission order
        pairs = np.array([(int(ei.name), int(ej.name)) for ei, ej in coll.collective], dtype=np.int64).reshape(-1, 2)
        t = f'case{n}'
        out[f'{t}_fixture'] = np.array(fixture)
        out[f'{t}_jumps_key'] = np.array(key)
        out[f'{t}_max_steps'] = np.array(max_steps)
        out[f'{t}_max_dist'] = np.array(max_dist)
        out[f'{t}_sorted_jumps'] = ordered.to_numpy()
        out[f'{t}_pairs'] = pairs
        out[f'{t}_coll_jumps'] = np.array(coll.coll_jumps, dtype=np.int64).reshape(-1, 2, 2)
        out[f'{t}_n_solo_jumps'] = np.array(int(coll.n_solo_jumps))
        out[f'{t}_n_coll_jumps'] = np.array(int(coll.n_coll_jumps))
        # collective.py:108-166; the label-pair order comes from a set, so it is stored too
        out[f'{t}_site_pair_labels'] = np.array(coll.site_pair_count_matrix_labels(), dtype=str).reshape(-1, 2)
        out[f'{t}_site_pair_count_matrix'] = coll.site_pair_count_matrix()
        if len(pairs):
            mj, mc = coll.multiple_collective()
            out[f'{t}_multiple_jumps'] = np.stack([mj['start'], mj['stop']], axis=-1).astype(np.int64)
            out[f'{t}_multiple_counts'] = mc.astype(np.int64)
        print(t, fixture, key, 'jumps', len(data), 'pairs', len(pairs), 'solo', int(coll.n_solo_jumps))
    out['n_cases'] = np.array(len(CASES))
    np.savez_compressed(HERE / 'collective.npz', **out)


if __name__ == '__main__':
    main()
