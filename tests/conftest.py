from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / 'golden'

PIPELINE_CASES = [
    'pipeline_c2_r09_inner05',
    'pipeline_c2_r09_inner10',
    'pipeline_c2_default_radius',
    'pipeline_c2_dict_radius',
    'pipeline_c2_overlap',
    'pipeline_box_mixed',
]


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


def load_golden(name: str) -> dict:
    with np.load(GOLDEN / f'{name}.npz', allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def load_pipeline_case(name: str) -> dict:
    g = load_golden(name)
    g['name'] = name
    keys = [str(k) for k in g['radius_keys']]
    g['site_radius'] = {k: float(v) for k, v in zip(keys, g['radius_values'])}
    g['inner_fraction'] = {k: float(v) for k, v in zip(keys, g['inner_values'])}
    g['labels'] = [str(x) for x in g['site_labels']]
    if 'species' in g:
        species = [str(s) for s in g['species']]
        g['floating_mask'] = np.array([s == 'Li' for s in species])
    else:
        g['floating_mask'] = np.ones(g['positions'].shape[1], dtype=bool)
    return g


@pytest.fixture(params=PIPELINE_CASES)
def pipeline_case(request):
    return load_pipeline_case(request.param)


@pytest.fixture(scope='session')
def displacement_golden():
    return load_golden('displacement_metrics')


@pytest.fixture(scope='session')
def utils_golden():
    return load_golden('utils')
