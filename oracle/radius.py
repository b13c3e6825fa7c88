"""Oracle for SiteRadius: src/gemdat/transitions.py:485-733.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""

from __future__ import annotations

from itertools import combinations_with_replacement

import numpy as np

from .pmg import Lattice


def radius_to_dict(radius, inner_fraction):
    """transitions.py:713-733."""
    if isinstance(radius, (int, float)):
        if isinstance(inner_fraction, dict):
            r = {label: radius for label in inner_fraction.keys()}
        elif isinstance(inner_fraction, (int, float)):
            r = {'': radius}
    elif isinstance(radius, dict):
        r = radius
    else:
        raise TypeError(f'Invalid type for `site_radius`: {type(radius)}')
    if isinstance(inner_fraction, (int, float)):
        f = {label: inner_fraction for label in r.keys()}
    elif isinstance(inner_fraction, dict):
        f = inner_fraction
    else:
        raise TypeError(f'Invalid type for `site_inner_fraction`: {type(inner_fraction)}')
    return r, f


def radius_from_vibration_amplitude(site_frac, lattice_matrix, vibration_amplitude: float) -> float:
    """transitions.py:590-629: 2*amplitude, clipped to half the closest site pair - 0.005 A."""
    pdist = Lattice(lattice_matrix).get_all_distances(site_frac, site_frac)
    min_dist = np.min(pdist[np.triu_indices_from(pdist, k=1)])
    site_radius = 2 * vibration_amplitude
    if min_dist < 2 * site_radius:
        site_radius = (0.5 * min_dist) - 0.005
        if site_radius * 2 < 0.5:
            raise ValueError('Two sites are within half an Angstrom of each other.')
    return float(site_radius)


def sites_are_overlapping(site_frac, site_labels, lattice_matrix, radius: dict, factor: float = 1.0) -> bool:
    """transitions.py:648-685."""
    pdist = Lattice(lattice_matrix).get_all_distances(site_frac, site_frac)
    labels = np.array(list(site_labels))
    for i, j in combinations_with_replacement(radius.keys(), 2):
        ii = np.where(labels == i)[0]
        jj = np.where(labels == j)[0]
        if ii.size == 0 or jj.size == 0:
            continue
        sub = pdist[np.ix_(ii, jj)].copy()
        _, a, b = np.intersect1d(ii, jj, return_indices=True)
        if a.size:
            sub[a, b] = np.inf
        if factor * float(sub.min(initial=np.inf)) < radius[i] + radius[j]:
            return True
    return False
