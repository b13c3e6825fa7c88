"""Oracle for site-state assignment: src/gemdat/transitions.py:736-809.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Two independent restatements that must agree bit-for-bit:

* ``calculate_atom_states`` follows the reference line by line on top of the restated
  ``PeriodicKDTree`` (``oracle.mda``; the ball query itself is real scipy cKDTree code);
* ``calculate_atom_states_bruteforce`` evaluates the same float32-staged geometry
  for every (site, atom-frame) pair with plain numpy, no tree, no image list.

Semantics that are observable in the output:
``any sphere that contains the atom, last write wins`` (highest site index in a
label group; later label groups overwrite earlier ones), ``integer_remap`` by rank
among the *visited* sites of a group, float32 rounding of Cartesian coordinates,
box lengths and periodic images, float64 ``dx*dx+dy*dy+dz*dz <= r*r``.
"""

from __future__ import annotations

import warnings

import numpy as np

from . import NOSITE
from .fill import integer_remap
from .mda import PeriodicKDTree, augment_ortho_f32, check_ortho_box, ortho_pbc_f32
from .pmg import Lattice


def label_groups(site_labels, site_radius: dict):
    """Yield ``(label, radius, key)`` in ``site_radius`` dict order.

    ``key`` is None for the catch-all label ``''`` (transitions.py:786-793), else the
    ascending global indices of the sites carrying that label.
    """
    labels = np.asarray(list(site_labels), dtype=object)
    for label, radius in site_radius.items():
        if label:
            key = np.nonzero(labels == label)[0]
            if key.size == 0:
                # the reference does zip(*empty) -> ValueError
                raise ValueError('not enough values to unpack (expected 2, got 0)')
            yield label, radius, key
        else:
            yield label, radius, None


def calculate_atom_states(
    *,
    site_frac: np.ndarray,
    site_labels,
    positions: np.ndarray,
    lattice_matrix: np.ndarray,
    site_radius: dict,
    site_inner_fraction: dict,
) -> np.ndarray:
    """transitions.py:769-809 on top of the restated PeriodicKDTree; int64 [T, A]."""
    lattice = Lattice(lattice_matrix)
    cutoff = max(list(site_radius.values()))
    positions = np.asarray(positions, dtype=np.float64)
    frac = positions.reshape(-1, 3)
    cart = lattice.get_cartesian_coords(frac)

    tree = PeriodicKDTree(box=np.array(lattice.parameters, dtype=np.float32))
    tree.set_coords(cart, cutoff=cutoff)

    shape = positions.shape[0:2]
    atom_sites = np.full((cart.shape[0]), NOSITE)
    site_frac = np.asarray(site_frac, dtype=np.float64)

    for label, radius, key in label_groups(site_labels, site_radius):
        group_frac = site_frac if key is None else site_frac[key]
        cart_sites = lattice.get_cartesian_coords(group_frac)
        pairs = tree.search_tree(cart_sites, radius * site_inner_fraction[label])
        if pairs.size == 0:
            warnings.warn(f'No floating species in range of {label} ({radius=})', stacklevel=2)
            continue
        siteno, index = pairs.T
        if key is not None:
            siteno = integer_remap(a=siteno, key=key, palette=np.unique(siteno))
        atom_sites[index] = siteno  # sorted by (siteno, index): last write = highest siteno
    return atom_sites.reshape(shape)


# --------------------------------------------------------------------------- brute force


def staged_atoms(positions, lattice_matrix, cutoff):
    """float32 staging of the atoms: wrapped coordinate + (optional) image per axis.

    Returns ``c0`` float32 [N,3] (wrapped) and ``c1`` float32 [N,3] (image coordinate,
    NaN where the atom has no image along that axis).
    """
    lattice = Lattice(lattice_matrix)
    box3 = check_ortho_box(np.array(lattice.parameters, dtype=np.float32))
    cart = lattice.get_cartesian_coords(np.asarray(positions, dtype=np.float64).reshape(-1, 3))
    c0 = ortho_pbc_f32(cart.astype(np.float32), box3)
    r32 = np.float32(cutoff)
    lo = c0 <= r32
    hi = ((box3[None, :] - c0).astype(np.float32) <= r32) & ~lo
    c1 = np.full_like(c0, np.nan)
    plus = (c0 + box3[None, :]).astype(np.float32)
    minus = (c0 - box3[None, :]).astype(np.float32)
    c1[lo] = plus[lo]
    c1[hi] = minus[hi]
    return c0, c1, box3


def staged_sites(site_frac, lattice_matrix):
    lattice = Lattice(lattice_matrix)
    box3 = check_ortho_box(np.array(lattice.parameters, dtype=np.float32))
    cart = lattice.get_cartesian_coords(np.asarray(site_frac, dtype=np.float64).reshape(-1, 3))
    return ortho_pbc_f32(cart, box3)


def min_d2(site32, c0, c1):
    """float64 ``dx*dx+dy*dy+dz*dz`` (x, y, z order) minimised over the atom's images.

    Rounded addition is monotone, so the minimum over image combinations is the
    combination of per-axis minima.
    """
    s = site32.astype(np.float64)
    d0 = s[None, :] - c0.astype(np.float64)
    q = d0 * d0
    d1 = s[None, :] - c1.astype(np.float64)
    q1 = d1 * d1
    q = np.where(np.isnan(q1), q, np.minimum(q, q1))
    return (q[:, 0] + q[:, 1]) + q[:, 2]


def calculate_atom_states_bruteforce(
    *,
    site_frac,
    site_labels,
    positions,
    lattice_matrix,
    site_radius: dict,
    site_inner_fraction: dict,
    return_visited: bool = False,
):
    positions = np.asarray(positions, dtype=np.float64)
    shape = positions.shape[0:2]
    cutoff = max(list(site_radius.values()))
    c0, c1, _ = staged_atoms(positions, lattice_matrix, cutoff)
    s32 = staged_sites(site_frac, lattice_matrix)
    out = np.full(c0.shape[0], NOSITE, dtype=np.int64)
    visited_all = np.zeros(len(s32), dtype=bool)
    for label, radius, key in label_groups(site_labels, site_radius):
        members = np.arange(len(s32)) if key is None else key
        r = radius * site_inner_fraction[label]
        if cutoff < r:
            raise RuntimeError('Set cutoff greater or equal to the radius.')
        r2 = r * r
        hit = np.zeros((len(members), c0.shape[0]), dtype=bool)
        for li, g in enumerate(members):
            hit[li] = min_d2(s32[g], c0, c1) <= r2
        visited = hit.any(axis=1)
        visited_all[members[visited]] = True
        if not visited.any():
            warnings.warn(f'No floating species in range of {label} ({radius=})', stacklevel=2)
            continue
        if key is None:
            value = np.arange(len(members))
        else:
            # integer_remap(siteno, key, palette=unique(siteno)): rank among visited
            rank = np.cumsum(visited) - 1
            value = key[np.clip(rank, 0, None)]
        for li in range(len(members)):  # ascending local index: last write wins
            out[hit[li]] = value[li]
    out = out.reshape(shape)
    if return_visited:
        return out, visited_all
    return out


def boundary_band_counts(
    *, site_frac, site_labels, positions, lattice_matrix, site_radius, site_inner_fraction,
    bands=(1e-12, 4e-6),
):
    """Number of atom-frames with ``| |d| - r | < band`` for at least one site, per band."""
    positions = np.asarray(positions, dtype=np.float64)
    cutoff = max(list(site_radius.values()))
    c0, c1, _ = staged_atoms(positions, lattice_matrix, cutoff)
    s32 = staged_sites(site_frac, lattice_matrix)
    flags = [np.zeros(c0.shape[0], dtype=bool) for _ in bands]
    for label, radius, key in label_groups(site_labels, site_radius):
        members = np.arange(len(s32)) if key is None else key
        r = radius * site_inner_fraction[label]
        r2 = r * r
        for g in members:
            d2 = min_d2(s32[g], c0, c1)
            near = np.abs(d2 - r2) < r2 * 1e-3
            if near.any():
                dd = np.abs(np.sqrt(d2[near]) - np.sqrt(r2))
                idx = np.nonzero(near)[0]
                for f, b in zip(flags, bands):
                    f[idx[dd < b]] = True
    return tuple(int(f.sum()) for f in flags)
