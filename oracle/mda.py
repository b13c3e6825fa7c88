"""Restatement of the MDAnalysis 2.6.1 pieces behind ``PeriodicKDTree``  [recalled].

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference calls (src/gemdat/transitions.py:776-779, 796)::

    tree = PeriodicKDTree(box=np.array(lattice.parameters, dtype=np.float32))
    tree.set_coords(cart_coords, cutoff=max_radius)
    pairs = tree.search_tree(cart_sites, radius * fraction)

MDAnalysis is not installed here and its source is not under /root/reference,
so its published algorithm is restated (orthorhombic boxes only -- SURVEY H2):

* ``set_coords``: cast to float32, wrap into the box in float32 (``_ortho_pbc`` of
  lib/src/calc_distances.h), create periodic images of every atom closer than
  ``cutoff`` to a face (lib/_augment.pyx, float32 arithmetic), build a
  ``scipy.spatial.cKDTree`` over originals + images.
* ``search_tree``: cast the query centres to float32, wrap them, build a second
  cKDTree, ``query_ball_tree`` (REAL scipy code, float64 compare of
  ``dx*dx+dy*dy+dz*dz <= r*r``), map image hits back to their atom
  (``undo_augment``) and return the lexicographically sorted unique
  ``(centre, atom)`` rows (lib.util.unique_rows).
"""

from __future__ import annotations

import numpy as np
from scipy.spatial import cKDTree


def check_ortho_box(box) -> np.ndarray:
    """float32 ``[a, b, c]`` if all three angles equal 90 in float32, else raise."""
    box = np.asarray(box, dtype=np.float32)
    if box.shape != (6,):
        raise ValueError('box must be [a, b, c, alpha, beta, gamma]')
    if not np.all(box[3:] == np.float32(90.0)):
        raise NotImplementedError(
            'triclinic boxes are out of scope for the oracle (SURVEY.md H2)'
        )
    return box[:3].copy()


def ortho_pbc_f32(coords, box3) -> np.ndarray:
    """``apply_PBC`` for an orthorhombic box: float32 in, float32 out.

    Single-box shift first, multi-box shift only when one shift is not enough,
    lower bound handled before the upper bound (``_ortho_pbc``).
    """
    c = np.array(coords, dtype=np.float32, copy=True).reshape(-1, 3)
    box3 = np.asarray(box3, dtype=np.float32)
    if not box3.any():
        return c
    inv = 1.0 / box3.astype(np.float64)
    for j in range(3):
        L = box3[j]
        col = c[:, j]
        # --- lower bound
        neg = col < np.float32(0.0)
        if neg.any():
            once = (col + L).astype(np.float32)  # float32 add (== rounded double add)
            ok = neg & ~(once < np.float32(0.0))
            multi = neg & (once < np.float32(0.0))
            col[ok] = once[ok]
            if multi.any():
                s = np.floor(col[multi].astype(np.float64) * inv[j]).astype(np.int32)
                shift = (s.astype(np.float32) * L).astype(np.float32)
                v = (col[multi] - shift).astype(np.float32)
                again = v < np.float32(0.0)
                v[again] = (v[again] + L).astype(np.float32)
                col[multi] = v
        # --- upper bound (always checked after the lower bound)
        hi = col >= L
        if hi.any():
            once = (col - L).astype(np.float32)
            ok = hi & ~(once >= L)
            multi = hi & (once >= L)
            col[ok] = once[ok]
            if multi.any():
                s = np.floor(col[multi].astype(np.float64) * inv[j]).astype(np.int32)
                shift = (s.astype(np.float32) * L).astype(np.float32)
                v = (col[multi] - shift).astype(np.float32)
                again = v >= L
                v[again] = (v[again] - L).astype(np.float32)
                col[multi] = v
        c[:, j] = col
    return c


def augment_ortho_f32(coords32, box3, r) -> tuple[np.ndarray, np.ndarray]:
    """Periodic images of atoms within ``r`` of a face (``augment_coordinates``).

    All arithmetic in float32 (``r`` is a C float in the Cython source).  Per axis an
    atom gets the ``+L`` image if ``x <= r``; *else* the ``-L`` image if ``L - x <= r``;
    every combination over the three axes is emitted (faces, edges, corners).

    Returns (images float32 [M,3], owner index int64 [M]).
    """
    c = np.asarray(coords32, dtype=np.float32)
    L = np.asarray(box3, dtype=np.float32)
    r32 = np.float32(r)
    lo = c <= r32  # dot(coord, unit reciprocal) == coord for an orthorhombic box
    hi = (L[None, :] - c).astype(np.float32) <= r32
    hi &= ~lo  # "elif hi_x"
    plus = (c + L[None, :]).astype(np.float32)
    minus = (c - L[None, :]).astype(np.float32)
    shifted = np.where(lo, plus, minus)  # only meaningful where lo|hi
    has = lo | hi

    images = []
    owners = []
    idx_all = np.arange(len(c))
    # all non-empty subsets of axes
    for mask in range(1, 8):
        axes = [j for j in range(3) if mask >> j & 1]
        sel = np.ones(len(c), dtype=bool)
        for j in axes:
            sel &= has[:, j]
        if not sel.any():
            continue
        img = c[sel].copy()
        for j in axes:
            img[:, j] = shifted[sel, j]
        images.append(img)
        owners.append(idx_all[sel])
    if images:
        return np.concatenate(images).astype(np.float32), np.concatenate(owners)
    return np.zeros((0, 3), dtype=np.float32), np.zeros((0,), dtype=np.int64)


def unique_rows(pairs: np.ndarray) -> np.ndarray:
    """Sorted unique rows (lexicographic), as ``MDAnalysis.lib.util.unique_rows``."""
    if pairs.size == 0:
        return pairs
    return np.unique(pairs, axis=0)


class PeriodicKDTree:
    """Drop-in for the subset of ``MDAnalysis.lib.pkdtree.PeriodicKDTree`` the reference uses."""

    def __init__(self, box=None, leafsize: int = 10):
        self.leafsize = leafsize
        self.dim = 3
        self.box = None if box is None else np.asarray(box, dtype=np.float32)
        self._built = False
        self.cutoff = None

    @property
    def pbc(self) -> bool:
        return self.box is not None

    def set_coords(self, coords, cutoff=None):
        if not self.pbc:
            raise NotImplementedError('oracle restates the periodic path only')
        if cutoff is None:
            raise RuntimeError('Provide a cutoff distance with tree.set_coords(...)')
        coords = np.asarray(coords, dtype=np.float32)
        self.cutoff = cutoff
        self._box3 = check_ortho_box(self.box)
        self.coords = ortho_pbc_f32(coords, self._box3)
        self.aug, self.mapping = augment_ortho_f32(self.coords, self._box3, cutoff)
        self.all_coords = np.concatenate([self.coords, self.aug])
        self.ckdt = cKDTree(self.all_coords, leafsize=self.leafsize)
        self._built = True

    def search_tree(self, centers, radius) -> np.ndarray:
        if not self._built:
            raise RuntimeError('Unbuilt tree. Run tree.set_coords(...)')
        centers = np.asarray(centers)
        if centers.shape == (self.dim,):
            centers = centers.reshape((1, self.dim))
        if self.cutoff < radius:
            raise RuntimeError('Set cutoff greater or equal to the radius.')
        wrapped = ortho_pbc_f32(centers, self._box3)
        other = cKDTree(wrapped, leafsize=self.leafsize)
        hits = other.query_ball_tree(self.ckdt, radius)
        n = np.fromiter((len(h) for h in hits), dtype=np.intp, count=len(hits))
        if n.sum() == 0:
            return np.zeros((0, 2), dtype=np.intp)
        first = np.repeat(np.arange(len(hits), dtype=np.intp), n)
        second = np.fromiter((j for h in hits for j in h), dtype=np.intp, count=int(n.sum()))
        # undo_augment: image index -> owning atom
        n_real = len(self.coords)
        is_img = second >= n_real
        second[is_img] = self.mapping[second[is_img] - n_real]
        pairs = np.stack([first, second], axis=1)
        return unique_rows(pairs)
