"""ctypes wrapper of oracle/c/states_bf.c -- TEST INFRASTRUCTURE ONLY.

Fast (multi-threaded) brute-force evaluation of the same float32-staged geometry as
``oracle.states``; used where the numpy/scipy oracle would take minutes.  Labelled
groups with never-visited sites (the integer_remap quirk) are handled by a second
pass: the visited flags of the first pass define the remapped values.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

from .states import label_groups, staged_sites

_HERE = Path(__file__).resolve().parent
_SO = _HERE / '_build' / 'liboracle.so'


def _lib():
    if not _SO.exists():
        subprocess.run(['make', '-C', str(_HERE), '--no-print-directory'], check=True)
    lib = C.CDLL(str(_SO))
    lib.oracle_states_bf.restype = None
    lib.oracle_states_bf.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_double, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_int]
    return lib


def calculate_atom_states_pair(*, site_frac, site_labels, positions, lattice_matrix,
                               site_radius: dict, site_inner_fraction: dict, n_threads: int | None = None):
    """(outer, inner) int64 [T, A]; outer uses fraction 1.0 (transitions.py:160-172)."""
    lattice_matrix = np.asarray(lattice_matrix, dtype=np.float64)
    off = lattice_matrix - np.diag(np.diag(lattice_matrix))
    if np.any(off != 0):
        raise NotImplementedError('C oracle handles diagonal lattices only')
    positions = np.ascontiguousarray(positions, dtype=np.float64)
    shape = positions.shape[:2]
    s32 = staged_sites(site_frac, lattice_matrix)
    order, r2o, r2i, keys = [], [], [], []
    for label, radius, key in label_groups(site_labels, site_radius):
        members = np.arange(len(s32)) if key is None else key
        ro = radius * 1.0
        ri = radius * site_inner_fraction[label]
        order.append(members)
        r2o.append(np.full(len(members), ro * ro))
        r2i.append(np.full(len(members), ri * ri))
        keys.append((key, len(members)))
    order = np.concatenate(order)
    r2o, r2i = np.concatenate(r2o), np.concatenate(r2i)
    site32 = np.ascontiguousarray(s32[order], dtype=np.float32)
    diag = np.ascontiguousarray(np.diag(lattice_matrix))
    cutoff = float(max(site_radius.values()))
    n = shape[0] * shape[1]
    lib = _lib()
    n_threads = n_threads or os.cpu_count() or 1

    def run(nat_o, nat_i):
        outs = []
        visited = np.zeros(2 * len(order), dtype=np.uint8)
        for which, nat in ((0, nat_o), (1, nat_i)):
            o = np.empty(n, dtype=np.int64)
            i = np.empty(n, dtype=np.int64)
            nat = np.ascontiguousarray(nat, dtype=np.int64)
            lib.oracle_states_bf(positions.ctypes.data, n, diag.ctypes.data, cutoff,
                                 site32.ctypes.data, nat.ctypes.data, r2o.ctypes.data,
                                 r2i.ctypes.data, len(order), o.ctypes.data, i.ctypes.data,
                                 visited.ctypes.data, n_threads)
            outs.append(o if which == 0 else i)
        return outs[0], outs[1], visited.reshape(2, -1).astype(bool)

    outer, inner, visited = run(order, order)
    # integer_remap quirk: i-th visited site of a labelled group is reported as key[i]
    nat = [order.copy(), order.copy()]
    remap = False
    pos0 = 0
    for key, m in keys:
        if key is not None:
            for w in (0, 1):
                v = visited[w, pos0:pos0 + m]
                if not v.all():
                    rank = np.cumsum(v) - 1
                    nat[w][pos0:pos0 + m][v] = key[rank[v]]
                    remap = True
        pos0 += m
    if remap:
        outer, inner, _ = run(nat[0], nat[1])
    return outer.reshape(shape), inner.reshape(shape)
