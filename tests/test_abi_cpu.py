"""CPU-side checks of the C ABI: every declared symbol is exported, and the host-built
site grid is a conservative superset of the exact matches (so K1's exact test decides)."""

from __future__ import annotations

import ctypes as C
import re
import struct
from pathlib import Path

import numpy as np
import pytest

from gemdat_b200 import _lib
from gemdat_b200.transitions import SiteGrid
from oracle import states as ostates

ROOT = Path(__file__).resolve().parents[1]


def header_symbols():
    text = (ROOT / 'include' / 'gemdat_b200.h').read_text()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(gd_[a-z0-9_]+)\s*\(', text)))


def test_every_declared_symbol_is_exported_and_bound():
    syms = header_symbols()
    assert len(syms) >= 20
    for name in syms:
        assert hasattr(_lib.lib, name), f'{name} declared in include/gemdat_b200.h but not exported'
    assert sorted(_lib.SIGNATURES) == syms


def test_version_and_error_string():
    assert _lib.lib.gd_version() == 100
    assert isinstance(_lib.last_error(), str)


def parse_blob(blob: np.ndarray):
    raw = blob.tobytes()
    magic, version, n_sites, n_total, n_groups = struct.unpack_from('<IIiii', raw, 0)
    dims = struct.unpack_from('<3i', raw, 20)
    n_entries, is_diag, max_count = struct.unpack_from('<3i', raw, 32)
    box = np.frombuffer(raw, dtype=np.float32, count=3, offset=44)
    cutoff32 = np.frombuffer(raw, dtype=np.float32, count=1, offset=56)[0]
    inv_cell = np.frombuffer(raw, dtype=np.float32, count=3, offset=60)
    off = 76 + 4  # pad0 -> 80
    inv_box = np.frombuffer(raw, dtype=np.float64, count=3, offset=off)
    matrix = np.frombuffer(raw, dtype=np.float64, count=9, offset=off + 24)
    o_sites, o_cells, o_entries, o_groups, total, o_excl = struct.unpack_from('<6Q', raw, off + 24 + 72)
    assert magic == 0x47534447 and version == 4 and total == len(raw)
    # 16-byte site records: float32 x, y, z (staged like the reference) + packed int32
    sites = np.frombuffer(raw, dtype=np.float32, count=4 * n_sites, offset=o_sites).reshape(-1, 4)
    natgrp = sites[:, 3].copy().view(np.int32)
    excl2 = np.frombuffer(raw, dtype=np.float32, count=n_sites, offset=o_excl)
    ncell = dims[0] * dims[1] * dims[2]
    cells = np.frombuffer(raw, dtype=np.uint32, count=ncell, offset=o_cells)
    entries = np.frombuffer(raw, dtype=np.uint32, count=max(n_entries, 1), offset=o_entries)
    grp = np.frombuffer(raw, dtype=np.uint8, count=48 * n_groups, offset=o_groups).reshape(-1, 48)
    r2 = grp[:, :16].copy().view(np.float64).reshape(-1, 2)  # r2_outer, r2_inner
    band = grp[:, 16:32].copy().view(np.int32).reshape(-1, 4)  # lo_o, span_o, lo_i, span_i
    return dict(n_sites=n_sites, n_total=n_total, dims=dims, box=box, cutoff32=cutoff32,
                inv_cell=inv_cell, matrix=matrix.reshape(3, 3), xyz=sites[:, :3], excl2=excl2,
                site_img=(natgrp >> 20) & 7,
                natural=natgrp & 0xFFFFF, group=natgrp >> 24, cells=cells, entries=entries,
                r2=r2, band=band, is_diag=is_diag, max_count=max_count, n_entries=n_entries)


def emulate_k1(b, c0, c1):
    """numpy restatement of the device walk: cell -> candidate words -> flagged-axis image ->
    exact test; returns (outer, inner, n_band_prefilter_misses)."""
    dims = b['dims']
    idx = np.empty((len(c0), 3), dtype=np.int64)
    for j in range(3):
        raw = (c0[:, j] * b['inv_cell'][j]).astype(np.float32).astype(np.int64)
        idx[:, j] = np.minimum(raw.astype(np.uint32), dims[j] - 1)
    cell = (idx[:, 0] * dims[1] + idx[:, 1]) * dims[2] + idx[:, 2]
    outer = np.full(len(c0), -1, dtype=np.int64)
    inner = np.full(len(c0), -1, dtype=np.int64)
    prefilter_misses = 0
    for n in range(len(c0)):
        word = int(b['cells'][cell[n]])
        cnt, payload = word >> 24, word & 0xFFFFFF
        words = [payload] if cnt == 1 else [int(w) for w in b['entries'][payload:payload + cnt]]
        for w in words:
            e, flags = w & 0xFFFFF, w >> 20
            s = b['xyz'][e].astype(np.float64)
            d = s - c0[n].astype(np.float64)
            for j in range(3):
                if flags >> j & 1 and not np.isnan(c1[n, j]):
                    alt = s[j] - np.float64(c1[n, j])
                    if abs(alt) < abs(d[j]):
                        d[j] = alt
            d2 = (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]
            r2o, r2i = b['r2'][b['group'][e]]
            if d2 <= r2o:
                outer[n] = b['natural'][e]
            if d2 <= r2i:
                inner[n] = b['natural'][e]
            # the integer pre-filter must pass everything the exact band filter would look at
            hi = int(np.float64(d2).view(np.int64) >> 32)
            lo_o, span_o, lo_i, span_i = (int(v) for v in b['band'][b['group'][e]])
            passed = 0 <= hi - lo_o <= span_o or 0 <= hi - lo_i <= span_i
            exact = abs(d2 - r2o) < r2o * 1e-3 or abs(d2 - r2i) < r2i * 1e-3
            prefilter_misses += exact and not passed
    return outer, inner, prefilter_misses


CASES = [
    # lattice diag, n_sites, radius dict, inner dict, label pattern
    ((19.848, 9.924, 9.924), 96, {'': 0.9}, {'': 0.5}, None),
    ((7.31, 5.07, 6.43), 24, {'Q': 1.1, 'P': 0.8}, {'Q': 0.7, 'P': 1.0}, 'QP'),
    ((38.91, 38.91, 25.94), 400, {'B': 0.5, 'A': 0.6}, {'B': 1.0, 'A': 0.4}, 'AB'),
    ((3.0, 40.0, 2.5), 12, {'': 1.2}, {'': 1.0}, None),  # box narrower than 2*cutoff along x, z
    ((2.0, 2.2, 1.9), 3, {'': 1.0}, {'': 0.3}, None),  # every cell sees every site through images
]


@pytest.mark.parametrize('grid_k', ['0.4', '0.75', '1.5'])
@pytest.mark.parametrize('lengths,n_sites,radius,inner,pattern', CASES)
def test_site_grid_walk_equals_bruteforce(lengths, n_sites, radius, inner, pattern, grid_k, monkeypatch):
    """The host-built grid (exact sphere/box test + per-axis image flags) walked exactly as the
    kernel walks it gives the brute-force oracle's states for atoms anywhere in the box."""
    monkeypatch.setenv('GD_GRID_K', grid_k)
    rng = np.random.default_rng(n_sites)
    lattice = np.diag(lengths)
    site_frac = rng.random((n_sites, 3)) * 1.2 - 0.1  # some outside [0, 1)
    labels = ['X'] * n_sites if pattern is None else [pattern[i % len(pattern)] for i in range(n_sites)]
    grid = SiteGrid(lattice_matrix=lattice, site_frac=site_frac, site_labels=labels,
                    site_radius=radius, site_inner_fraction=inner)
    b = parse_blob(grid.h_blob)
    assert b['n_total'] == n_sites and b['is_diag'] == 1
    # staged site centres equal the oracle's float32 staging, in overwrite (rank) order
    s32 = ostates.staged_sites(site_frac, lattice)
    np.testing.assert_array_equal(b['xyz'], s32[b['natural']])
    np.testing.assert_array_equal(b['box'], np.array(lengths, dtype=np.float32))
    assert b['cutoff32'] == np.float32(max(radius.values()))
    # rank order: groups in dict order, ascending natural index inside a group
    order = []
    for g, label in enumerate(radius):
        order += [i for i in range(n_sites) if (label == '' or labels[i] == label)]
    np.testing.assert_array_equal(b['natural'], order)

    # atoms: uniform + hugging faces / edges / site surfaces (where the float32 staging decides)
    pos = rng.random((4000, 1, 3))
    pos[:500, 0] = rng.choice([0.0, 1e-9, 1 - 1e-9, 0.5, 1 - 2**-24, 2**-25], size=(500, 3))
    near = site_frac[rng.integers(n_sites, size=1500)] + rng.normal(scale=0.03, size=(1500, 3))
    pos[500:2000, 0] = np.mod(near, 1)
    # points on the sphere surfaces, to rounding
    pick = rng.integers(n_sites, size=700)
    u = rng.normal(size=(700, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    rr = np.array([radius[labels[i]] if pattern is not None else radius[''] for i in pick])
    ff = np.array([inner[labels[i]] if pattern is not None else inner[''] for i in pick])
    rr = np.where(rng.random(700) < 0.5, rr, rr * ff)
    pos[2000:2700, 0] = np.mod(site_frac[pick] + (u * rr[:, None]) / np.array(lengths), 1)
    cutoff = max(radius.values())
    c0, c1, box3 = ostates.staged_atoms(pos, lattice, cutoff)

    got_o, got_i, misses = emulate_k1(b, c0, c1)
    assert misses == 0
    kw = dict(site_frac=site_frac, site_labels=labels, positions=pos, lattice_matrix=lattice,
              site_radius=radius)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        # every site visited -> integer_remap is the identity, the natural indices compare
        exp_o, vis_o = ostates.calculate_atom_states_bruteforce(
            **kw, site_inner_fraction={k: 1.0 for k in radius}, return_visited=True)
        exp_i, vis_i = ostates.calculate_atom_states_bruteforce(
            **kw, site_inner_fraction=inner, return_visited=True)
    lut_o = grid.remap_lut(vis_o)
    lut_i = grid.remap_lut(vis_i)
    if lut_o is not None:
        got_o = np.where(got_o >= 0, lut_o[np.clip(got_o, 0, None)], -1)
    if lut_i is not None:
        got_i = np.where(got_i >= 0, lut_i[np.clip(got_i, 0, None)], -1)
    np.testing.assert_array_equal(got_o, exp_o[:, 0])
    np.testing.assert_array_equal(got_i, exp_i[:, 0])
    # candidate lists ascend in rank (last match = overwrite winner)
    multi = np.nonzero((b['cells'] >> 24) > 1)[0][:2000]
    for c in multi:
        cnt, payload = int(b['cells'][c]) >> 24, int(b['cells'][c]) & 0xFFFFFF
        lst = b['entries'][payload:payload + cnt] & 0xFFFFF
        assert np.all(np.diff(lst.astype(np.int64)) > 0)


def test_sitegrid_error_paths():
    frac = np.random.default_rng(0).random((5, 3))
    with pytest.raises(NotImplementedError):
        SiteGrid(lattice_matrix=np.array([[5.0, 0, 0], [1.0, 5.0, 0], [0, 0, 5.0]]), site_frac=frac,
                 site_labels=['a'] * 5, site_radius={'': 0.5}, site_inner_fraction={'': 1.0})
    with pytest.raises(RuntimeError, match='Set cutoff greater or equal to the radius'):
        SiteGrid(lattice_matrix=np.eye(3) * 5, site_frac=frac, site_labels=['a'] * 5,
                 site_radius={'': 0.5}, site_inner_fraction={'': 1.5})
    with pytest.raises(ValueError):
        SiteGrid(lattice_matrix=np.eye(3) * 5, site_frac=frac, site_labels=['a'] * 5,
                 site_radius={'zzz': 0.5}, site_inner_fraction={'zzz': 1.0})


def test_compute_entry_points_fail_loudly_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from gemdat_b200 import Trajectory

    t = Trajectory(species=['Li'], coords=np.zeros((4, 1, 3)), lattice=np.eye(3), time_step=1.0)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        t.distances_from_base_position()


@pytest.mark.parametrize('lengths,n_sites,radius,inner,pattern', CASES + [
    ((25.94, 25.94, 12.97), 160, {'': 0.6}, {'': 0.5}, None),  # well separated sites: big balls
])
def test_exclusive_ball_decides_alone(lengths, n_sites, radius, inner, pattern):
    """K1's register cache: inside the exclusive ball of a site (periodic metric) the site's own
    record, evaluated with its SITE-level image flags, gives the brute-force oracle's states."""
    rng = np.random.default_rng(7 * n_sites + 1)
    lattice = np.diag(lengths)
    if n_sites == 160:  # jittered lattice, >= 2.2 A apart
        g = np.stack(np.meshgrid(np.arange(8), np.arange(5), np.arange(4), indexing='ij'), -1).reshape(-1, 3)
        site_frac = (g + 0.5 + rng.uniform(-0.12, 0.12, g.shape)) / np.array([8, 5, 4])
    else:
        site_frac = rng.random((n_sites, 3)) * 1.2 - 0.1
    labels = ['X'] * n_sites if pattern is None else [pattern[i % len(pattern)] for i in range(n_sites)]
    grid = SiteGrid(lattice_matrix=lattice, site_frac=site_frac, site_labels=labels,
                    site_radius=radius, site_inner_fraction=inner)
    b = parse_blob(grid.h_blob)
    has_ball = np.nonzero(b['excl2'] > 0)[0]
    if n_sites == 160:
        assert len(has_ball) == n_sites and b['excl2'].min() > 0.6**2
    if len(has_ball) == 0:
        return
    # points in the balls: uniform directions, radii up to (just inside) the exclusive radius
    pick = has_ball[rng.integers(len(has_ball), size=6000)]
    u = rng.normal(size=(6000, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    rad = np.sqrt(b['excl2'][pick].astype(np.float64)) * np.minimum(1.0, rng.random(6000) * 1.05) * (1 - 1e-6)
    centre = b['xyz'][pick].astype(np.float64)
    pos = np.mod((centre + u * rad[:, None]) / np.array(lengths), 1.0).reshape(-1, 1, 3)
    cutoff = max(radius.values())
    c0, c1, _ = ostates.staged_atoms(pos, lattice, cutoff)
    # the fast path's arithmetic: per-axis image on the site-level flagged axes
    s64 = b['xyz'][pick].astype(np.float64)
    d = s64 - c0.astype(np.float64)
    alt = s64 - c1.astype(np.float64)
    use = ((b['site_img'][pick][:, None] >> np.arange(3)) & 1).astype(bool) & ~np.isnan(c1) & (np.abs(alt) < np.abs(d))
    d = np.where(use, alt, d)
    d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    inside = d2 <= b['excl2'][pick].astype(np.float64)
    assert inside.mean() > 0.9
    r2o, r2i = b['r2'][b['group'][pick]].T
    got_o = np.where(d2 <= r2o, b['natural'][pick], -1)
    got_i = np.where(d2 <= r2i, b['natural'][pick], -1)
    kw = dict(site_frac=site_frac, site_labels=labels, positions=pos, lattice_matrix=lattice, site_radius=radius)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        exp_o, vis_o = ostates.calculate_atom_states_bruteforce(
            **kw, site_inner_fraction={k: 1.0 for k in radius}, return_visited=True)
        exp_i, vis_i = ostates.calculate_atom_states_bruteforce(
            **kw, site_inner_fraction=inner, return_visited=True)
    lut_o, lut_i = grid.remap_lut(vis_o), grid.remap_lut(vis_i)
    if lut_o is not None:
        got_o = np.where(got_o >= 0, lut_o[np.clip(got_o, 0, None)], -1)
    if lut_i is not None:
        got_i = np.where(got_i >= 0, lut_i[np.clip(got_i, 0, None)], -1)
    np.testing.assert_array_equal(got_o[inside], exp_o[inside, 0])
    np.testing.assert_array_equal(got_i[inside], exp_i[inside, 0])


def test_rdf_state_table_matches_the_reference_names():
    """Host side of N4: gemdat_b200.rdf._state_names (dense code -> name index) against the
    code -> name dict the reference's _get_states produced for the golden run, through the
    reference's own integer codes (rdf.py:30-50)."""
    from conftest import load_golden
    from gemdat_b200 import rdf

    g = load_golden('rdf')
    uniq = [str(x) for x in g['unique_labels']]
    lut, names = rdf._state_names(uniq)
    ref = dict(zip(g['states2str_codes'].tolist(), [str(x) for x in g['states2str_names']]))
    n = len(uniq)
    assert len(lut) == (n + 1) ** 3 and len(ref) == len(lut)
    for i in range(-1, n):
        for j in range(-1, n):
            for k in range(-1, n):
                dense = ((i + 1) * (n + 1) + (j + 1)) * (n + 1) + (k + 1)
                assert names[lut[dense]] == ref[int(i * 1e6 + j * 1e3 + k)]
    # dense codes order states like the reference's signed codes do (key order of the result)
    codes = [int(i * 1e6 + j * 1e3 + k) for i in range(-1, n) for j in range(-1, n) for k in range(-1, n)]
    assert codes == sorted(codes)


def test_collective_and_rdf_entry_points_reject_bad_arguments_without_a_gpu():
    """Argument checks run before any CUDA call, so they are testable here."""
    lib = _lib.lib
    assert lib.gd_collective_work_bytes(-1) == -1
    n_pairs, n_solo = C.c_int64(7), C.c_int64(7)
    assert lib.gd_collective_count(None, 0, None, 0, 5, None, C.byref(n_pairs), C.byref(n_solo), None) == 0
    assert (n_pairs.value, n_solo.value) == (0, 0)
    assert lib.gd_collective_count(None, 3, None, 1, 5, None, C.byref(n_pairs), C.byref(n_solo), None) != 0
    assert 'null pointer' in _lib.last_error()
    assert lib.gd_collective_sort(None, -1, None, None, None) != 0
    lat = np.eye(3) * 5.0
    assert lib.gd_rdf_hist(None, 4, 3, lat.ctypes.data, None, 2, None, 1, None, None, 2, 1, None, 1, 0.1, 0, 0,
                           None, None) != 0  # fewer than two edges
    assert 'gd_rdf_hist' in _lib.last_error()
    assert lib.gd_rdf_hist(None, 0, 3, lat.ctypes.data, None, 2, None, 1, None, None, 2, 1, None, 5, 0.1, 0, 0,
                           None, None) == 0  # no frames: nothing to do
    assert lib.gd_rdf_classes(None, None, None, 0, 4, None, 3, 2, None, None, None, None) == 0
    assert lib.gd_rdf_classes(None, None, None, 2, 4, None, 3, 1000, None, None, None, None) != 0


def test_sqrt_threshold_argument_of_the_rdf_binning():
    """gd_rdf.cu bins on the squared distance: for every edge e it walks from fl(e*e) to the
    smallest double x with fl(sqrt(x)) > e (or >= e).  The same walk in numpy (np.sqrt is correctly
    rounded like the device's) must make `x <= d2` agree with the comparison on the rounded root
    for every double around the threshold."""
    def threshold(e, strict):
        if not e > 0.0:
            return np.nextafter(0.0, 1.0) if strict else 0.0
        ok = (lambda v: np.sqrt(v) > e) if strict else (lambda v: np.sqrt(v) >= e)
        x = np.float64(e) * np.float64(e)
        while x > 0.0 and ok(x):
            x = np.nextafter(x, -np.inf)
        while not ok(x):
            x = np.nextafter(x, np.inf)
        return x

    for edges in (np.arange(0, 5.0 + 0.1, 0.1), np.arange(0, 3.3 + 0.25, 0.25), np.array([0.0, 1e-9, 2.0, 1e4])):
        for strict in (True, False):
            for e in edges:
                x = threshold(float(e), strict)
                around = [x]
                for _ in range(6):
                    around = [np.nextafter(around[0], -np.inf)] + around + [np.nextafter(around[-1], np.inf)]
                for d2 in around:
                    if d2 < 0:
                        continue
                    d = np.sqrt(d2)
                    assert (d > e if strict else d >= e) == (x <= d2), (e, strict, d2)


def test_drift_selection_and_atomic_masses_host_logic():
    """Host side of Trajectory.drift / center_of_mass (no device needed): which atoms the drift is
    averaged over (trajectory.py:778-789; `floating_species` follows the docstring, see
    tests/golden/make_golden_drift.py) and the mass look-up."""
    from gemdat_b200 import Trajectory
    from gemdat_b200.elements import atomic_mass

    species = ['Li', 'Li+', 'P', 'S', 'S2-', 'Br']
    traj = Trajectory(species=species, coords=np.zeros((2, 6, 3)), lattice=np.eye(3) * 10, time_step=1e-15)
    assert traj._selected_atoms(None, None).tolist() == [0, 1, 2, 3, 4, 5]
    assert traj._selected_atoms('S', None).tolist() == [3, 4]
    assert traj._selected_atoms(('P', 'Br'), None).tolist() == [2, 5]
    assert traj._selected_atoms(None, 'Li').tolist() == [2, 3, 4, 5]
    assert traj._selected_atoms(None, ('Li', 'S')).tolist() == [2, 5]
    assert traj._selected_atoms('P', 'Li').tolist() == [2]  # fixed_species wins, like the reference
    assert traj._selected_atoms('Xe', None).size == 0
    assert atomic_mass('Li') == atomic_mass('Li+') == 6.941
    assert atomic_mass('S2-') == 32.065

    class WithMass:
        atomic_mass = 12.5

    assert atomic_mass(WithMass()) == 12.5
    with pytest.raises(KeyError):
        atomic_mass('Qq')


def test_trajectory_cache_round_trip(tmp_path):
    """Trajectory.to_cache / from_cache (trajectory.py:195-217): a pickle of the host data; the
    cached device copy does not travel."""
    from gemdat_b200 import Trajectory

    rng = np.random.default_rng(3)
    traj = Trajectory(species=['Li', 'Li', 'S'], coords=rng.random((7, 3, 3)), lattice=np.diag([9.0, 10.0, 11.0]),
                      metadata={'temperature': 600.0}, time_step=2e-15)
    traj.__dict__['_dev_holder'] = {'pos': ('would be a device tensor',)}
    path = tmp_path / 'traj.cache'
    traj.to_cache(path)
    back = Trajectory.from_cache(path)
    assert '_dev_holder' not in back.__dict__
    assert back.species == traj.species and back.metadata == traj.metadata and back.time_step == traj.time_step
    np.testing.assert_array_equal(back.coords, traj.coords)
    np.testing.assert_array_equal(back.lattice, traj.lattice)
    np.testing.assert_array_equal(back.base_positions, traj.base_positions)
    assert back.constant_lattice and not back.coords_are_displacement and len(back) == 7
    lat = np.stack([np.eye(3) * (10 + t) for t in range(7)])
    per_frame = Trajectory(species=['Li', 'Li', 'S'], coords=rng.random((7, 3, 3)), lattice=lat,
                           constant_lattice=False, time_step=2e-15)
    per_frame.to_cache(path)
    assert Trajectory.from_cache(path).get_lattice(4).matrix[0, 0] == 14.0


def test_trajectory_with_one_lattice_per_frame_host_logic():
    """Trajectory(constant_lattice=False): the lattice array follows frame slices and filters,
    get_lattice(idx) picks a frame (trajectory.py:686-703), and whatever needs THE lattice says so
    before touching the device."""
    from gemdat_b200 import Trajectory

    rng = np.random.default_rng(5)
    lat = np.stack([np.diag([9.0, 10.0, 11.0]) * (1 + 0.01 * t) for t in range(12)])
    traj = Trajectory(species=['Li', 'S', 'Li'], coords=rng.random((12, 3, 3)), lattice=lat, constant_lattice=False,
                      time_step=2e-15)
    assert traj.lattice.shape == (12, 3, 3) and not traj.constant_lattice
    np.testing.assert_array_equal(traj.get_lattice(5).matrix, lat[5])
    part = traj[3:9]
    assert len(part) == 6 and not part.constant_lattice
    np.testing.assert_array_equal(part.lattice, lat[3:9])
    li = traj.filter('Li')
    assert li.species == ['Li', 'Li'] and li.lattice.shape == (12, 3, 3) and not li.constant_lattice
    assert [len(p) for p in traj.split(3)] == [3, 4, 4] and all(not p.constant_lattice for p in traj.split(3))
    for call in (traj.get_lattice, traj.distances_from_base_position, lambda: traj.metrics().tracer_diffusivity()):
        with pytest.raises(NotImplementedError):
            call()
    with pytest.raises(ValueError):
        Trajectory(species=['Li'], coords=rng.random((4, 1, 3)), lattice=lat[:3], constant_lattice=False)
    assert 'constant_lattice=False' in repr(traj)


def test_drift_oracle_edge_cases():
    """oracle/drift.py on the edges the kernels are tested against: one frame, one atom, steps that
    cross the cell boundary, an empty selection (numpy's NaN with a warning, like the reference)."""
    import warnings

    from oracle import drift as odrift

    one = np.array([[[0.2, 0.4, 0.9]]])
    assert odrift.drift(one, [0]).shape == (1, 1, 3) and not odrift.drift(one, [0]).any()
    np.testing.assert_array_equal(odrift.corrected_positions(one, [0]), one)
    pos = np.array([[[0.95, 0.5, 0.05]], [[0.05, 0.5, 0.95]], [[0.15, 0.5, 0.85]]])  # wraps in x and z
    np.testing.assert_allclose(odrift.displacements(pos)[:, 0], [[0, 0, 0], [0.1, 0, -0.1], [0.1, 0, -0.1]], atol=1e-15)
    # drift of the only atom is its own step: the corrected atom stays where it started
    np.testing.assert_allclose(odrift.corrected_positions(pos, [0]), np.broadcast_to(pos[0], pos.shape), atol=1e-15)
    np.testing.assert_allclose(odrift.center_of_mass(pos, [2.0])[:, 0], [[0.95, 0.5, 0.05], [1.05, 0.5, -0.05],
                                                                         [1.15, 0.5, -0.15]], atol=1e-15)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        assert np.isnan(odrift.drift(pos, [])).all()


def test_reference_arm_contract_without_gpu_and_without_the_library():
    """`bench.py --impl reference` (the CPU arm the driver runs beside ours): prints ONE JSON line with
    the contract's keys, needs no GPU, and never maps libgemdat_b200.so (it times the oracle port only)."""
    import json
    import subprocess
    import sys

    root = Path(__file__).resolve().parents[1]
    code = (
        "import sys, runpy\n"
        "sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--cpu-frames', '60']\n"
        "try:\n"
        "    runpy.run_path('bench.py', run_name='__main__')\n"
        "except SystemExit as e:\n"
        "    assert not e.code, e.code\n"
        "print('MAPPED', 'libgemdat_b200' in open('/proc/self/maps').read())\n"
    )
    out = subprocess.run([sys.executable, '-c', code], cwd=root, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['n_gpus'] == 1 and d['steps'] == 1 and d['warmup'] == 0
    assert d['metric'] == 'site_assignment_plus_jumps_atom_frames_per_s' and d['unit'] == 'atom-frames/s'
    assert d['higher_is_better'] is True and d['value'] > 0 and d['ms_per_step'] > 0
    assert d['config']['workload'].startswith('C3 ') and d['dtype'] == 'f64' and d['data'] == 'synthetic'
    assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1
    assert d['cpu_baseline']['value'] == d['value']
    assert d['e2e'] == {'value': d['value'], 'unit': d['unit'], 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert 'MAPPED False' in out.stdout
