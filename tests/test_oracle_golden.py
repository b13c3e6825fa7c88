"""The oracle against fixtures produced by executing the reference (tests/golden/make_golden.py)
and against the reference's own unit-test constants.  CPU only."""

from __future__ import annotations

import warnings

import numpy as np
import pytest
from numpy.testing import assert_allclose, assert_array_equal

from oracle import displ, events, fill, jumps, msd, radius, states, volume
from oracle.pmg import Lattice


def _states(case, fn, fraction):
    pos = case['positions'][:, case['floating_mask']]
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return fn(
            site_frac=case['site_frac'], site_labels=case['labels'], positions=pos,
            lattice_matrix=case['lattice'], site_radius=case['site_radius'],
            site_inner_fraction=fraction,
        )


def test_states_kdtree_oracle_matches_reference(pipeline_case):
    c = pipeline_case
    outer = {k: 1.0 for k in c['site_radius']}
    assert_array_equal(_states(c, states.calculate_atom_states, outer), c['states'])
    assert_array_equal(_states(c, states.calculate_atom_states, c['inner_fraction']), c['inner_states'])


def test_states_bruteforce_oracle_matches_reference(pipeline_case):
    c = pipeline_case
    outer = {k: 1.0 for k in c['site_radius']}
    assert_array_equal(_states(c, states.calculate_atom_states_bruteforce, outer), c['states'])
    assert_array_equal(
        _states(c, states.calculate_atom_states_bruteforce, c['inner_fraction']), c['inner_states']
    )


def test_events_match_reference(pipeline_case):
    c = pipeline_case
    ev = events.calculate_transition_events(c['states'], c['inner_states'])
    assert_array_equal(ev, c['events'])
    assert list(c['event_columns']) == events.EVENT_COLUMNS


def test_fill_matches_reference(pipeline_case):
    c = pipeline_case
    assert_array_equal(fill.ffill(c['states'], axis=0), c['states_prev'])
    assert_array_equal(fill.bfill(c['states'], axis=0), c['states_next'])


def test_jumps_match_reference(pipeline_case):
    c = pipeline_case
    n_frames = c['positions'].shape[0]
    for key in [k for k in c if k.startswith('jumps_') and k[6:].isdigit()]:
        mr = int(key[6:])
        j = jumps.transitions_to_jumps(c['events'], n_frames, mr)
        assert_array_equal(j, c[key])
        n_sites = len(c['site_frac'])
        m = jumps.transitions_matrix(j[:, 1:3], n_sites)
        assert_array_equal(m, c[f'{key}_matrix'])
        pdist = Lattice(c['lattice']).get_all_distances(c['site_frac'], c['site_frac'])
        d = jumps.jump_diffusivity(
            pdist=pdist, matrix=m, dimensions=3, n_floating=int(c['floating_mask'].sum()),
            total_time=n_frames * 2e-15,
        )
        assert_allclose(d, float(c[f'{key}_diffusivity']), rtol=1e-12)
    assert list(c['jump_columns']) == jumps.JUMP_COLUMNS


def test_event_matrix_matches_reference(pipeline_case):
    c = pipeline_case
    assert_array_equal(jumps.transitions_matrix(c['events'][:, 1:3], len(c['site_frac'])), c['matrix'])


def test_default_radius_matches_reference():
    from conftest import load_golden

    c = load_golden('pipeline_c2_default_radius')
    vib = displ.vibration_amplitude(c['positions'], c['lattice'])
    r = radius.radius_from_vibration_amplitude(c['site_frac'], c['lattice'], vib)
    assert_allclose(r, float(c['radius_values'][0]), rtol=1e-13)


@pytest.mark.parametrize('tag', ['unit', 'walk', 'c2'])
def test_displacement_path_matches_reference(displacement_golden, tag):
    g = displacement_golden
    pos, lat, ts = g[f'{tag}_positions'], g[f'{tag}_lattice'], float(g[f'{tag}_time_step'])
    assert_array_equal(displ.cumulative_displacements(pos), g[f'{tag}_cumulative'])
    assert_allclose(displ.distances_from_base_position(pos, lat), g[f'{tag}_distances'], rtol=1e-14, atol=1e-15)
    assert_allclose(displ.speed(pos, lat), g[f'{tag}_speed'], rtol=1e-12, atol=1e-14)
    assert_allclose(displ.amplitudes(g[f'{tag}_speed']), g[f'{tag}_amplitudes'], rtol=1e-13, atol=1e-15)
    assert_allclose(displ.vibration_amplitude(pos, lat), float(g[f'{tag}_vibration_amplitude']), rtol=1e-12, atol=1e-15)
    assert_allclose(
        displ.tracer_diffusivity(pos, lat, time_step=ts), float(g[f'{tag}_tracer_diffusivity']), rtol=1e-12
    )


@pytest.mark.parametrize('tag', ['unit', 'walk', 'c2'])
def test_msd_matches_reference(displacement_golden, tag):
    g = displacement_golden
    pos, lat = g[f'{tag}_positions'], g[f'{tag}_lattice']
    ref = g[f'{tag}_msd']
    assert_allclose(msd.msd_fft(pos, lat), ref, rtol=1e-12, atol=1e-12)
    # the direct definition agrees with the reference's FFT form to the FFT form's own accuracy
    direct = msd.msd_direct(pos, lat)
    assert_allclose(direct, ref, rtol=1e-9, atol=1e-11 * max(1.0, float(np.abs(ref).max())))


@pytest.mark.parametrize('tag,res', [('unit', 0.2), ('walk', 0.2), ('walk', 0.5), ('c2', 0.2), ('c2', 0.5)])
def test_volume_matches_reference(displacement_golden, tag, res):
    g = displacement_golden
    v = volume.trajectory_to_volume(g[f'{tag}_positions'], g[f'{tag}_lattice'], res)
    assert_array_equal(v, g[f'{tag}_volume_{res}'])
    assert v.sum() == g[f'{tag}_positions'].shape[0] * g[f'{tag}_positions'].shape[1]


def test_reference_unit_test_constants(displacement_golden):
    """Constants asserted by the reference's own tests (tests/trajectory_test.py:65-163,
    tests/metrics_test.py:35,41)."""
    pos, lat = displacement_golden['unit_positions'], displacement_golden['unit_lattice']
    assert_allclose(displ.displacements(pos[:, [0, 3]])[:, 0, 0], [0.0, 0.2, 0.2, 0.2, 0.3])
    assert_allclose(displ.distances_from_base_position(pos[:, :2], lat),
                    [[0.0, 0.2, 0.4, 0.6, 0.9], [0.0, 0.0, 0.0, 0.0, 0.0]])
    m = msd.msd_fft(pos, lat)
    assert_allclose(m[0], [0.0, 0.0525, 0.19, 0.425, 0.81], atol=1e-15)
    assert_allclose(m.mean(), 0.073875)
    assert_allclose(displ.speed(pos[:, :1], lat), [[0.0, 0.2, 0.2, 0.2, 0.3]])
    assert_allclose(displ.amplitudes(displ.speed(pos[:, :1], lat)), [0.9])


def test_utils_match_reference(utils_golden):
    """tests/utils_test.py:17-85 and tests/sites_test.py:9-26 of the reference."""
    u = utils_golden
    assert_array_equal(fill.ffill(u['arr']), [[5, 5, 5, 7, 2], [3, 3, 1, 8, 8], [4, 9, 6, 6, 6]])
    assert_array_equal(fill.bfill(u['arr']), [[5, 7, 7, 7, 2], [3, 1, 1, 8, -1], [4, 9, 6, -1, -1]])
    assert_array_equal(fill.ffill(u['arr'], axis=0), [[5, -1, -1, 7, 2], [3, -1, 1, 8, 2], [4, 9, 6, 8, 2]])
    assert_array_equal(fill.bfill(u['arr'], axis=0), [[5, 9, 1, 7, 2], [3, 9, 1, 8, -1], [4, 9, 6, -1, -1]])
    for key, fn, kw in (('rnd_ffill0', fill.ffill, {'axis': 0}), ('rnd_bfill0', fill.bfill, {'axis': 0}),
                        ('rnd_ffill1', fill.ffill, {}), ('rnd_bfill1', fill.bfill, {})):
        assert_array_equal(fn(u['rnd'], **kw), u[key])
    assert_array_equal(fill.integer_remap(np.array([4, 2, 1, 3]), key=np.array([10, 20, 30, 40])), u['remap'])
    assert_array_equal(jumps.transitions_matrix(u['matrix_events'][:, 1:3], 3), [[0, 1, 0], [2, 0, 0], [1, 1, 0]])
    assert_array_equal(u['matrix'], [[0, 1, 0], [2, 0, 0], [1, 1, 0]])


def test_event_quirks():
    """transitions.py:442-443, 451, 470."""
    const = np.zeros((6, 2), dtype=np.int64)
    with pytest.raises(ValueError):
        events.calculate_transition_events(const, const)
    outer = const.copy()
    outer[3:, 0] = 1
    with pytest.raises(IndexError):
        events.calculate_transition_events(outer, const)
    inner = const.copy()
    inner[2, 1] = 5  # inner-only change on an atom whose outer state is constant: skipped
    inner[4:, 0] = 1
    ev = events.calculate_transition_events(outer, inner)
    assert_array_equal(ev, [[0, 0, 1, 0, 0, 2], [0, 1, 1, 0, 1, 3]])


@pytest.mark.parametrize('tag', ['unit', 'walk', 'c2'])
def test_oracle_attempt_frequency_vs_reference_golden(tag):
    """the numpy restatement of scipy.signal.periodogram + meanfreq against values
    from executing the reference's metrics.py / utils.py (tests/golden/make_golden_attempt.py)."""
    from conftest import load_golden
    from oracle import spectrum

    g, a = load_golden('displacement_metrics'), load_golden('attempt_frequency')
    fs = 1.0 / float(g[f'{tag}_time_step'])
    np.testing.assert_allclose(spectrum.meanfreq(g[f'{tag}_speed'], fs), a[f'{tag}_meanfreq'],
                               rtol=1e-12, equal_nan=True)
    freq, std = spectrum.attempt_frequency(g[f'{tag}_speed'], fs)
    np.testing.assert_allclose(freq, float(a[f'{tag}_attempt_frequency']), rtol=1e-12, equal_nan=True)
    np.testing.assert_allclose(std, float(a[f'{tag}_attempt_frequency_std']), rtol=1e-10, equal_nan=True)


@pytest.mark.parametrize('case', [0, 1, 2, 3])
def test_oracle_collective_vs_reference_golden(case):
    """oracle/collective.py against pairs from executing the reference's
    collective.py:61-106 (tests/golden/make_golden_collective.py), including the sort order and
    the early ``break`` of the inner loop."""
    from conftest import load_golden
    from oracle import collective as oc

    g = load_golden('collective')
    t = f'case{case}'
    src = load_golden(str(g[f'{t}_fixture']))
    jumps = src[str(g[f'{t}_jumps_key'])]
    ev, pairs = oc.collective_pairs(jumps, src['site_frac'], src['lattice'], int(g[f'{t}_max_steps']),
                                    float(g[f'{t}_max_dist']))
    np.testing.assert_array_equal(ev, g[f'{t}_sorted_jumps'])
    np.testing.assert_array_equal(pairs, g[f'{t}_pairs'])
    np.testing.assert_array_equal(oc.coll_jumps(ev, pairs), g[f'{t}_coll_jumps'])
    assert oc.solo_count(len(ev), pairs) == int(g[f'{t}_n_solo_jumps'])
    assert len(ev) - oc.solo_count(len(ev), pairs) == int(g[f'{t}_n_coll_jumps'])
    # derived tables (collective.py:108-166)
    cj = oc.coll_jumps(ev, pairs)
    labels = [str(x) for x in src['site_labels']]
    site_pairs = [tuple(p) for p in g[f'{t}_site_pair_labels'].tolist()]
    np.testing.assert_array_equal(oc.site_pair_count_matrix(cj, labels, site_pairs),
                                  g[f'{t}_site_pair_count_matrix'])
    mj, mc = oc.multiple_collective(cj)
    np.testing.assert_array_equal(np.stack([mj['start'], mj['stop']], axis=-1), g[f'{t}_multiple_jumps'])
    np.testing.assert_array_equal(mc, g[f'{t}_multiple_counts'])


def test_oracle_rdf_vs_reference_golden():
    """oracle/rdf.py against rdf.py executed by tests/golden/make_golden_rdf.py --
    the label off-by-one of _uniqify_labels, the state names, the per-state histograms in the
    reference's key order, and the species-species RDF."""
    from conftest import load_golden
    from oracle import rdf as orf

    g = load_golden('rdf')
    src = load_golden(str(g['fixture']))
    labels = [str(x) for x in src['site_labels']]
    uniq = [str(x) for x in g['unique_labels']]
    species = [str(s) for s in src['species']]
    np.testing.assert_array_equal(orf.uniqify(src['states'], labels, uniq), g['uniqify_states'])
    names = orf.state_names(uniq)
    assert list(names.keys()) == g['states2str_codes'].tolist()
    assert list(names.values()) == [str(x) for x in g['states2str_names']]
    codes = orf.states_array(src['states'], src['states_prev'], src['states_next'], labels, uniq)
    np.testing.assert_array_equal(codes, g['states_array'])
    for n in (0, 1):
        t = f'rdf{n}'
        res = orf.radial_distribution(positions=src['positions'], symbols=species, floating_specie='Li',
                                      codes=codes, unique_labels=uniq, lattice=src['lattice'],
                                      max_dist=float(g[f'{t}_max_dist']), resolution=float(g[f'{t}_resolution']))
        # the reference groups its curves by state (dict of lists): same grouping here
        grouped = {}
        for (state, sym), (x, y) in res.items():
            grouped.setdefault(state, []).append((sym, x, y))
        flat = [(state, sym, x, y) for state, curves in grouped.items() for sym, x, y in curves]
        assert [f[0] for f in flat] == [str(s) for s in g[f'{t}_states']]
        assert [f[1] for f in flat] == [str(s) for s in g[f'{t}_symbols']]
        np.testing.assert_array_equal(flat[0][2], g[f'{t}_x'])
        np.testing.assert_array_equal(np.stack([f[3] for f in flat]), g[f'{t}_y'])
    pos = src['positions']
    sel = lambda want: pos[:, [s in want for s in species]]  # noqa: E731
    for n in (0, 1, 2):
        t = f'between{n}'
        s1 = g[f'{t}_specie_1'].tolist()
        s2 = g[f'{t}_specie_2'].tolist()
        w1 = [s1] if isinstance(s1, str) else s1
        w2 = [s2] if isinstance(s2, str) else s2
        x, y = orf.between_species(coords_1=sel(w1), coords_2=sel(w2), same=(s1 == s2), lattice=src['lattice'],
                                   max_dist=float(g[f'{t}_max_dist']), resolution=float(g[f'{t}_resolution']))
        np.testing.assert_array_equal(x, g[f'{t}_x'])
        np.testing.assert_allclose(y, g[f'{t}_y'], rtol=1e-13)


def test_oracle_collective_edge_cases():
    """Empty and one-row tables, a window of zero frames, and the early break: a partner whose
    start lies beyond the window ends the walk of i even if a later row would qualify."""
    from oracle import collective as oc

    lat = np.eye(3) * 10.0
    frac = np.array([[0.1, 0.1, 0.1], [0.15, 0.1, 0.1], [0.6, 0.6, 0.6]])
    empty = np.zeros((0, 6), dtype=np.int64)
    ev, pairs = oc.collective_pairs(empty, frac, lat, 5, 1.0)
    assert ev.shape == (0, 6) and pairs.shape == (0, 2) and oc.solo_count(0, pairs) == 0
    one = np.array([[0, 0, 1, 3, 4, 0]])
    ev, pairs = oc.collective_pairs(one, frac, lat, 5, 1.0)
    assert pairs.shape == (0, 2) and oc.solo_count(1, pairs) == 1
    # rows sorted by (stop, start): B starts late and stops at 30, C starts early and stops at 31
    #   A: atom 0, start 0, stop 1;  B: atom 1, start 29, stop 30;  C: atom 2, start 2, stop 31
    table = np.array([[0, 0, 1, 0, 1, 0], [1, 2, 2, 29, 30, 0], [2, 1, 0, 2, 31, 0]])
    ev, pairs = oc.collective_pairs(table, frac, lat, 5, 1.0)
    # A-C would qualify (start_C - stop_A = 1, sites 0/1 are 0.5 A apart) but B breaks A's walk first
    assert pairs.tolist() == []
    ev, pairs = oc.collective_pairs(table[[0, 2]], frac, lat, 5, 1.0)
    assert pairs.tolist() == [[0, 1]] and oc.solo_count(2, pairs) == 0
    # same atom is never paired with itself
    same = table[[0, 2]].copy()
    same[:, 0] = 7
    assert oc.collective_pairs(same, frac, lat, 5, 1.0)[1].shape == (0, 2)


def test_oracle_rdf_edge_cases():
    """Self distance lands in bin 0 of the state-resolved RDF; distances beyond max_dist fall in the
    dropped overflow bin; a distance exactly on an edge belongs to the bin it closes (right=True)
    and, for np.histogram, to the bin it opens."""
    from oracle import rdf as orf

    lat = np.eye(3) * 10.0
    pos = np.array([[[0.0, 0.0, 0.0], [0.05, 0.0, 0.0], [0.5, 0.5, 0.5]]])  # 0.5 A and 8.66 A away
    codes = np.array([[orf.state_code(0, 0, 0)]])
    res = orf.radial_distribution(positions=pos, symbols=['Li', 'O', 'O'], floating_specie='Li', codes=codes,
                                  unique_labels=['a'], lattice=lat, max_dist=1.0, resolution=0.5)
    (x_li, y_li), (x_o, y_o) = res[('@a', 'Li')], res[('@a', 'O')]
    np.testing.assert_array_equal(x_li, [0.0, 0.5, 1.0])
    np.testing.assert_array_equal(y_li, [1, 0, 0])  # the atom itself, d = 0
    np.testing.assert_array_equal(y_o, [0, 1, 0])   # d = 0.5 closes bin (0, 0.5]; 8.66 A is dropped
    x, y = orf.between_species(coords_1=pos[:, :1], coords_2=pos[:, 1:], same=False, lattice=lat, max_dist=1.0,
                               resolution=0.5)
    np.testing.assert_array_equal(x, [0.0, 0.5])
    assert y[0] == 0.0 and y[1] > 0.0  # np.histogram: d = 0.5 opens [0.5, 1.0]


# ------------------------------------------------------------------ N1 (occupancy, residence, split)

N1_CASES = ['pipeline_c2_r09_inner05', 'pipeline_c2_dict_radius', 'pipeline_box_mixed']


@pytest.fixture(scope='module')
def n1_golden():
    from conftest import load_golden

    return load_golden('n1')


@pytest.mark.parametrize('name', N1_CASES)
def test_oracle_n1_vs_reference_golden(n1_golden, name):
    """oracle/sitestats.py against the outputs of the reference's transitions.py:280-410, 837-879."""
    from conftest import load_pipeline_case
    from oracle import sitestats

    g, c = n1_golden, load_pipeline_case(name)
    st, n_sites = c['states'], len(c['labels'])
    n_floating = int(c['floating_mask'].sum())
    occ = sitestats.occupancy(st, n_sites)
    assert_array_equal(occ, g[f'{name}/occupancy'])
    by_type = sitestats.by_label(occ, c['labels'], True)
    assert list(by_type) == [str(k) for k in g[f'{name}/by_type_keys']]
    assert_array_equal(np.array(list(by_type.values())), g[f'{name}/by_type_values'])
    loc = sitestats.by_label(occ, c['labels'], False, n_floating)
    assert list(loc) == [str(k) for k in g[f'{name}/locations_keys']]
    assert_array_equal(np.array(list(loc.values())), g[f'{name}/locations_values'])
    assert_array_equal(sitestats.residence_time(st), g[f'{name}/residence'])
    for n in g[f'{name}/splits']:
        bounds, _ = sitestats.split_bounds(len(st), int(n))
        parts = sitestats.split_events(c['events'], len(st), int(n))
        for i, (rows, src) in enumerate(parts):
            tag = f'{name}/split{n}/{i}'
            assert_array_equal(st[bounds[i]:bounds[i + 1]], g[f'{tag}/states'])
            assert_array_equal(rows, g[f'{tag}/events'])
            assert_array_equal(src, g[f'{tag}/events_index'])
            assert_array_equal(sitestats.occupancy(g[f'{tag}/states'], n_sites), g[f'{tag}/occupancy'])
            assert_array_equal(sitestats.residence_time(g[f'{tag}/states']), g[f'{tag}/residence'])
            mr = int(g[f'{name}/minimal_residence'])
            if f'{tag}/jumps' in g:
                got = jumps.transitions_to_jumps(rows, int(g[f'{tag}/n_frames_trajectory']), mr)
                assert_array_equal(got, g[f'{tag}/jumps'])
            else:
                assert str(g[f'{tag}/jumps_error']) == 'No jumps found'
                with pytest.raises(ValueError, match='No jumps found'):
                    jumps.transitions_to_jumps(rows, int(g[f'{tag}/n_frames_trajectory']), mr)
    with pytest.raises(ValueError) as exc:
        sitestats.split_events(c['events'], len(st), len(c['events']) + 1)
    assert str(exc.value) == str(g[f'{name}/split_error'])


# ------------------------------------------------------------------ minimum image in skewed cells

SKEWED_CELLS = {
    'gamma40': np.array([[10.0, 0, 0], [9.5 * np.cos(np.radians(40)), 9.5 * np.sin(np.radians(40)), 0], [1.0, 2.0, 8.0]]),
    'non_reduced_supercell': np.array([[5.0, 0, 0], [10.0, 5.0, 0], [15.0, 10.0, 7.0]]),
    'argyrodite_2x1x1': np.diag([19.848, 9.924, 9.924]),
}


def _brute_force_distances(m, f1, f2, reach=6):
    imgs = np.array([[i, j, k] for i in range(-reach, reach + 1) for j in range(-reach, reach + 1)
                     for k in range(-reach, reach + 1)], dtype=float)
    diff = f2[None, :, :] - f1[:, None, :]
    cart = (diff[:, :, None, :] + imgs[None, None, :, :]) @ m
    return np.sqrt((cart**2).sum(-1).min(-1))


@pytest.mark.parametrize('cell', list(SKEWED_CELLS))
def test_oracle_distances_are_minimum_image_in_skewed_cells(cell):
    """Lattice.get_all_distances searches the 27 images of the LLL-reduced basis (pymatgen
    pbc_shortest_vectors): equal to a brute-force search over 13^3 images of the raw basis,
    which the 27 raw-basis images are not for these cells."""
    m = SKEWED_CELLS[cell]
    rng = np.random.default_rng(3)
    f1, f2 = rng.random((25, 3)) * 3 - 1, rng.random((31, 3)) * 3 - 1
    got = Lattice(m).get_all_distances(f1, f2)
    assert_allclose(got, _brute_force_distances(m, f1, f2), rtol=0, atol=1e-12)
    if cell == 'non_reduced_supercell':
        a, b = f1 - np.floor(f1), f2 - np.floor(f2)
        raw = np.full(got.shape, np.inf)
        for img in np.array([[i, j, k] for i in (-1, 0, 1) for j in (-1, 0, 1) for k in (-1, 0, 1)], dtype=float):
            v = ((b[None, :, :] + img) - a[:, None, :]) @ m
            raw = np.minimum(raw, np.sqrt((v * v).sum(-1)))
        assert (raw > got + 1e-9).any(), 'this cell should defeat the raw-basis 27-image search'


def test_library_lll_equals_oracle_lll():
    """gd_lll_reduce_host (C++, the product's reduction) against oracle/pmg.py's restatement of
    pymatgen's _calculate_lll: same unimodular mapping on random, skewed and non-reduced cells."""
    from gemdat_b200 import _lib

    rng = np.random.default_rng(11)
    cells = list(SKEWED_CELLS.values())
    while len(cells) < 400:
        m = rng.normal(size=(3, 3)) * rng.uniform(1, 10)
        if len(cells) % 3 == 0:
            m = rng.integers(-3, 4, size=(3, 3)) @ m
        if abs(np.linalg.det(m)) > 1e-2:
            cells.append(m)
    for m in cells:
        m = np.ascontiguousarray(m, dtype=np.float64)
        lll, mapping = np.zeros(9), np.zeros(9)
        assert _lib.lib.gd_lll_reduce_host(m.ctypes.data, lll.ctypes.data, mapping.ctypes.data) == 0
        want_lll, want_map = Lattice(m)._calculate_lll()
        assert_array_equal(mapping.reshape(3, 3), want_map)
        assert_allclose(lll.reshape(3, 3), want_lll, rtol=1e-12, atol=1e-12)
        assert abs(abs(np.linalg.det(mapping.reshape(3, 3))) - 1.0) < 1e-9


# ------------------------------------------------------------------ drift / centre of mass
DRIFT_SELECTIONS = {'all': None, 'fixed_P_S': {'P', 'S'}, 'fixed_Br': {'Br'}, 'fixed_P_S_Br': {'P', 'S', 'Br'}}


def test_oracle_drift_equals_reference_golden():
    """oracle/drift.py against the reference's Trajectory.drift / apply_drift_correction /
    center_of_mass (trajectory.py:732-830; tests/golden/make_golden_drift.py): bit-exact."""
    from conftest import load_golden
    from gemdat_b200.elements import ATOMIC_MASS
    from oracle import drift as odrift

    g = load_golden('drift')
    species = [str(s) for s in g['species']]
    pos = g['coords']
    for tag, names in DRIFT_SELECTIONS.items():
        sel = [i for i, s in enumerate(species) if names is None or s in names]
        assert_array_equal(odrift.drift(pos, sel), g[f'drift/{tag}'])
        assert_array_equal(odrift.corrected_steps(pos, sel), g[f'corrected/{tag}/displacements'])
        assert_array_equal(odrift.corrected_positions(pos, sel), g[f'corrected/{tag}/positions'])
    assert_array_equal(odrift.center_of_mass(pos, [ATOMIC_MASS[s] for s in species]), g['com/coords'])
    # the reference's floating_species route averages over zero atoms (see the generator's header)
    assert np.isnan(g['drift/floating_Li'][1:]).all() and np.isnan(g['drift/floating_Li_Br'][1:]).all()


# ------------------------------------------------------------------ species RDF, one lattice per frame
RDF_FRAME_CASES = ['diag', 'sheared', 'skewed']


@pytest.mark.parametrize('name', RDF_FRAME_CASES)
def test_oracle_species_rdf_with_per_frame_lattices_equals_reference_golden(name):
    """oracle/rdf.py::between_species with one cell per frame against the reference's
    radial_distribution_between_species on a constant_lattice=False trajectory
    (rdf.py:232-261; tests/golden/make_golden_rdf_frames.py)."""
    from conftest import load_golden
    from oracle import rdf as orf

    g = load_golden('rdf_frames')
    species = [str(s) for s in g['species']]
    pos = g['positions']
    sel = lambda want: pos[:, [s in want for s in species]]  # noqa: E731
    for n in (0, 1, 2):
        t = f'{name}_between{n}'
        s1, s2 = g[f'{t}_specie_1'].tolist(), g[f'{t}_specie_2'].tolist()
        w1 = [s1] if isinstance(s1, str) else s1
        w2 = [s2] if isinstance(s2, str) else s2
        x, y = orf.between_species(coords_1=sel(w1), coords_2=sel(w2), same=(s1 == s2),
                                   lattice=g[f'{name}_lattices'], max_dist=float(g[f'{t}_max_dist']),
                                   resolution=float(g[f'{t}_resolution']))
        np.testing.assert_array_equal(x, g[f'{t}_x'])
        np.testing.assert_allclose(y, g[f'{t}_y'], rtol=1e-13)
