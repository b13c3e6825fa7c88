"""Parity of the CUDA path (through the C ABI / ctypes, via the public API) against the
fixtures produced by executing the reference and against the oracle.  Integer and
index results are compared bit-exactly; floating-point results within the
tolerance written at each assertion (north_star: 1e-10 relative)."""

from __future__ import annotations

import warnings

import numpy as np
import pytest
from numpy.testing import assert_allclose, assert_array_equal

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')


@pytest.fixture(scope='module', autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')


def _product(case):
    from gemdat_b200 import Sites, Trajectory

    species = [str(s) for s in case['species']] if 'species' in case else ['Li'] * case['positions'].shape[1]
    traj = Trajectory(species=species, coords=case['positions'], lattice=case['lattice'],
                      metadata={'temperature': 650.0}, time_step=2e-15)
    sites = Sites(case['lattice'], 'Li', case['site_frac'], labels=case['labels'])
    return traj, sites


def _radius_args(case):
    name = case['name']
    if name == 'pipeline_c2_default_radius':
        return None, 1.0
    if list(case['site_radius']) == ['']:
        return case['site_radius'][''], case['inner_fraction']['']
    return case['site_radius'], case['inner_fraction']


def test_pipeline_matches_reference(pipeline_case):
    c = pipeline_case
    traj, sites = _product(c)
    radius, inner = _radius_args(c)
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter('always')
        tr = traj.transitions_between_sites(
            sites, 'Li', site_radius=radius, site_inner_fraction=inner,
            fraction_of_overlap=float(c['fraction_of_overlap']),
        )
    assert sorted(str(w.message) for w in caught) == sorted(str(w) for w in c['warnings'])
    if radius is None:
        assert_allclose(tr.site_radius.radius[''], c['site_radius'][''], rtol=1e-12)
    assert tr.states.dtype == np.int64 and tr.states.shape == c['states'].shape
    assert_array_equal(tr.states, c['states'])
    assert_array_equal(tr.inner_states, c['inner_states'])
    assert list(tr.events.columns) == [str(x) for x in c['event_columns']]
    assert_array_equal(tr.events.to_numpy(), c['events'])
    assert tr.n_events == len(c['events']) and tr.n_states == len(c['states'])
    assert_array_equal(tr.matrix(), c['matrix'])
    assert_array_equal(tr.states_next(), c['states_next'])
    assert_array_equal(tr.states_prev(), c['states_prev'])
    for key in [k for k in c if k.startswith('jumps_') and k[6:].isdigit()]:
        mr = int(key[6:])
        jumps = tr.jumps(minimal_residence=mr)
        assert list(jumps.data.columns) == [str(x) for x in c['jump_columns']]
        assert_array_equal(jumps.data.to_numpy(), c[key])
        assert_array_equal(jumps.matrix(), c[f'{key}_matrix'])
        assert_allclose(float(jumps.jump_diffusivity(3)), float(c[f'{key}_diffusivity']), rtol=1e-10)


def test_private_seams_match_reference(pipeline_case):
    """_calculate_atom_states / _calculate_transition_events keep the reference signatures."""
    from gemdat_b200.transitions import _calculate_atom_states, _calculate_transition_events

    c = pipeline_case
    traj, sites = _product(c)
    diff = traj.filter('Li')
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        inner = _calculate_atom_states(sites, diff, c['site_radius'], c['inner_fraction'])
    assert_array_equal(inner, c['inner_states'])
    ev = _calculate_transition_events(atom_sites=c['states'], atom_inner_sites=c['inner_states'])
    assert_array_equal(ev.to_numpy(), c['events'])


@pytest.mark.parametrize('tag', ['unit', 'walk', 'c2'])
def test_displacement_path(displacement_golden, tag):
    from gemdat_b200 import Trajectory, TrajectoryMetrics
    from oracle import msd as omsd

    g = displacement_golden
    pos, lat, ts = g[f'{tag}_positions'], g[f'{tag}_lattice'], float(g[f'{tag}_time_step'])
    traj = Trajectory(species=['Li'] * pos.shape[1], coords=pos, lattice=lat, time_step=ts)
    # integer-winding unwrap vs the reference's sequential cumsum: 1e-12 absolute (fractional units)
    assert_allclose(traj.cumulative_displacements, g[f'{tag}_cumulative'], rtol=0, atol=1e-12)
    scale = max(1.0, float(np.abs(g[f'{tag}_distances']).max()))
    assert_allclose(traj.distances_from_base_position(), g[f'{tag}_distances'], rtol=1e-10, atol=1e-11 * scale)
    m = TrajectoryMetrics(traj)
    assert_allclose(m.speed(), g[f'{tag}_speed'], rtol=0, atol=1e-11 * scale)
    assert_allclose(float(m.tracer_diffusivity(dimensions=3)), float(g[f'{tag}_tracer_diffusivity']), rtol=1e-10)
    amp = m.amplitudes()
    assert amp.shape == g[f'{tag}_amplitudes'].shape
    assert_allclose(amp, g[f'{tag}_amplitudes'], rtol=0, atol=1e-10 * scale)
    assert_allclose(float(m.vibration_amplitude()), float(g[f'{tag}_vibration_amplitude']), rtol=1e-10, atol=1e-12)
    # MSD: vs the compensated direct oracle at 1e-10 relative; vs the reference's FFT form at the
    # FFT form's own accuracy (it loses ~<r^2>/msd digits at small lags)
    msd = traj.mean_squared_displacement()
    direct = omsd.msd_direct(pos, lat)
    assert msd.shape == g[f'{tag}_msd'].shape
    assert_allclose(msd, direct, rtol=1e-10, atol=1e-18)
    peak = max(1.0, float(np.abs(g[f'{tag}_msd']).max()))
    assert_allclose(msd, g[f'{tag}_msd'], rtol=1e-9, atol=1e-11 * peak)


@pytest.mark.parametrize('tag,res', [('unit', 0.2), ('walk', 0.2), ('walk', 0.5), ('c2', 0.2), ('c2', 0.5)])
def test_density_volume(displacement_golden, tag, res):
    from gemdat_b200 import Trajectory, trajectory_to_volume

    g = displacement_golden
    pos, lat = g[f'{tag}_positions'], g[f'{tag}_lattice']
    traj = Trajectory(species=['Li'] * pos.shape[1], coords=pos, lattice=lat, time_step=1.0)
    vol = trajectory_to_volume(traj, resolution=res)
    assert vol.data.shape == g[f'{tag}_volume_{res}'].shape
    assert_array_equal(vol.data, g[f'{tag}_volume_{res}'])
    assert_array_equal(traj.to_volume(resolution=res).data, vol.data)


@pytest.mark.parametrize('tag', ['unit', 'walk', 'c2'])
def test_attempt_frequency(displacement_golden, tag):
    """metrics.py:167-188 / utils.py:128-160 (scipy.signal.periodogram) --
    goldens from executing the reference (tests/golden/make_golden_attempt.py).  The spectrum is
    fp64 cuFFT vs pocketfft: 1e-10 relative; atoms that never move give 0/0 = NaN in both."""
    from gemdat_b200 import Trajectory
    from gemdat_b200.metrics import TrajectoryMetrics

    g = displacement_golden
    from conftest import load_golden

    a = load_golden('attempt_frequency')
    p = g[f'{tag}_positions']
    traj = Trajectory(species=['Li'] * p.shape[1], coords=p, lattice=g[f'{tag}_lattice'],
                      time_step=float(g[f'{tag}_time_step']))
    m = TrajectoryMetrics(traj)
    with np.errstate(invalid='ignore'):
        freq, std = m.attempt_frequency()
    assert_allclose(m._meanfreq(), a[f'{tag}_meanfreq'], rtol=1e-10, equal_nan=True)
    assert_allclose(float(freq), float(a[f'{tag}_attempt_frequency']), rtol=1e-10, equal_nan=True)
    assert_allclose(float(std), float(a[f'{tag}_attempt_frequency_std']), rtol=1e-9, equal_nan=True)


def test_density_rejects_out_of_range():
    from gemdat_b200 import Trajectory

    pos = np.full((3, 2, 3), 0.5)
    # np.mod(-1e-20, 1) == 1.0: the one way a coordinate survives to_positions outside [0, 1)
    # (a plain 1.0 is wrapped to 0.0 by trajectory.py:126-127 before volume.py:476-477 sees it)
    pos[1, 1, 2] = -1e-20
    traj = Trajectory(species=['Li'] * 2, coords=pos, lattice=np.eye(3) * 4, time_step=1.0)
    traj.coords_are_displacement = False
    with pytest.raises(AssertionError):
        traj.to_volume(0.5)


def test_fill_and_matrix_helpers(utils_golden):
    from gemdat_b200.transitions import _calculate_transitions_matrix, bfill, ffill
    import pandas as pd

    u = utils_golden
    assert_array_equal(ffill(u['arr']), u['ffill'])
    assert_array_equal(bfill(u['arr']), u['bfill'])
    assert_array_equal(ffill(u['arr'], axis=0), u['ffill0'])
    assert_array_equal(bfill(u['arr'], axis=0), u['bfill0'])
    assert_array_equal(ffill(u['rnd'], axis=0), u['rnd_ffill0'])
    assert_array_equal(bfill(u['rnd'], axis=0), u['rnd_bfill0'])
    assert_array_equal(ffill(u['rnd']), u['rnd_ffill1'])
    assert_array_equal(bfill(u['rnd']), u['rnd_bfill1'])
    ev = pd.DataFrame(u['matrix_events'], columns=['atom index', 'start site', 'destination site', 'time'])
    m = _calculate_transitions_matrix(ev, n_sites=3)
    assert m.dtype == int
    assert_array_equal(m, u['matrix'])


def test_event_and_jump_error_paths():
    """transitions.py:442-443, 451, 470 and jumps.py:107-108."""
    from gemdat_b200.transitions import _calculate_transition_events

    const = np.zeros((6, 2), dtype=np.int64)
    with pytest.raises(ValueError):
        _calculate_transition_events(atom_sites=const, atom_inner_sites=const)
    outer = const.copy()
    outer[3:, 0] = 1
    with pytest.raises(IndexError):
        _calculate_transition_events(atom_sites=outer, atom_inner_sites=const)
    inner = const.copy()
    inner[2, 1] = 5
    inner[4:, 0] = 1
    ev = _calculate_transition_events(atom_sites=outer, atom_inner_sites=inner)
    assert_array_equal(ev.to_numpy(), [[0, 0, 1, 0, 0, 2], [0, 1, 1, 0, 1, 3]])


def test_no_jumps_found():
    from gemdat_b200 import Sites, Trajectory

    # one atom leaves its site and never arrives anywhere: events but no jumps
    pos = np.full((8, 1, 3), 0.1)
    pos[4:, 0, 0] = 0.5
    traj = Trajectory(species=['Li'], coords=pos, lattice=np.eye(3) * 10.0, time_step=1e-15)
    sites = Sites(np.eye(3) * 10.0, 'Li', [[0.1, 0.1, 0.1], [0.9, 0.9, 0.9]], labels=['a', 'a'])
    tr = traj.transitions_between_sites(sites, 'Li', site_radius=0.5)
    assert_array_equal(tr.events.to_numpy(), [[0, 0, -1, 0, -1, 3]])
    with pytest.raises(ValueError, match='No jumps found'):
        tr.jumps()


@pytest.mark.parametrize('n_frames,n_atoms', [(1, 1), (2, 3), (7, 1), (511, 1), (513, 3), (1025, 7)])
def test_ragged_sizes_against_oracle(n_frames, n_atoms):
    """Odd atom-frame counts exercise the 16-byte tail of the TMA bulk copies and partial chunks."""
    from gemdat_b200 import Sites, Trajectory
    from oracle import cstates

    rng = np.random.default_rng(n_frames * 31 + n_atoms)
    lat = np.diag([6.0, 7.5, 5.25])
    site_frac = np.array([[0.1, 0.1, 0.1], [0.6, 0.6, 0.6], [0.98, 0.02, 0.5], [0.3, 0.8, 0.99]])
    pos = rng.random((n_frames, n_atoms, 3))
    traj = Trajectory(species=['Li'] * n_atoms, coords=pos, lattice=lat, time_step=1e-15)
    sites = Sites(lat, 'Li', site_frac, labels=['s'] * 4)
    from gemdat_b200.transitions import _site_states_device, widen_to_host

    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        o, i, _ = _site_states_device(sites=sites, trajectory=traj, site_radius={'': 1.4},
                                      site_inner_fraction={'': 0.6})
        eo, ei = cstates.calculate_atom_states_pair(
            site_frac=site_frac, site_labels=['s'] * 4, positions=pos, lattice_matrix=lat,
            site_radius={'': 1.4}, site_inner_fraction={'': 0.6})
    assert_array_equal(widen_to_host(o), eo)
    assert_array_equal(widen_to_host(i), ei)


def test_synth_device_equals_host_replay():
    from gemdat_b200 import synth

    rec = synth.config_c2(n_frames=3000, seed=5)
    dev = rec.device_positions().cpu().numpy()
    assert_array_equal(dev, rec.host_positions())
    window = rec.device_positions(1999, 300).cpu().numpy()
    assert_array_equal(window, rec.host_positions(1999, 300))


def test_c2_20k_frames_against_oracles():
    """BASELINE config 2 shape (48 Li, 96 argyrodite sites) at 2*10^4 frames: states vs the C
    oracle, events / jumps vs the numpy oracle, boundary-band counts vs the numpy oracle."""
    from gemdat_b200 import synth
    from oracle import cstates, events as oevents, jumps as ojumps, states as ostates

    rec = synth.config_c2(n_frames=20_000, seed=20260101)
    traj = rec.trajectory(device=True)
    tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=0.9, site_inner_fraction=0.5)
    pos = traj.positions.cpu().numpy()
    kw = dict(site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=pos,
              lattice_matrix=rec.lattice, site_radius={'': 0.9})
    eo, ei = cstates.calculate_atom_states_pair(**kw, site_inner_fraction={'': 0.5})
    assert_array_equal(tr.states, eo)
    assert_array_equal(tr.inner_states, ei)
    ev = oevents.calculate_transition_events(eo, ei)
    assert_array_equal(tr.events.to_numpy(), ev)
    for mr in (0, 3, 100):
        assert_array_equal(tr.jumps(minimal_residence=mr).data.to_numpy(),
                           ojumps.transitions_to_jumps(ev, len(traj), mr))
    sub = dict(kw, positions=pos[:2000])
    b12, bf32 = ostates.boundary_band_counts(**sub, site_inner_fraction={'': 1.0})
    tr_sub = traj[:2000].transitions_between_sites(rec.sites, 'Li', site_radius=0.9, site_inner_fraction=1.0)
    assert tr_sub.info['band_1e-12'][0] == b12
    assert tr_sub.info['band_4e-6'][0] == bf32


def test_c3_two_label_radii_full_parity():
    """BASELINE config 3 shape (1008 Li, 2160 sites), per-label radii (SURVEY.md Q1/Q2 paths):
    whole-trajectory comparison with the C oracle, including the integer_remap quirk."""
    from gemdat_b200 import synth
    from oracle import cstates

    rec = synth.config_c3(n_frames=1500)
    traj = rec.trajectory(device=True)
    radius = {'24d': 0.6, '96h': 0.5}
    inner = {'24d': 0.5, '96h': 0.8}
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=radius, site_inner_fraction=inner)
        eo, ei = cstates.calculate_atom_states_pair(
            site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels,
            positions=traj.positions.cpu().numpy(), lattice_matrix=rec.lattice,
            site_radius=radius, site_inner_fraction=inner)
    assert not tr.info['visited_outer'].all()  # short run: the remap quirk is active
    assert_array_equal(tr.states, eo)
    assert_array_equal(tr.inner_states, ei)


def test_k1_repeatable_under_pipeline_pressure():
    """The TMA ring of K1 is refilled by the async proxy while warps read it through the
    generic proxy; a missing cross-proxy fence showed up as whole warps reading the data of a
    later chunk, a few times in ten runs.  Many short launches + one long one, all identical."""
    from gemdat_b200 import synth
    from gemdat_b200.transitions import _site_states_device
    from oracle import cstates

    rec = synth.config_c3(n_frames=1500)
    traj = rec.trajectory(device=True)
    kw = dict(sites=rec.sites, trajectory=traj, site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
    eo, ei = cstates.calculate_atom_states_pair(
        site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels,
        positions=traj.positions.cpu().numpy(), lattice_matrix=rec.lattice,
        site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
    eo_d = torch.from_numpy(eo.astype(np.int32)).cuda()
    ei_d = torch.from_numpy(ei.astype(np.int32)).cuda()
    for _ in range(25):
        o, i, _ = _site_states_device(**kw)
        assert bool((o == eo_d).all()) and bool((i == ei_d).all())
    big = synth.config_c3(n_frames=40_000).trajectory(device=True)
    kw['trajectory'] = big
    o1, i1, _ = _site_states_device(**kw)
    o1, i1 = o1.clone(), i1.clone()
    for _ in range(3):
        o2, i2, _ = _site_states_device(**kw)
        assert bool((o1 == o2).all()) and bool((i1 == i2).all())


def test_c3_shape_properties_and_window_parity():
    """BASELINE config 3 shape on a device-resident trajectory: sampled frame windows vs the C
    oracle (replayed on the host from the recipe), plus size-independent invariants of the
    event / jump tables."""
    from gemdat_b200 import synth
    from oracle import cstates

    rec = synth.config_c3(n_frames=20_000)
    traj = rec.trajectory(device=True)
    tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=0.6, site_inner_fraction=0.5)
    states, inner_states = tr.states, tr.inner_states
    n_sites = len(rec.sites)
    assert states.min() >= -1 and states.max() < n_sites
    # inner sphere inside the outer sphere of the same site -> inner state is the outer one or -1
    assert np.all((inner_states == -1) | (inner_states == states))
    for frame0 in (0, 7777, 19_900):
        pos = rec.host_positions(frame0, 100)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            eo, ei = cstates.calculate_atom_states_pair(
                site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=pos,
                lattice_matrix=rec.lattice, site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
        assert_array_equal(states[frame0:frame0 + 100], eo)
        assert_array_equal(inner_states[frame0:frame0 + 100], ei)
    ev = tr.events.to_numpy()
    # every event row is consistent with the state arrays it was cut from
    a, t = ev[:, 0], ev[:, 5]
    assert_array_equal(states[t, a], ev[:, 1])
    assert_array_equal(states[t + 1, a], ev[:, 2])
    assert_array_equal(inner_states[t, a], ev[:, 3])
    assert_array_equal(inner_states[t + 1, a], ev[:, 4])
    assert np.all((ev[:, 1] != ev[:, 2]) | (ev[:, 3] != ev[:, 4]))
    assert np.all(np.diff(a) >= 0) and np.all(np.diff(t)[np.diff(a) == 0] > 0)
    changed = (states[1:] != states[:-1]) | (inner_states[1:] != inner_states[:-1])
    keep = (states[1:] != states[:-1]).any(axis=0)  # atoms whose outer state never changes are skipped
    assert len(ev) == int(changed[:, keep].sum())
    jumps = tr.jumps().data.to_numpy()
    assert np.all(jumps[:, 1] != jumps[:, 2]) and np.all(jumps[:, 4] > jumps[:, 3])
    assert int(tr.jumps().matrix().sum()) == len(jumps)
    # the whole event table and the jump tables against the oracle's restatement of
    # transitions.py:413-482 and jumps.py:24-119 on the downloaded states
    from oracle import events as oevents
    from oracle import jumps as ojumps

    assert_array_equal(ev, oevents.calculate_transition_events(states, inner_states))
    for mr in (0, 3, 100):
        assert_array_equal(tr.jumps(minimal_residence=mr).data.to_numpy(),
                           ojumps.transitions_to_jumps(ev, len(traj), mr))


def test_c3_full_size_properties():
    """BASELINE config 3 at its FULL size (1008 Li x 10^6 frames, 24 GB of positions resident in
    HBM) through the device-resident route the bench times.  Size-independent properties checked on
    the device: every event row agrees with the state arrays it was cut from, the row count is the
    number of state changes of the non-skipped atoms, rows are sorted by (atom, time), jumps are
    well formed; the checked, optimistic and graph-replayed passes give identical tables; sampled
    frame windows equal the C oracle; the jump rows of sampled atoms equal the oracle's state
    machine run on their event slices."""
    from gemdat_b200 import synth
    from gemdat_b200.pipeline import HotPath
    from oracle import cstates

    free, _ = torch.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip('needs ~50 GB of free HBM')
    T = 1_000_000
    rec = synth.config_c3(n_frames=T)
    pos = rec.device_positions(0, T)
    hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=rec.n_atoms, site_radius=0.6,
                 site_inner_fraction=0.5, minimal_residence=3, n_frames_total=T)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        first = hp.run(pos)
        ev, jumps = first['events'].clone(), first['jumps'].clone()
        outer, inner = first['outer'], first['inner']
        assert hp.last_events_route == 'staged'
        for expect in ('eager', 'graph', 'graph'):
            again = hp.run(pos)
            assert hp.last_pass == 'optimistic' and hp.last_launch_mode == expect, hp.last_optimistic_reject
            assert torch.equal(again['events'], ev) and torch.equal(again['jumps'], jumps)
    # jump rows of sampled atoms against the oracle's state machine (jumps.py:24-119) on their
    # event slices (an atom's jumps depend on its own events only)
    from oracle import jumps as ojumps

    off = first['atom_offsets'].cpu().numpy()
    jumps_h = jumps.cpu().numpy().astype(np.int64)
    for atom in (0, 1, 333, 640, rec.n_atoms - 1):
        rows = ev[off[atom]:off[atom + 1]].cpu().numpy().astype(np.int64)
        assert len(rows) > 1000
        try:
            want = ojumps.transitions_to_jumps(rows, T, 3)
        except ValueError:
            want = np.zeros((0, 6), dtype=np.int64)
        assert_array_equal(jumps_h[jumps_h[:, 0] == atom], want)
    a, t = ev[:, 0].long(), ev[:, 5].long()
    # rows agree with the states they were cut from, and each is a real change
    assert torch.equal(outer[t, a], ev[:, 1]) and torch.equal(outer[t + 1, a], ev[:, 2])
    assert torch.equal(inner[t, a], ev[:, 3]) and torch.equal(inner[t + 1, a], ev[:, 4])
    assert bool(((ev[:, 1] != ev[:, 2]) | (ev[:, 3] != ev[:, 4])).all())
    # sorted by (atom, time), strictly
    assert bool(((a[1:] > a[:-1]) | ((a[1:] == a[:-1]) & (t[1:] > t[:-1]))).all())
    del a, t
    # row count = state changes of the atoms whose outer state changes at all (transitions.py:442)
    d_o = outer[1:] != outer[:-1]
    keep = d_o.any(dim=0)
    n_changes = int(((d_o | (inner[1:] != inner[:-1]))[:, keep]).sum())
    del d_o
    assert n_changes == ev.shape[0]
    # jumps: start != destination, stop > start, residence >= minimal_residence, sorted by atom
    j = jumps.long()
    assert bool((j[:, 1] != j[:, 2]).all()) and bool((j[:, 4] > j[:, 3]).all()) and bool((j[:, 5] >= 3).all())
    assert bool((j[1:, 0] >= j[:-1, 0]).all())
    # inner sphere inside the outer sphere of the same site
    assert bool(((inner == -1) | (inner == outer)).all())
    for frame0 in (0, 499_990, 999_950):
        p = rec.host_positions(frame0, 50)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            eo, ei = cstates.calculate_atom_states_pair(
                site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=p,
                lattice_matrix=rec.lattice, site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
        assert_array_equal(outer[frame0:frame0 + 50].cpu().numpy(), eo)
        assert_array_equal(inner[frame0:frame0 + 50].cpu().numpy(), ei)


def _jumps_via_abi(events: np.ndarray, n_atoms: int, n_frames: int, minimal_residence: int,
                   atom_base: int = 0) -> np.ndarray:
    """gd_jumps_count / gd_jumps_fill on a raw [E, 6] event table (sorted by atom, time).
    ``atom_base``: the table is an atom block whose first atom has that global index."""
    import ctypes as C

    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib

    ev = torch.from_numpy(np.ascontiguousarray(events, dtype=np.int32)).cuda()
    counts = np.bincount(events[:, 0] - atom_base, minlength=n_atoms)
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)).cuda()
    work = torch.empty(int(_lib.lib.gd_jumps_work_bytes(len(events), n_atoms)), dtype=torch.uint8, device='cuda')
    n = C.c_int64(0)
    s = dv.stream_ptr()
    _lib.check(_lib.lib.gd_jumps_count(ev.data_ptr(), off.data_ptr(), len(events), n_atoms, n_frames,
                                       minimal_residence, work.data_ptr(), C.byref(n), s))
    out = torch.empty((max(n.value, 1), 6), dtype=torch.int32, device='cuda')
    if n.value:
        _lib.check(_lib.lib.gd_jumps_fill(ev.data_ptr(), off.data_ptr(), len(events), n_atoms, n_frames,
                                          minimal_residence, work.data_ptr(), out.data_ptr(), s))
    return out[: n.value].cpu().numpy().astype(np.int64)


def _random_event_stream(rng, n_events, n_sites, p_stay, consistent=True):
    """Event rows of ONE atom: a random walk over outer/inner states (NOSITE = -1 included)."""
    rows = []
    s, i = int(rng.integers(-1, n_sites)), -1
    t = 0
    for _ in range(n_events):
        t += int(rng.integers(1, 4))
        if rng.random() < p_stay:  # inner-only change at the same outer state
            s1 = s
            i1 = -1 if (i != -1 or s == -1) else s
            if i1 == i:
                s1 = -1 if s != -1 else int(rng.integers(0, n_sites))
                i1 = -1
        else:
            s1 = int(rng.integers(-1, n_sites))
            if s1 == s:
                s1 = -1 if s != -1 else 0
            i1 = s1 if (s1 != -1 and rng.random() < 0.5) else -1
        s0 = s if consistent else int(rng.integers(-1, n_sites))
        rows.append((0, s0, s1, i, i1, t))
        s, i = s1, i1
    return rows, t + 2


@pytest.mark.parametrize('mr', [0, 2, 7])
def test_jump_machine_time_parallel_segments(mr):
    """K3 runs 512-event segments of an atom in parallel (speculate / chain / fill).  Streams
    that cross many segment boundaries -- frequent departures, no departure at all for
    thousands of events (chain falls back to sequential), and inconsistent rows (start state
    != previous destination) -- must all equal the sequential oracle."""
    from oracle import jumps as ojumps

    rng = np.random.default_rng(99 + mr)
    streams = [
        _random_event_stream(rng, 5000, 6, 0.5),
        _random_event_stream(rng, 3000, 3, 0.97),     # long stretches without a departure
        _random_event_stream(rng, 1537, 4, 0.2),
        _random_event_stream(rng, 2100, 5, 0.6, consistent=False),
        _random_event_stream(rng, 512, 4, 0.5),       # exactly one segment
        _random_event_stream(rng, 513, 4, 0.5),
        _random_event_stream(rng, 1, 4, 0.5),
    ]
    # atom 2 has no events at all
    atom_ids = [0, 1, 3, 4, 5, 6, 7]
    rows, n_frames = [], 0
    for a, (r, t_end) in zip(atom_ids, streams):
        rows += [(a,) + x[1:] for x in r]
        n_frames = max(n_frames, t_end)
    events = np.array(rows, dtype=np.int64)
    got = _jumps_via_abi(events, 8, n_frames, mr)
    try:
        exp = ojumps.transitions_to_jumps(events, n_frames, mr)
    except ValueError:
        exp = np.zeros((0, 6), dtype=np.int64)
    assert len(exp) > 50
    assert_array_equal(got, exp)


def test_jump_rows_of_an_atom_block_keep_global_atom_index():
    """After the event exchange a rank holds an atom BLOCK (block-local offsets, global atom
    column); candidates inherited across a segment boundary must still report the global index."""
    from oracle import jumps as ojumps

    rng = np.random.default_rng(77)
    rows = []
    for a in (24, 25, 26):
        r, t_end = _random_event_stream(rng, 1500, 5, 0.5)
        rows += [(a,) + x[1:] for x in r]
    events = np.array(rows, dtype=np.int64)
    n_frames = int(events[:, 5].max()) + 2
    exp = ojumps.transitions_to_jumps(events, n_frames, 2)
    assert len(exp) > 100
    assert_array_equal(_jumps_via_abi(events, 3, n_frames, 2, atom_base=24), exp)


def _jumps_frame_shards_via_abi(events, n_atoms, n_frames, minimal_residence, cuts):
    """The frame-sharded K3 protocol of include/gemdat_b200.h on ONE device: the event table is
    cut at the frame numbers ``cuts`` into time slices (each sorted by atom, time), every slice
    runs gd_jumps_count_shard from the empty state, the per-atom records are "gathered" (stacked),
    gd_jumps_resolve gives each slice its entry records (repeating while some are unknown), the
    final pass counts and fills, and the slices' jumps are merged per atom.  Returns
    (jumps, resolve rounds, OR of the mismatch flags)."""
    import ctypes as C

    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from gemdat_b200.pipeline import merge_plan

    lib, s = _lib.lib, dv.stream_ptr()
    nb = _lib.GD_JUMP_STATE_BYTES
    edges = [-(1 << 40)] + list(cuts) + [1 << 40]
    W = len(edges) - 1
    shards = []
    for r in range(W):
        ev = events[(events[:, 5] >= edges[r]) & (events[:, 5] < edges[r + 1])]
        d_ev = torch.from_numpy(np.ascontiguousarray(ev, dtype=np.int32)).cuda()
        off = np.concatenate([[0], np.cumsum(np.bincount(ev[:, 0], minlength=n_atoms))]).astype(np.int64)
        work = torch.empty(int(lib.gd_jumps_work_bytes(len(ev), n_atoms)), dtype=torch.uint8, device='cuda')
        shards.append(dict(ev=d_ev, n=len(ev), off=torch.from_numpy(off).cuda(), work=work,
                           exit=torch.zeros((n_atoms, nb), dtype=torch.uint8, device='cuda'),
                           entry=torch.zeros((n_atoms, nb), dtype=torch.uint8, device='cuda')))

    def count(r, entry, published=None, mode='async', reuse=1):
        """mode: 'async' (first pass: exit records only, no read-back), 'checked' (later rounds:
        the chain's proof flag is read back), 'final' (counts + per-atom offsets + published check)."""
        sh = shards[r]
        n, fl = C.c_int64(0), C.c_int32(0)
        final = mode == 'final'
        atom_off = torch.zeros(n_atoms + 1, dtype=torch.int64, device='cuda') if final else None
        _lib.check(lib.gd_jumps_count_shard(
            dv.ptr(sh['ev']) if sh['n'] else None, dv.ptr(sh['off']), sh['n'], n_atoms, n_frames,
            minimal_residence, dv.ptr(entry), 1 if r == W - 1 else 0, reuse, dv.ptr(published), dv.ptr(sh['work']),
            dv.ptr(sh['exit']), dv.ptr(atom_off), C.byref(n) if final else None,
            None if mode == 'async' else C.byref(fl), s))
        return n.value, fl.value, atom_off

    for r in range(W):  # shard 0 truly starts empty; zeroed records = "empty, not known"
        count(r, shards[r]['entry'] if r else None, reuse=0)
    published = None
    rounds = 0
    while True:
        allrec = torch.stack([sh['exit'] for sh in shards]).contiguous()
        if published is None:
            published = allrec.clone()
        unresolved = []
        for r in range(W):
            # the count is left in the shard's work area and comes back with the next pass's flags
            d_unres = shards[r]['work'].data_ptr() + int(lib.gd_jumps_work_unresolved_offset(shards[r]['n'], n_atoms))
            u = C.c_int32(0)
            _lib.check(lib.gd_jumps_resolve(dv.ptr(allrec), W, r, n_atoms, dv.ptr(shards[r]['entry']),
                                            d_unres, C.byref(u), s))
            unresolved.append(u.value)
        assert len(set(unresolved)) == 1, 'every rank must see the same number of unresolved entries'
        rounds += 1
        if unresolved[0] == 0:
            break
        assert rounds < W, 'states must resolve in at most world - 1 extra rounds'
        for r in range(W):
            _, fl, _ = count(r, shards[r]['entry'], mode='checked')
            assert fl & 4, 'the unresolved count must come back with the flags of the next pass'
    tables, offs, mismatch = [], [], 0
    for r in range(W):
        sh = shards[r]
        n, fl, atom_off = count(r, sh['entry'], published[r].contiguous(), mode='final')
        assert not fl & 4
        mismatch |= fl & 2
        out = torch.empty((max(n, 1), 6), dtype=torch.int32, device='cuda')
        if n:
            _lib.check(lib.gd_jumps_fill(dv.ptr(sh['ev']) if sh['n'] else None, dv.ptr(sh['off']), sh['n'],
                                         n_atoms, n_frames, minimal_residence, dv.ptr(sh['work']),
                                         dv.ptr(out), s))
        tables.append(out[:n].cpu().numpy().astype(np.int64))
        offs.append(atom_off.cpu().numpy())
    offs = np.array(offs)
    assert [len(t) for t in tables] == offs[:, -1].tolist()
    total_off, dst = merge_plan(np.diff(offs, axis=1))
    merged = np.zeros((int(total_off[-1]), 6), dtype=np.int64)
    for r in range(W):
        for a in range(n_atoms):
            k = offs[r, a + 1] - offs[r, a]
            merged[dst[r, a]: dst[r, a] + k] = tables[r][offs[r, a]: offs[r, a + 1]]
    return merged, rounds, mismatch


@pytest.mark.parametrize('mr', [0, 3])
@pytest.mark.parametrize('n_shards', [2, 3, 8])
def test_jump_machine_frame_shards(mr, n_shards):
    """K3 over frame shards: only the machine's per-atom state crosses a shard boundary.  Atoms
    with thousands of events per shard (one round), atoms whose shard holds events but no
    departure (extra rounds), atoms without events in some or all shards, and candidates that
    stay open across several shards must all give the sequential oracle's table."""
    from oracle import jumps as ojumps

    rng = np.random.default_rng(5 + 17 * mr + n_shards)
    streams = [
        _random_event_stream(rng, 6000, 6, 0.5),
        _random_event_stream(rng, 4000, 3, 0.97),    # long stretches without a departure
        _random_event_stream(rng, 40, 4, 0.9),       # a few events per shard, mostly inner-only
        _random_event_stream(rng, 9, 4, 0.3),
        _random_event_stream(rng, 1, 4, 0.5),
        _random_event_stream(rng, 2500, 5, 0.6),
    ]
    atom_ids = [0, 1, 3, 4, 5, 7]                    # atoms 2 and 6 have no events at all
    rows, n_frames = [], 0
    for k, (a, (r, t_end)) in enumerate(zip(atom_ids, streams)):
        stretch = [1, 1, 150, 700, 1, 2][k]          # sparse atoms span the whole time axis
        rows += [(a,) + x[1:5] + (x[5] * stretch,) for x in r]
        n_frames = max(n_frames, t_end * stretch)
    events = np.array(rows, dtype=np.int64)
    cuts = [n_frames * (k + 1) // n_shards for k in range(n_shards - 1)]
    got, rounds, mismatch = _jumps_frame_shards_via_abi(events, 8, n_frames, mr, cuts)
    exp = ojumps.transitions_to_jumps(events, n_frames, mr)
    assert len(exp) > 50 and mismatch == 0
    assert_array_equal(got, exp)
    assert_array_equal(_jumps_via_abi(events, 8, n_frames, mr), exp)
    if n_shards == 8:
        assert rounds > 1, 'the sparse atoms should have needed a second resolve round'


def test_jump_machine_frame_shards_flag_inconsistent_stream():
    """Rows whose start state is not the previous destination cannot come from K2; if such a
    table is sharded the published records may be wrong -- the final pass must either flag it
    or still be right."""
    from oracle import jumps as ojumps

    rng = np.random.default_rng(1234)
    flagged = wrong_unflagged = 0
    for trial in range(6):
        r, t_end = _random_event_stream(rng, 3000, 4, 0.5, consistent=False)
        events = np.array(r, dtype=np.int64)
        got, _, mismatch = _jumps_frame_shards_via_abi(events, 1, t_end, 2, [t_end // 3, 2 * t_end // 3])
        exp = ojumps.transitions_to_jumps(events, t_end, 2)
        if mismatch:
            flagged += 1
        elif not np.array_equal(got, exp):
            wrong_unflagged += 1
    assert wrong_unflagged == 0


@pytest.mark.parametrize('n_frames,n_atoms', [(300, 33), (400, 65), (1000, 31), (600, 2), (260, 97), (40, 5)])
def test_k1_strip_walker_shapes(n_frames, n_atoms):
    """Odd atom counts take the non-TMA variant (rows only 8-byte aligned), partial strips and
    several frame chunks; states, fused event counts and events vs the oracles."""
    from gemdat_b200 import Sites, Trajectory
    from oracle import cstates, events as oevents

    rng = np.random.default_rng(n_frames * 131 + n_atoms)
    lat = np.diag([6.0, 7.5, 5.25])
    # 14 random sites at least 1.3 A apart (minimum image): the spheres still overlap (tie rule
    # Q1) but pass the reference's check at fraction_of_overlap=1.0 (2 * min_dist >= r_i + r_j)
    site_frac = np.zeros((0, 3))
    while len(site_frac) < 14:
        cand = rng.random(3)
        d = site_frac - cand
        d -= np.round(d)
        if len(site_frac) == 0 or np.sqrt((((d @ lat)) ** 2).sum(axis=1)).min() >= 1.3:
            site_frac = np.vstack([site_frac, cand])
    labels = ['p', 'q'] * 7
    # slow random walks so that states persist for a while and change now and then
    pos = np.mod(rng.random((1, n_atoms, 3)) + np.cumsum(rng.normal(scale=0.04, size=(n_frames, n_atoms, 3)), axis=0), 1)
    traj = Trajectory(species=['Li'] * n_atoms, coords=pos, lattice=lat, time_step=1e-15)
    sites = Sites(lat, 'Li', site_frac, labels=labels)
    radius, inner = {'p': 1.3, 'q': 1.1}, {'p': 0.5, 'q': 0.8}
    kw = dict(site_radius=radius, site_inner_fraction=inner, fraction_of_overlap=1.0)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eo, ei = cstates.calculate_atom_states_pair(
            site_frac=site_frac, site_labels=labels, positions=pos, lattice_matrix=lat,
            site_radius=radius, site_inner_fraction=inner)
        try:
            exp_events = oevents.calculate_transition_events(eo, ei)
        except (IndexError, ValueError) as exc:  # reference quirks Q4: same exception expected
            with pytest.raises(type(exc)):
                traj.transitions_between_sites(sites, 'Li', **kw)
            return
        tr = traj.transitions_between_sites(sites, 'Li', **kw)
    assert_array_equal(tr.states, eo)
    assert_array_equal(tr.inner_states, ei)
    assert_array_equal(tr.events.to_numpy(), exp_events)
    # the stand-alone count pass (states given as arrays) agrees with the count fused into K1
    from gemdat_b200.transitions import _calculate_transition_events

    ev2 = _calculate_transition_events(atom_sites=eo, atom_inner_sites=ei)
    assert_array_equal(ev2.to_numpy(), tr.events.to_numpy())


@pytest.mark.parametrize('n_frames,n_atoms,scale,cap,route', [
    (3000, 48, 0.004, 32, 'staged'), (1000, 33, 0.004, 32, 'staged'), (4100, 96, 0.003, 32, 'staged'),
    (700, 64, 0.05, 0, 'reread'),   # events nearly every frame: default staging (8 rows / 32 frames) overflows
])
def test_hotpath_staged_events_equal_reread_and_oracle(monkeypatch, n_frames, n_atoms, scale, cap, route):
    """K1 stages the event rows while the states are in registers and K2 only compacts them
    (gd_site_states_staged + gd_events_compact).  Same table as the re-read route
    (gd_events_fill) and as the oracle, incl. frame0 of a shard, odd atom counts (non-TMA K1),
    several chunks and the overflow fallback.  One site per atom, atoms start on their site:
    every site is visited, so the Q2 remap (which also forces the re-read route) stays out."""
    from gemdat_b200 import Sites
    from gemdat_b200.pipeline import HotPath
    from oracle import cstates, events as oevents

    rng = np.random.default_rng(n_frames + n_atoms)
    lat = np.diag([16.0, 17.5, 15.25])
    site_frac = np.zeros((0, 3))
    while len(site_frac) < n_atoms:
        cand = rng.random(3)
        d = site_frac - cand
        d -= np.round(d)
        if len(site_frac) == 0 or np.sqrt((((d @ lat)) ** 2).sum(axis=1)).min() >= 2.0:
            site_frac = np.vstack([site_frac, cand])
    walk = np.cumsum(rng.normal(scale=scale, size=(n_frames, n_atoms, 3)), axis=0)
    pos = np.mod(site_frac[None] + walk - walk[:1], 1)
    labels = ['p'] * n_atoms
    sites = Sites(lat, 'Li', site_frac, labels=labels)
    frame0 = 12345
    d_pos = torch.from_numpy(pos).cuda()
    if cap:
        monkeypatch.setenv('GD_EVENTS_STAGE_CAP', str(cap))
    tables = {}
    for staged in ('1', '0'):
        monkeypatch.setenv('GD_EVENTS_STAGED', staged)
        hp = HotPath(sites=sites, lattice=lat, n_atoms=n_atoms, site_radius=0.9, site_inner_fraction=0.6,
                     minimal_residence=0, n_frames_total=frame0 + n_frames, frame0=frame0)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            res = hp.run(d_pos)
        assert hp.last_events_route == (route if staged == '1' else 'reread')
        tables[staged] = res['events'].cpu().numpy()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eo, ei = cstates.calculate_atom_states_pair(
            site_frac=site_frac, site_labels=labels, positions=pos, lattice_matrix=lat,
            site_radius={'p': 0.9}, site_inner_fraction={'p': 0.6})
    exp = oevents.calculate_transition_events(eo, ei)
    exp[:, 5] += frame0
    assert len(exp) > 200
    assert_array_equal(tables['0'], exp)
    assert_array_equal(tables['1'], exp)


def test_hotpath_optimistic_pass_equals_checked_pass(monkeypatch):
    """With table sizes remembered from an earlier run HotPath queues the whole pass without
    reading anything back (one synchronisation at the end).  Same tables as the checked pass;
    a table that outgrew its remembered size, or jumps that do, send it back to the checked pass."""
    from gemdat_b200 import synth
    from gemdat_b200.pipeline import HotPath

    monkeypatch.setenv('GD_EVENTS_STAGE_CAP', '32')  # chunks are 32 frames here: staging never overflows
    rec = synth.config_c3(n_frames=6000)  # (departures are frequent: the time-parallel chain proves itself)
    pos = rec.device_positions(0, 6000)
    hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=rec.n_atoms, site_radius=0.6,
                 site_inner_fraction=0.5, minimal_residence=3, n_frames_total=6000)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        # reference tables from the checked pass (every count read back where it is needed)
        monkeypatch.setenv('GD_FIRST_CALL_BOUND', '0')
        hp0 = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=rec.n_atoms, site_radius=0.6,
                      site_inner_fraction=0.5, minimal_residence=3, n_frames_total=6000)
        checked = hp0.run(pos)
        assert hp0.last_pass == 'checked'
        ref = {k: checked[k].cpu().numpy().copy() for k in ('outer', 'inner', 'events', 'jumps')}
        assert len(ref['events']) > 500 and len(ref['jumps']) > 5
        monkeypatch.delenv('GD_FIRST_CALL_BOUND')
        # the very first call of a fresh object already takes the one-synchronisation pass, with
        # upper bounds for the table sizes
        first = hp.run(pos)
        assert hp.last_pass == 'optimistic' and hp.last_launch_mode == 'eager', hp.last_optimistic_reject
        for k, v in ref.items():
            assert_array_equal(first[k].cpu().numpy(), v)
        second = hp.run(pos)
        assert hp.last_pass == 'optimistic', hp.last_optimistic_reject
        for k, v in ref.items():
            assert_array_equal(second[k].cpu().numpy(), v)
        assert second['n_events'] == len(ref['events']) and second['n_jumps'] == len(ref['jumps'])
        assert hp.last_launch_mode == 'eager'
        for _ in range(3):  # same buffers and sizes again: captured as one CUDA graph, then replayed
            again = hp.run(pos)
            assert hp.last_pass == 'optimistic' and hp.last_launch_mode == 'graph'
            for k, v in ref.items():
                assert_array_equal(again[k].cpu().numpy(), v)
        hp._cap_events = len(ref['events']) - 7     # the event table no longer fits
        third = hp.run(pos)
        assert hp.last_pass == 'checked' and 'status 8' in hp.last_optimistic_reject
        assert_array_equal(third['events'].cpu().numpy(), ref['events'])
        assert_array_equal(third['jumps'].cpu().numpy(), ref['jumps'])
        hp._jump_cap = len(ref['jumps']) - 1        # the jump table no longer fits
        fourth = hp.run(pos)
        assert hp.last_pass == 'checked' and 'jump table' in hp.last_optimistic_reject
        assert_array_equal(fourth['jumps'].cpu().numpy(), ref['jumps'])
        fifth = hp.run(pos)
        assert hp.last_pass == 'optimistic', hp.last_optimistic_reject
        assert_array_equal(fifth['jumps'].cpu().numpy(), ref['jumps'])
        assert_array_equal(fifth['events'].cpu().numpy(), ref['events'])


@pytest.mark.parametrize('pinned', [False, True])
@pytest.mark.parametrize('n_atoms', [48, 33])
def test_pipelined_host_path_equals_resident_path(monkeypatch, n_atoms, pinned):
    """Host numpy input large enough for the upload / K1 / download pipeline (forced here) gives
    the same states, events and jumps as the single-launch path on a device-resident copy;
    pageable input goes through the threaded page-locked staging buffers."""
    import gemdat_b200.transitions as gt
    from gemdat_b200 import Sites, Trajectory, synth

    rec = synth.config_c2(n_frames=4001, seed=5)
    pos = rec.host_positions(0, 4001)[:, :n_atoms].copy()
    if pinned:
        pos = torch.from_numpy(pos).pin_memory().numpy()
    monkeypatch.setattr(gt, 'COPY_MIN_BYTES', 1 << 14)  # several host threads per part even here
    kw = dict(site_radius=0.9, site_inner_fraction=0.5)
    ref = Trajectory(species=['Li'] * n_atoms, coords=torch.from_numpy(pos).cuda(), lattice=rec.lattice,
                     time_step=2e-15).transitions_between_sites(rec.sites, 'Li', **kw)
    monkeypatch.setattr(gt, 'PIPELINE_MIN_BYTES', 0)
    tr = Trajectory(species=['Li'] * n_atoms, coords=pos, lattice=rec.lattice,
                    time_step=2e-15).transitions_between_sites(rec.sites, 'Li', **kw)
    assert tr.info['host_future'] is not None and ref.info['host_future'] is None
    assert_array_equal(tr.states, ref.states)
    assert_array_equal(tr.inner_states, ref.inner_states)
    assert tr.states.dtype == np.int64
    assert_array_equal(tr.events.to_numpy(), ref.events.to_numpy())
    assert_array_equal(tr.jumps(minimal_residence=3).data.to_numpy(), ref.jumps(minimal_residence=3).data.to_numpy())
    assert tr.info['band_4e-6'] == ref.info['band_4e-6']


# ------------------------------------------------------------------ N3: collective jumps
@pytest.mark.parametrize('case', [0, 1, 2, 3])
def test_collective_vs_reference_golden(case):
    """Collective (collective.py:61-166) against pairs from executing the
    reference (tests/golden/make_golden_collective.py): sort order, emission order of the pairs,
    solo count, site-pair matrix, multiple_collective."""
    import pandas as pd
    from types import SimpleNamespace

    from conftest import load_golden
    from gemdat_b200 import Sites
    from gemdat_b200.collective import Collective
    from gemdat_b200.lattice import Lattice

    g = load_golden('collective')
    t = f'case{case}'
    src = load_golden(str(g[f'{t}_fixture']))
    data = pd.DataFrame(src[str(g[f'{t}_jumps_key'])], columns=[str(x) for x in src['jump_columns']])
    labels = [str(x) for x in src['site_labels']]
    sites = Sites(src['lattice'], 'Li', src['site_frac'], labels=labels)
    coll = Collective(jumps=SimpleNamespace(data=data), sites=sites, lattice=Lattice(src['lattice']),
                      max_steps=int(g[f'{t}_max_steps']), max_dist=float(g[f'{t}_max_dist']))
    assert_array_equal(coll.events.to_numpy(), g[f'{t}_sorted_jumps'])
    assert_array_equal(coll.pairs, g[f'{t}_pairs'])
    assert coll.n_solo_jumps == int(g[f'{t}_n_solo_jumps'])
    assert coll.n_coll_jumps == int(g[f'{t}_n_coll_jumps'])
    assert_array_equal(np.array(coll.coll_jumps, dtype=np.int64).reshape(-1, 2, 2), g[f'{t}_coll_jumps'])
    if case < 3:  # one pandas row per pair, like the reference's `collective` list
        rows = coll.collective
        assert len(rows) == len(g[f'{t}_pairs'])
        for (ei, ej), (i, j) in zip(rows, g[f'{t}_pairs']):
            assert_array_equal(ei.to_numpy(), g[f'{t}_sorted_jumps'][i])
            assert_array_equal(ej.to_numpy(), g[f'{t}_sorted_jumps'][j])
    # label pairs come from a set: compare through the stored order
    mine = coll.site_pair_count_matrix_labels()
    ref_pairs = [tuple(p) for p in g[f'{t}_site_pair_labels'].tolist()]
    assert sorted(mine) == sorted(ref_pairs)
    perm = [mine.index(p) for p in ref_pairs]
    assert_array_equal(coll.site_pair_count_matrix()[np.ix_(perm, perm)], g[f'{t}_site_pair_count_matrix'])
    mj, mc = coll.multiple_collective()
    assert_array_equal(np.stack([mj['start'], mj['stop']], axis=-1), g[f'{t}_multiple_jumps'])
    assert_array_equal(mc, g[f'{t}_multiple_counts'])


@pytest.mark.parametrize('n_jumps,max_steps,max_dist', [(1, 5, 2.0), (2, 5, 50.0), (700, 8, 2.5),
                                                        (3000, 40, 1.8), (3000, 0, 3.0)])
def test_collective_random_tables_vs_oracle(n_jumps, max_steps, max_dist):
    """Random jump tables with many equal (stop, start) keys and long jumps, so that the stable
    sort and the early `break` on a non-monotonic key (collective.py:76-77) both matter."""
    from gemdat_b200 import _device as dv
    from gemdat_b200.collective import _close_matrix, collective_pairs_device
    from oracle import collective as oc

    rng = np.random.default_rng(n_jumps * 7 + max_steps)
    n_sites, n_atoms = 23, 9
    lat = np.array([[7.0, 0.0, 0.0], [1.0, 6.5, 0.0], [0.5, -0.8, 8.0]])
    site_frac = rng.random((n_sites, 3))
    start = rng.integers(0, max(4, n_jumps // 3), n_jumps)
    dur = np.where(rng.random(n_jumps) < 0.15, rng.integers(1, 60, n_jumps), rng.integers(1, 4, n_jumps))
    jumps = np.stack([rng.integers(0, n_atoms, n_jumps), rng.integers(0, n_sites, n_jumps),
                      rng.integers(0, n_sites, n_jumps), start, start + dur,
                      rng.integers(0, 9, n_jumps)], axis=1).astype(np.int64)
    ev, pairs = oc.collective_pairs(jumps, site_frac, lat, max_steps, max_dist)
    close = _close_matrix(lat, site_frac, max_dist)
    d_ev, d_pairs, n_solo = collective_pairs_device(dv.to_device(jumps.astype(np.int32)), close, max_steps)
    assert_array_equal(dv.to_host(d_ev), ev)
    assert_array_equal(dv.to_host(d_pairs).reshape(-1, 2), pairs)
    assert n_solo == oc.solo_count(n_jumps, pairs)


def test_jumps_collective_through_the_api():
    """Jumps.collective (jumps.py:223-251) end to end: attempt frequency -> max_steps -> pairs,
    against the oracle run on the same jump table."""
    from math import ceil

    from oracle import collective as oc
    from oracle import spectrum

    from conftest import load_pipeline_case

    case = load_pipeline_case('pipeline_box_mixed')
    traj, sites = _product(case)
    radius, inner = _radius_args(case)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(sites, 'Li', site_radius=radius, site_inner_fraction=inner,
                                            fraction_of_overlap=float(case['fraction_of_overlap']))
    jumps = tr.jumps()
    coll = jumps.collective(max_dist=2.5)
    from gemdat_b200.metrics import TrajectoryMetrics

    freq, _ = TrajectoryMetrics(jumps.trajectory).attempt_frequency()
    assert coll.max_steps == ceil(1.0 / (float(freq) * jumps.trajectory.time_step))
    ev, pairs = oc.collective_pairs(jumps.data.to_numpy(), case['site_frac'], case['lattice'], coll.max_steps, 2.5)
    assert_array_equal(coll.events.to_numpy(), ev)
    assert_array_equal(coll.pairs, pairs)
    assert jumps.n_solo_jumps == oc.solo_count(len(ev), pairs)
    assert jumps.solo_fraction == jumps.n_solo_jumps / jumps.n_jumps


# ------------------------------------------------------------------ N4: radial distribution
def _rdf_case():
    from conftest import load_golden, load_pipeline_case

    g = load_golden('rdf')
    case = load_pipeline_case(str(g['fixture']))
    traj, sites = _product(case)
    radius, inner = _radius_args(case)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(sites, 'Li', site_radius=radius, site_inner_fraction=inner,
                                            fraction_of_overlap=float(case['fraction_of_overlap']))
    return g, case, traj, tr


@pytest.mark.parametrize('n', [0, 1])
def test_radial_distribution_vs_reference_golden(n, monkeypatch):
    """rdf.py:111-190 against curves from executing the reference
    (tests/golden/make_golden_rdf.py); integer counts, bit-exact, in the reference's key order.
    ``list(set(labels))`` is pinned to the order the golden run saw (it names the '~>' states)."""
    import gemdat_b200.rdf as rdf_mod

    g, case, traj, tr = _rdf_case()
    uniq = [str(x) for x in g['unique_labels']]

    class _PinnedSet(set):
        def __iter__(self):
            pinned = len(self) == len(uniq) and all(u in self for u in uniq)
            return iter(uniq) if pinned else super().__iter__()

    monkeypatch.setattr(rdf_mod, 'set', _PinnedSet, raising=False)
    t = f'rdf{n}'
    res = tr.radial_distribution(floating_specie='Li', max_dist=float(g[f'{t}_max_dist']),
                                 resolution=float(g[f'{t}_resolution']))
    flat = [(state, d) for state, coll in res.items() for d in coll]
    assert [f[0] for f in flat] == [str(s) for s in g[f'{t}_states']]
    assert [f[1].state for f in flat] == [str(s) for s in g[f'{t}_states']]
    assert [f[1].label for f in flat] == [str(s) for s in g[f'{t}_symbols']]
    assert_array_equal(flat[0][1].x, g[f'{t}_x'])
    assert_array_equal(np.stack([f[1].y for f in flat]), g[f'{t}_y'])


@pytest.mark.parametrize('n', [0, 1, 2])
def test_radial_distribution_between_species_vs_reference_golden(n):
    """rdf.py:193-268: integer histogram on the device, fp64 normalisation on the host; the
    reference divides frame by frame and sums, here the sum is divided once: 1e-12 relative."""
    g, case, traj, tr = _rdf_case()
    t = f'between{n}'
    s1, s2 = g[f'{t}_specie_1'].tolist(), g[f'{t}_specie_2'].tolist()
    d = traj.radial_distribution_between_species(specie_1=s1, specie_2=s2, max_dist=float(g[f'{t}_max_dist']),
                                                 resolution=float(g[f'{t}_resolution']))
    assert d.label == str(g[f'{t}_label']) and d.state == ''
    assert_array_equal(d.x, g[f'{t}_x'])
    assert_allclose(d.y, g[f'{t}_y'], rtol=1e-12, atol=0)


@pytest.mark.parametrize('device_input', [False, True])
@pytest.mark.parametrize('name', ['diag', 'sheared', 'skewed'])
def test_species_rdf_with_per_frame_lattices_vs_reference_golden(name, device_input):
    """radial_distribution_between_species on a constant_lattice=False trajectory (rdf.py:232-261):
    gd_rdf_hist_frames counts every frame in its own cell (axis-aligned, sheared, strongly
    skewed), the host normalises each frame with its own volume.  Integer counts exact, curve to
    1e-12 relative against the reference's own output (tests/golden/make_golden_rdf_frames.py)."""
    from conftest import load_golden
    from gemdat_b200 import Trajectory

    g = load_golden('rdf_frames')
    coords = torch.from_numpy(g['positions']).cuda() if device_input else g['positions']
    traj = Trajectory(species=[str(s) for s in g['species']], coords=coords, lattice=g[f'{name}_lattices'],
                      constant_lattice=False, time_step=2e-15)
    for n in (0, 1, 2):
        t = f'{name}_between{n}'
        s1, s2 = g[f'{t}_specie_1'].tolist(), g[f'{t}_specie_2'].tolist()
        d = traj.radial_distribution_between_species(specie_1=s1, specie_2=s2, max_dist=float(g[f'{t}_max_dist']),
                                                     resolution=float(g[f'{t}_resolution']))
        assert d.label == str(g[f'{t}_label']) and d.state == ''
        assert_array_equal(d.x, g[f'{t}_x'])
        assert_allclose(d.y, g[f'{t}_y'], rtol=1e-12, atol=0)
    # the rest of the path asks for THE lattice, which such a trajectory does not have
    with pytest.raises(NotImplementedError):
        traj.get_lattice()
    with pytest.raises(NotImplementedError):
        traj.distances_from_base_position()
    assert traj.get_lattice(3).matrix.shape == (3, 3) and traj[2:10].lattice.shape == (8, 3, 3)


def test_rdf_hist_frames_equals_one_launch_per_frame():
    """gd_rdf_hist_frames through the C ABI: the table of frame t equals gd_rdf_hist run on that
    frame alone with its lattice (bit-exact), with several groups and a skip table."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib

    rng = np.random.default_rng(12)
    T, N = 9, 70
    pos = rng.random((T, N, 3)) * 1.6 - 0.3  # also outside [0, 1)
    lattices = np.stack([np.diag([9.0, 10.0, 11.0]) @ (np.eye(3) + rng.uniform(-0.05, 0.05, (3, 3))) for _ in range(T)])
    lattices[0] = np.diag([9.0, 10.0, 11.0])  # an axis-aligned frame among general ones
    idx1 = np.arange(0, 30, dtype=np.int32)
    idx2 = np.arange(20, N, dtype=np.int32)
    group = (idx2 % 3).astype(np.int32)
    skip = np.zeros((T, len(idx1)), dtype=np.int32)
    skip[rng.random(skip.shape) < 0.2] = -1
    bins = np.arange(0, 4.0 + 0.2, 0.2)
    lib, s = _lib.lib, dv.stream_ptr()
    d_pos, d_idx1, d_idx2, d_group = dv.to_device(pos), dv.to_device(idx1), dv.to_device(idx2), dv.to_device(group)
    d_skip, d_edges = dv.to_device(skip), dv.to_device(bins)
    flat = np.ascontiguousarray(lattices.reshape(T, 9))
    for mode in (0, 1):
        work = dv.empty((int(lib.gd_rdf_frames_work_bytes(T)),), torch.uint8)
        got = dv.zeros((T, 3, len(bins) + 1), torch.int64)
        _lib.check(lib.gd_rdf_hist_frames(dv.ptr(d_pos), T, N, flat.ctypes.data, dv.ptr(work), dv.ptr(d_idx1),
                                          len(idx1), dv.ptr(d_skip), dv.ptr(d_idx2), dv.ptr(d_group), len(idx2), 3,
                                          dv.ptr(d_edges), len(bins), 0.2, mode, 1, dv.ptr(got), s))
        for t in range(T):
            want = dv.zeros((1, 3, len(bins) + 1), torch.int64)
            cls = dv.to_device(skip[t:t + 1].copy())
            _lib.check(lib.gd_rdf_hist(dv.ptr(d_pos[t:t + 1]), 1, N, flat[t].ctypes.data, dv.ptr(d_idx1), len(idx1),
                                       dv.ptr(cls), 1, dv.ptr(d_idx2), dv.ptr(d_group), len(idx2), 3,
                                       dv.ptr(d_edges), len(bins), 0.2, mode, 1, dv.ptr(want), s))
            assert_array_equal(got[t].cpu().numpy(), want[0].cpu().numpy())
        n_pairs = int((skip >= 0).sum()) * len(idx2) - int(sum((skip[t] >= 0)[20:30].sum() for t in range(T)))
        assert int(got.sum().item()) == n_pairs


@pytest.mark.parametrize('triclinic', [False, True])
def test_rdf_hist_random_vs_oracle(triclinic):
    """gd_rdf_hist on a random mixed-species trajectory (general and diagonal cell, several target
    tiles' worth of atoms, coordinates outside [0,1)) against oracle/rdf.py."""
    from gemdat_b200 import Trajectory
    from oracle import rdf as orf

    rng = np.random.default_rng(11 + triclinic)
    lat = np.array([[9.0, 0.0, 0.0], [1.5, 8.0, 0.0], [-0.7, 0.9, 10.0]]) if triclinic else np.diag([9.0, 8.0, 10.0])
    species = ['Li'] * 37 + ['O'] * 150 + ['La'] * 20
    pos = rng.random((12, len(species), 3)) * 1.4 - 0.2
    traj = Trajectory(species=species, coords=pos, lattice=lat, time_step=1e-15)
    for s1, s2, md, res in [('Li', 'O', 5.0, 0.1), ('O', 'O', 4.0, 0.25), (['Li', 'La'], 'O', 6.0, 0.3)]:
        d = traj.radial_distribution_between_species(specie_1=s1, specie_2=s2, max_dist=md, resolution=res)
        w1, w2 = ([s1] if isinstance(s1, str) else s1), ([s2] if isinstance(s2, str) else s2)
        sel = lambda want: pos[:, [s in want for s in species]]  # noqa: E731
        x, y = orf.between_species(coords_1=sel(w1), coords_2=sel(w2), same=(s1 == s2), lattice=lat, max_dist=md,
                                   resolution=res)
        assert_array_equal(d.x, x)
        assert_allclose(d.y, y, rtol=1e-12, atol=0)


# ------------------------------------------------------------------ N1: occupancy, residence, split

N1_CASES = ['pipeline_c2_r09_inner05', 'pipeline_c2_dict_radius', 'pipeline_box_mixed']


def _n1_transitions(name):
    from conftest import load_pipeline_case

    c = load_pipeline_case(name)
    traj, sites = _product(c)
    r, f = _radius_args(c)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(sites, 'Li', site_radius=r, site_inner_fraction=f,
                                            fraction_of_overlap=float(c['fraction_of_overlap']))
    return c, tr


@pytest.mark.parametrize('name', N1_CASES)
def test_n1_vs_reference_golden(name):
    """occupancy / occupancy_by_site_type / atom_locations / residence_time / split / Jumps.split
    against the outputs of the reference's transitions.py:280-410, 837-879 and jumps.py:412-433
    (tests/golden/make_golden_n1.py)."""
    from conftest import load_golden
    from gemdat_b200 import Jumps

    g = load_golden('n1')
    c, tr = _n1_transitions(name)
    occ = tr.occupancy()
    assert_array_equal(np.asarray(occ.occupancies), g[f'{name}/occupancy'])
    assert list(occ.labels) == c['labels']
    by_type = tr.occupancy_by_site_type()
    assert list(by_type) == [str(k) for k in g[f'{name}/by_type_keys']]
    assert_array_equal(np.array(list(by_type.values())), g[f'{name}/by_type_values'])
    loc = tr.atom_locations()
    assert list(loc) == [str(k) for k in g[f'{name}/locations_keys']]
    assert_array_equal(np.array(list(loc.values())), g[f'{name}/locations_values'])
    rt = tr.residence_time()
    assert list(rt.columns) == ['atom index', 'site', 'label', 'time']
    assert_array_equal(rt[['atom index', 'site', 'time']].to_numpy(dtype=np.int64).reshape(-1, 3), g[f'{name}/residence'])
    assert [str(x) for x in rt['label']] == [str(x) for x in g[f'{name}/residence_labels']]
    mr = int(g[f'{name}/minimal_residence'])
    for n in g[f'{name}/splits']:
        n = int(n)
        parts = tr.split(n)
        assert len(parts) == n
        for i, part in enumerate(parts):
            tag = f'{name}/split{n}/{i}'
            assert_array_equal(part.states, g[f'{tag}/states'])
            assert_array_equal(part.inner_states, g[f'{tag}/inner_states'])
            assert_array_equal(part.events.to_numpy().reshape(-1, 6), g[f'{tag}/events'])
            assert_array_equal(part.events.index.to_numpy(), g[f'{tag}/events_index'])
            assert len(part.trajectory) == int(g[f'{tag}/n_frames_trajectory'])
            assert_array_equal(np.asarray(part.occupancy().occupancies), g[f'{tag}/occupancy'])
            assert_array_equal(part.residence_time()[['atom index', 'site', 'time']].to_numpy(dtype=np.int64).reshape(-1, 3),
                               g[f'{tag}/residence'])
            if f'{tag}/jumps' in g:
                assert_array_equal(Jumps(part, minimal_residence=mr).data.to_numpy(), g[f'{tag}/jumps'])
            else:
                with pytest.raises(ValueError, match='No jumps found'):
                    Jumps(part, minimal_residence=mr)
        # Jumps.split (jumps.py:412-433) on a part size where every part has jumps
        if all(f'{name}/split{n}/{i}/jumps' in g for i in range(n)):
            js = tr.jumps(minimal_residence=mr).split(n)
            assert [j.n_jumps for j in js] == [len(g[f'{name}/split{n}/{i}/jumps']) for i in range(n)]
    with pytest.raises(ValueError) as exc:
        tr.split(tr.n_events + 1)
    assert str(exc.value) == str(g[f'{name}/split_error'])
    # the private seam on a DataFrame (transitions.py:837-879), with a shuffled-back index
    from gemdat_b200.transitions import _split_transitions_events

    n = int(g[f'{name}/splits'][0])
    for i, df in enumerate(_split_transitions_events(tr.events, tr.n_states, n)):
        assert_array_equal(df.to_numpy().reshape(-1, 6), g[f'{name}/split{n}/{i}/events'])
        assert_array_equal(df.index.to_numpy(), g[f'{name}/split{n}/{i}/events_index'])


@pytest.mark.parametrize('name', N1_CASES)
def test_rates_activation_energies_graph_vs_reference_golden(name):
    """The consumers of N1 / N2: Jumps.rates / activation_energies / to_graph /
    activation_energy_between_sites against the outputs of the reference's jumps.py:253-308,
    340-462 (tests/golden/make_golden_rates.py).  Floating point: 1e-10 relative (the attempt
    frequency comes from cuFFT instead of scipy's periodogram; everything else is exact)."""
    from conftest import load_golden
    from gemdat_b200 import Jumps

    g = load_golden('rates')
    c, tr = _n1_transitions(name)
    jumps = Jumps(tr, minimal_residence=int(g[f'{name}/minimal_residence']))
    freq, freq_std = jumps.trajectory.metrics().attempt_frequency()
    assert_allclose([float(freq), float(freq_std)], g[f'{name}/attempt_frequency'], rtol=1e-9)
    for n in g[f'{name}/part_counts'].tolist():
        if f'{name}/parts{n}/error' in g:  # some part holds no jump: ValueError('No jumps found')
            with pytest.raises(ValueError, match=str(g[f'{name}/parts{n}/error'])):
                jumps.rates(n_parts=n)
            with pytest.raises(ValueError, match=str(g[f'{name}/parts{n}/error'])):
                jumps.activation_energies(n_parts=n)
            continue
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')  # log(0) for pairs without jumps, like the reference
            rates, eact = jumps.rates(n_parts=n), jumps.activation_energies(n_parts=n)
        assert list(rates.columns) == ['rates', 'std'] and list(eact.columns) == ['energy', 'std']
        assert [list(k) for k in rates.index] == g[f'{name}/rates{n}/keys'].tolist()
        assert [list(k) for k in eact.index] == g[f'{name}/eact{n}/keys'].tolist()
        assert_allclose(rates.to_numpy(dtype=float), g[f'{name}/rates{n}/values'], rtol=1e-12)
        assert_allclose(eact.to_numpy(dtype=float), g[f'{name}/eact{n}/values'], rtol=1e-10, equal_nan=True)
    graph = jumps.to_graph()
    assert [d['label'] for _, d in graph.nodes(data=True)] == [str(x) for x in g[f'{name}/graph_nodes']]
    edges = np.array([(a, b, d['e_act']) for a, b, d in graph.edges(data=True)], dtype=float).reshape(-1, 3)
    want = g[f'{name}/graph_edges']
    assert_array_equal(edges[:, :2], want[:, :2])  # same edges in the same order
    assert_allclose(edges[:, 2], want[:, 2], rtol=1e-10)
    lo, hi = g[f'{name}/graph_cut'].tolist()
    cut = jumps.to_graph(min_e_act=lo, max_e_act=hi)
    cut_edges = np.array([(a, b, d['e_act']) for a, b, d in cut.edges(data=True)], dtype=float).reshape(-1, 3)
    # an edge whose energy equals a bound to the last bit may fall on either side at 1e-10
    want_cut = g[f'{name}/graph_cut_edges']
    inner = (want_cut[:, 2] > lo * (1 + 1e-9)) & (want_cut[:, 2] < hi * (1 - 1e-9))
    have = {(int(a), int(b)) for a, b, _ in cut_edges}
    assert {(int(a), int(b)) for a, b, _ in want_cut[inner]} <= have
    assert have <= {(int(a), int(b)) for a, b, _ in want[(want[:, 2] >= lo * (1 - 1e-9)) & (want[:, 2] <= hi * (1 + 1e-9))]}
    a, b, e = g[f'{name}/between'].tolist()
    assert_allclose(jumps.activation_energy_between_sites(int(a), int(b)), e, rtol=1e-10)
    with pytest.raises(IndexError):
        jumps.activation_energy_between_sites(int(a), int(a))


# ------------------------------------------------------------------ drift / centre of mass
def _mod1_close(a, b, atol):
    """positions agree modulo 1 (a value next to 0 may come out next to 1)."""
    d = np.abs(np.asarray(a) - np.asarray(b))
    return np.minimum(d, 1.0 - d).max() <= atol


DRIFT_CASES = {
    # tag -> (kwargs of the call, golden tag).  floating_species: the reference averages over
    # zero atoms there (tests/golden/make_golden_drift.py); the product follows the docstring,
    # i.e. equals the fixed_species call with the complementary species
    'all': ({}, 'all'),
    'fixed_P_S': ({'fixed_species': ('P', 'S')}, 'fixed_P_S'),
    'fixed_Br': ({'fixed_species': 'Br'}, 'fixed_Br'),
    'floating_Li': ({'floating_species': 'Li'}, 'fixed_P_S_Br'),
    'floating_Li_Br': ({'floating_species': ('Li', 'Br')}, 'fixed_P_S'),
}


@pytest.mark.parametrize('device_input', [False, True])
@pytest.mark.parametrize('tag', list(DRIFT_CASES))
def test_drift_correction_vs_reference_golden(tag, device_input):
    """Trajectory.drift / apply_drift_correction (trajectory.py:751-830) against the reference's
    outputs.  drift: 1e-12 relative (a mean over atoms in another order); corrected positions:
    1e-11 absolute modulo 1 (the reference sums d - drift per atom sequentially, the kernels
    subtract the summed drift from K0's cumulative displacement)."""
    from conftest import load_golden
    from gemdat_b200 import Trajectory, TrajectoryMetrics

    g = load_golden('drift')
    kw, gtag = DRIFT_CASES[tag]
    coords = torch.from_numpy(g['coords']).cuda() if device_input else g['coords']
    traj = Trajectory(species=[str(s) for s in g['species']], coords=coords, lattice=g['lattice'],
                      metadata={'temperature': float(g['temperature'])}, time_step=float(g['time_step']))
    drift = traj.drift(**kw)
    assert drift.shape == g[f'drift/{gtag}'].shape
    assert_allclose(drift, g[f'drift/{gtag}'], rtol=1e-12, atol=1e-18)
    corrected = traj.apply_drift_correction(**kw)
    pos = corrected.positions
    assert dv_is_device(pos) == device_input
    pos = pos.cpu().numpy() if device_input else pos
    assert pos.min() >= 0.0 and pos.max() < 1.0
    assert _mod1_close(pos, g[f'corrected/{gtag}/positions'], 1e-11)
    d_li = TrajectoryMetrics(corrected.filter('Li')).tracer_diffusivity(dimensions=3)
    assert_allclose(corrected.displacements, g[f'corrected/{gtag}/displacements'], rtol=0, atol=1e-11)
    assert_allclose(float(d_li), float(g[f'corrected/{gtag}/tracer_diffusivity_Li']), rtol=1e-9)


def dv_is_device(x):
    return isinstance(x, torch.Tensor) and x.is_cuda


def test_center_of_mass_and_scalar_metrics_vs_reference_golden():
    """Trajectory.center_of_mass (trajectory.py:732-749) and TrajectoryMetrics.mol_per_liter /
    tracer_diffusivity_center_of_mass / haven_ratio / tracer_conductivity (metrics.py:64-165)
    against the reference's outputs, 1e-10 relative."""
    from conftest import load_golden
    from gemdat_b200 import Trajectory, TrajectoryMetrics

    g = load_golden('drift')
    species = [str(s) for s in g['species']]
    traj = Trajectory(species=species, coords=g['coords'], lattice=g['lattice'],
                      metadata={'temperature': float(g['temperature'])}, time_step=float(g['time_step']))
    com = traj.center_of_mass()
    assert com.species == ['X'] and com.coords.shape == g['com/coords'].shape
    assert_allclose(com.coords, g['com/coords'], rtol=1e-12, atol=1e-13)
    assert _mod1_close(traj.center_of_mass().positions, g['com/positions'], 1e-12)
    li = traj.filter('Li')
    assert_allclose(li.center_of_mass().coords, g['com_Li/coords'], rtol=1e-12, atol=1e-13)
    m = TrajectoryMetrics(li)
    assert_allclose(float(m.mol_per_liter()), float(g['metrics/mol_per_liter']), rtol=1e-12)
    assert_allclose(float(m.tracer_diffusivity(dimensions=3)), float(g['metrics/tracer_diffusivity']), rtol=1e-10)
    assert_allclose(float(m.tracer_diffusivity_center_of_mass(dimensions=3)),
                    float(g['metrics/tracer_diffusivity_com']), rtol=1e-9)
    assert_allclose(float(m.haven_ratio(dimensions=3)), float(g['metrics/haven_ratio']), rtol=1e-9)
    assert_allclose(float(m.tracer_conductivity(z_ion=1, dimensions=3)), float(g['metrics/tracer_conductivity']),
                    rtol=1e-10)
    assert_allclose(float(m.tracer_conductivity(z_ion=2, dimensions=2)),
                    float(g['metrics/tracer_conductivity_2d']), rtol=1e-10)
    assert m.tracer_conductivity(z_ion=1).unit == 'S m^-1' and m.mol_per_liter().unit == 'mol l^-1'


@pytest.mark.parametrize('n_frames,n_atoms', [(1, 5), (2, 1), (3001, 37), (20000, 300)])
def test_drift_kernels_random_vs_oracle(n_frames, n_atoms):
    """gd_drift / gd_cumsum_frames3 / gd_drift_correct / gd_frame_mean through the C ABI against
    oracle/drift.py on random walks, incl. one frame, one atom, a ragged atom count."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from oracle import drift as odrift

    rng = np.random.default_rng(n_frames + n_atoms)
    steps = rng.normal(scale=0.01, size=(n_frames, n_atoms, 3)) + rng.normal(scale=0.003, size=(n_frames, 1, 3))
    steps[0] = 0
    pos = np.mod(rng.random((n_atoms, 3)) + np.cumsum(steps, axis=0), 1.0)
    sel = np.sort(rng.choice(n_atoms, size=max(1, n_atoms // 2), replace=False)).astype(np.int32)
    lib, s = _lib.lib, dv.stream_ptr()
    # (every device buffer keeps a name until the end: a temporary would be freed, and its block
    # handed to the next allocation, before the kernel that reads it runs)
    d_pos = dv.to_device(pos)
    d_sel = dv.to_device(sel)
    d_base = dv.to_device(pos[0].copy())
    d_drift = dv.empty((n_frames, 3), torch.float64)
    _lib.check(lib.gd_drift(dv.ptr(d_pos), n_frames, n_atoms, dv.ptr(d_sel), len(sel), None, dv.ptr(d_drift), s))
    want = odrift.drift(pos, sel)[:, 0]
    assert_allclose(d_drift.cpu().numpy(), want, rtol=1e-12, atol=1e-18)
    d_sum = dv.empty((n_frames, 3), torch.float64)
    d_scan = dv.empty((int(lib.gd_cumsum_frames3_work_bytes(n_frames)),), torch.uint8)
    _lib.check(lib.gd_cumsum_frames3(dv.ptr(d_drift), n_frames, dv.ptr(d_scan), dv.ptr(d_sum), s))
    assert_allclose(d_sum.cpu().numpy(), np.cumsum(d_drift.cpu().numpy(), axis=0), rtol=1e-12, atol=1e-15)
    # a frame shard: the row before the shard supplies frame 0's step
    if n_frames > 10:
        cut = n_frames // 3
        part = dv.empty((n_frames - cut, 3), torch.float64)
        _lib.check(lib.gd_drift(dv.ptr(d_pos[cut:]), n_frames - cut, n_atoms, dv.ptr(d_sel), len(sel),
                                dv.ptr(d_pos[cut - 1]), dv.ptr(part), s))
        assert_array_equal(part.cpu().numpy(), d_drift[cut:].cpu().numpy())
    # corrected positions from the cumulative displacements
    cum = np.cumsum(odrift.displacements(pos), axis=0)
    d_cum = dv.to_device(cum)
    d_out = dv.empty((n_frames, n_atoms, 3), torch.float64)
    _lib.check(lib.gd_drift_correct(dv.ptr(d_cum), dv.ptr(d_base), dv.ptr(d_sum), n_frames, n_atoms, dv.ptr(d_out), s))
    got = d_out.cpu().numpy()
    assert got.min() >= 0.0 and got.max() < 1.0
    assert _mod1_close(got, odrift.corrected_positions(pos, sel), 1e-11)
    # weighted mean over the atoms of a frame
    w = rng.random(n_atoms) + 0.5
    d_w = dv.to_device(w)
    d_mean = dv.empty((n_frames, 3), torch.float64)
    _lib.check(lib.gd_frame_mean(dv.ptr(d_cum), dv.ptr(d_base), dv.ptr(d_w), n_frames, n_atoms, dv.ptr(d_mean), s))
    assert_allclose(d_mean.cpu().numpy(), odrift.center_of_mass(pos, w)[:, 0], rtol=1e-12, atol=1e-13)
    # in place
    _lib.check(lib.gd_drift_correct(dv.ptr(d_cum), dv.ptr(d_base), dv.ptr(d_sum), n_frames, n_atoms, dv.ptr(d_cum), s))
    assert_array_equal(d_cum.cpu().numpy(), got)
    _lib.check(lib.gd_frame_mean(dv.ptr(d_pos), None, None, n_frames, n_atoms, dv.ptr(d_mean), s))
    assert_allclose(d_mean.cpu().numpy(), pos.mean(axis=1), rtol=1e-12, atol=1e-15)


@pytest.mark.parametrize('n_frames,n_atoms,n_sites', [(1, 1, 3), (257, 5, 7), (4099, 33, 64), (1000, 17, 60_000)])
def test_n1_kernels_random_vs_oracle(n_frames, n_atoms, n_sites):
    """gd_occupancy (shared-memory bins; global atomics beyond 48k bins; unaligned input) and the
    residence rows from re-extracted outer changes against oracle/sitestats.py."""
    import ctypes as C

    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from gemdat_b200.transitions import _events_device
    from oracle import sitestats

    rng = np.random.default_rng(n_frames + n_atoms)
    # piecewise-constant states with NOSITE gaps
    st = rng.integers(-1, min(n_sites, 40), size=(n_frames, n_atoms))
    hold = rng.random((n_frames, n_atoms)) < 0.8
    for t in range(1, n_frames):
        st[t] = np.where(hold[t], st[t - 1], st[t])
    lib, s = _lib.lib, dv.stream_ptr()
    for shift in (0, 1):  # 16-byte aligned and not
        buf = dv.empty((st.size + 1,), torch.int32)
        d = buf[shift: shift + st.size]
        d.copy_(torch.from_numpy(st.astype(np.int32).ravel()))
        counts = dv.zeros((n_sites + 1,), torch.int64)
        _lib.check(lib.gd_occupancy(dv.ptr(d), st.size, n_sites, dv.ptr(counts), s))
        want = np.bincount(st.ravel() + 1, minlength=n_sites + 1)
        assert_array_equal(counts.cpu().numpy(), want)
    assert_array_equal(counts.cpu().numpy()[1:] / n_frames, sitestats.occupancy(st, n_sites))
    d_st = dv.to_device(st.astype(np.int32))
    want_rows = sitestats.residence_time(st)
    try:
        rows = _events_device(d_st, d_st).rows
    except ValueError:
        assert len(want_rows) == 0
        return
    n_rows = C.c_int64(0)
    work = dv.empty((int(lib.gd_residence_work_bytes(len(rows))),), torch.uint8)
    _lib.check(lib.gd_residence_count(dv.ptr(rows), len(rows), dv.ptr(work), C.byref(n_rows), s))
    got = dv.empty((max(n_rows.value, 1), 3), torch.int32)
    _lib.check(lib.gd_residence_fill(dv.ptr(rows), len(rows), dv.ptr(work), dv.ptr(got), s))
    assert_array_equal(got[: n_rows.value].cpu().numpy(), want_rows)


# ------------------------------------------------------------------ a4: amplitudes on the device


@pytest.mark.parametrize('n_frames,n_atoms', [(1, 1), (2, 3), (3, 1), (31, 2), (513, 5), (4100, 37)])
def test_amplitude_kernels_vs_oracle(n_frames, n_atoms):
    """gd_amplitudes_count/fill + gd_mean_std on random speed rows (with exact zeros and long
    constant-sign runs) against the numpy restatement of metrics.py:190-231.  Sums run left to
    right on the device, numpy's np.sum pairs differently beyond 8 elements: 1e-12 absolute on
    values of order 1."""
    import ctypes as C

    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from oracle import displ

    rng = np.random.default_rng(7 * n_frames + n_atoms)
    speed = rng.normal(size=(n_atoms, n_frames))
    speed[rng.random(speed.shape) < 0.1] = 0.0
    if n_frames > 100:
        speed[0, 10:90] = np.abs(speed[0, 10:90])  # one long run
        speed[-1] = 0.0                             # an atom that never moves
    want = displ.amplitudes(speed)
    lib, s = _lib.lib, dv.stream_ptr()
    tm = dv.to_device(np.ascontiguousarray(speed.T))
    work = dv.empty((int(lib.gd_amplitudes_work_bytes(n_frames, n_atoms)),), torch.uint8)
    off = dv.empty((n_atoms + 1,), torch.int64)
    n_amp = C.c_int64(0)
    _lib.check(lib.gd_amplitudes_count(dv.ptr(tm), n_frames, n_atoms, dv.ptr(work), dv.ptr(off), C.byref(n_amp), s))
    assert n_amp.value == len(want)
    amp = dv.empty((n_amp.value,), torch.float64)
    _lib.check(lib.gd_amplitudes_fill(dv.ptr(tm), n_frames, n_atoms, dv.ptr(work), dv.ptr(off), dv.ptr(amp), s))
    assert_allclose(amp.cpu().numpy(), want, rtol=0, atol=1e-12)
    mean, std = C.c_double(0), C.c_double(0)
    mwork = dv.empty((int(lib.gd_mean_std_work_bytes()),), torch.uint8)
    _lib.check(lib.gd_mean_std(dv.ptr(amp), n_amp.value, dv.ptr(mwork), C.byref(mean), C.byref(std), s))
    assert_allclose(mean.value, np.mean(want), rtol=1e-12, atol=1e-14)
    assert_allclose(std.value, np.std(want), rtol=1e-12, atol=1e-14)


def test_default_radius_route_keeps_the_time_axis_on_the_device(monkeypatch):
    """site_radius=None (transitions.py:141-149): vibration amplitude from K0 + the amplitude
    kernels; no [A, T] array may be downloaded on the way."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import synth

    rec = synth.config_c2(n_frames=3000, seed=5)
    traj = rec.trajectory(device=True)
    big = []
    real = dv.to_host

    def spy(t):
        if 3000 in t.shape or t.numel() >= 3000 * 48:  # anything with a frame axis
            big.append(tuple(t.shape))
        return real(t)

    monkeypatch.setattr(dv, 'to_host', spy)
    import gemdat_b200.metrics as gm
    import gemdat_b200.trajectory as gt

    monkeypatch.setattr(gm.dv, 'to_host', spy)
    monkeypatch.setattr(gt.dv, 'to_host', spy)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=None)
    assert big == [], big
    from oracle import displ
    from oracle import radius as oradius

    vib = displ.vibration_amplitude(rec.host_positions(), rec.lattice)
    want = oradius.radius_from_vibration_amplitude(rec.sites.frac_coords, rec.lattice, vib)
    assert_allclose(tr.site_radius.radius[''], want, rtol=1e-10)


# ------------------------------------------------------------------ np.mod(coords, 1) of to_positions


@pytest.mark.parametrize('route', ['resident', 'pipelined', 'device'])
def test_coordinates_outside_the_unit_cell_are_wrapped(monkeypatch, route):
    """trajectory.py:121-127: positions = np.mod(coords, 1).  Input in [-1, 2) gives the same
    positions, states, events and volume as the wrapped input (and as the oracle on it)."""
    import gemdat_b200.transitions as gtr
    from gemdat_b200 import Trajectory, synth, trajectory_to_volume
    from oracle import states as ostates
    from oracle import volume as ovolume

    rec = synth.config_c2(n_frames=400, seed=9)
    wrapped = rec.host_positions()
    rng = np.random.default_rng(0)
    shifted = wrapped + rng.integers(-1, 2, size=wrapped.shape)
    assert shifted.min() < 0 and shifted.max() >= 1
    want_pos = np.mod(shifted, 1)
    if route == 'pipelined':
        monkeypatch.setattr(gtr, 'PIPELINE_MIN_BYTES', 0)
    coords = torch.from_numpy(shifted).cuda() if route == 'device' else shifted
    traj = Trajectory(species=['Li'] * rec.n_atoms, coords=coords, lattice=rec.lattice, time_step=2e-15)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=0.9, site_inner_fraction=0.5)
        kw = dict(site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=want_pos,
                  lattice_matrix=rec.lattice, site_radius={'': 0.9})
        assert_array_equal(tr.states, ostates.calculate_atom_states(**kw, site_inner_fraction={'': 1.0}))
        assert_array_equal(tr.inner_states, ostates.calculate_atom_states(**kw, site_inner_fraction={'': 0.5}))
    pos = traj.positions
    pos = pos.cpu().numpy() if route == 'device' else pos
    assert_array_equal(pos, want_pos)
    vol = trajectory_to_volume(traj, resolution=0.5)
    assert_array_equal(vol.data, ovolume.trajectory_to_volume(want_pos, rec.lattice, 0.5))


# ------------------------------------------------------------------ minimum image in skewed cells


@pytest.mark.parametrize('cell', ['gamma40', 'non_reduced_supercell', 'argyrodite_2x1x1'])
def test_pbc_distances_in_skewed_cells(cell):
    """gd_pbc_distances (LLL-reduced basis + 27 images, pymatgen pbc_shortest_vectors) against the
    oracle operation by operation (1e-13 relative: the LLL coordinates come from np.dot there)
    and against a brute-force search over 13^3 images of the raw basis."""
    from gemdat_b200 import Lattice
    from test_oracle_golden import SKEWED_CELLS, _brute_force_distances

    from oracle.pmg import Lattice as OLattice

    m = SKEWED_CELLS[cell]
    rng = np.random.default_rng(3)
    f1, f2 = rng.random((25, 3)) * 3 - 1, rng.random((31, 3)) * 3 - 1
    got = Lattice(m).get_all_distances(f1, f2)
    assert_allclose(got, OLattice(m).get_all_distances(f1, f2), rtol=1e-13, atol=1e-13)
    assert_allclose(got, _brute_force_distances(m, f1, f2), rtol=0, atol=1e-12)


# ------------------------------------------------------------------ shard arguments of K0 and the fills


@pytest.mark.parametrize('n_shards', [2, 3, 8])
def test_unwrap_frame_shards_equal_one_shard(n_shards):
    """gd_unwrap_windings -> (all-gather + exclusive sum over shards) -> gd_unwrap_apply with
    d_prev_row / d_base_winding / d_base_pos: every output bit-identical to one call over the whole
    trajectory (windings are integers: the shard scan is exact)."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib

    rng = np.random.default_rng(n_shards)
    T, A = 1003, 29
    lat = np.array([[9.0, 0, 0], [1.0, 10.5, 0], [0.5, -0.7, 12.25]])
    pos_h = np.mod(0.3 + np.cumsum(rng.normal(scale=0.15, size=(T, A, 3)), axis=0), 1)
    pos = dv.to_device(pos_h)
    lib, s = _lib.lib, dv.stream_ptr()

    def outputs(n):
        return (dv.empty((n, A, 3), torch.float64), dv.empty((n, A, 3), torch.float64),
                dv.empty((A, n), torch.float64), dv.empty((n, A), torch.float64))

    work = dv.empty((int(lib.gd_unwrap_chunks(T, A)) * A * 3,), torch.int32)
    whole = outputs(T)
    wind = dv.empty((A, 3), torch.int32)
    _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None, dv.ptr(work),
                             *(dv.ptr(x) for x in whole), dv.ptr(wind), s))
    cuts = np.linspace(0, T, n_shards + 1).astype(int)
    totals, works = [], []
    for r in range(n_shards):
        f0, f1 = cuts[r], cuts[r + 1]
        w = dv.empty((int(lib.gd_unwrap_chunks(f1 - f0, A)) * A * 3,), torch.int32)
        tot = dv.empty((A, 3), torch.int32)
        prev = pos[f0 - 1].contiguous() if r else None
        _lib.check(lib.gd_unwrap_windings(dv.ptr(pos[f0:f1]), f1 - f0, A, dv.ptr(prev), dv.ptr(w), dv.ptr(tot), s))
        totals.append(tot)
        works.append(w)
    base = torch.cumsum(torch.stack(totals), dim=0) - torch.stack(totals)  # exclusive over shards
    assert torch.equal((base[-1] + totals[-1]).int(), wind)
    for r in range(n_shards):
        f0, f1 = cuts[r], cuts[r + 1]
        out = outputs(f1 - f0)
        prev = pos[f0 - 1].contiguous() if r else None
        _lib.check(lib.gd_unwrap_apply(dv.ptr(pos[f0:f1]), f1 - f0, A, lat.ctypes.data, dv.ptr(prev),
                                       dv.ptr(base[r].int().contiguous()), dv.ptr(pos[0].contiguous()),
                                       dv.ptr(works[r]), *(dv.ptr(x) for x in out), s))
        assert torch.equal(out[0], whole[0][f0:f1]) and torch.equal(out[1], whole[1][f0:f1])
        assert torch.equal(out[2], whole[2][:, f0:f1]) and torch.equal(out[3], whole[3][f0:f1])
        # the one-call form with the same shard arguments
        out2 = outputs(f1 - f0)
        _lib.check(lib.gd_unwrap(dv.ptr(pos[f0:f1]), f1 - f0, A, lat.ctypes.data, dv.ptr(prev),
                                 dv.ptr(base[r].int().contiguous()), dv.ptr(pos[0].contiguous()),
                                 dv.ptr(works[r]), *(dv.ptr(x) for x in out2), None, s))
        assert all(torch.equal(a, b) for a, b in zip(out, out2))


@pytest.mark.parametrize('n_shards', [2, 3, 8])
@pytest.mark.parametrize('direction', [1, -1])
def test_fill_frame_shards_equal_one_shard(n_shards, direction):
    """ffill / bfill over frame shards: gd_fill_last_valid -> nearest valid entry of the shards
    before (after) -> gd_fill_apply, and gd_fill_states with d_carry_in / d_carry_out: equal to one
    call over all frames and to the oracle (utils.py:59-98)."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from oracle import fill as ofill

    rng = np.random.default_rng(10 * n_shards + direction)
    T, A = 777, 41
    st = rng.integers(0, 9, size=(T, A))
    st[rng.random(st.shape) < 0.85] = -1
    st[:, 0] = -1          # an atom that never has a site
    st[: T // 2, 1] = -1   # the first shards of this atom are empty
    want = (ofill.ffill if direction > 0 else ofill.bfill)(st, axis=0)
    d = dv.to_device(st.astype(np.int32))
    lib, s = _lib.lib, dv.stream_ptr()
    cuts = np.linspace(0, T, n_shards + 1).astype(int)
    last, works = [], []
    for r in range(n_shards):
        f0, f1 = cuts[r], cuts[r + 1]
        w = dv.empty((int(lib.gd_events_work_elems(f1 - f0, A)),), torch.int32)
        lv = dv.empty((A,), torch.int32)
        _lib.check(lib.gd_fill_last_valid(dv.ptr(d[f0:f1]), f1 - f0, A, direction, dv.ptr(w), dv.ptr(lv), s))
        last.append(lv.cpu().numpy())
        works.append(w)
    order = range(n_shards) if direction > 0 else range(n_shards - 1, -1, -1)
    carry = np.full(A, -1, dtype=np.int32)
    chained = None
    for r in order:
        f0, f1 = cuts[r], cuts[r + 1]
        d_carry = dv.to_device(carry.copy())
        out = dv.empty((f1 - f0, A), torch.int32)
        _lib.check(lib.gd_fill_apply(dv.ptr(d[f0:f1]), f1 - f0, A, direction, dv.ptr(d_carry), dv.ptr(works[r]),
                                     dv.ptr(out), s))
        assert_array_equal(out.cpu().numpy(), want[f0:f1])
        # the one-call form, chaining d_carry_out -> d_carry_in
        w2 = dv.empty((int(lib.gd_events_work_elems(f1 - f0, A)),), torch.int32)
        out2, c_out = dv.empty((f1 - f0, A), torch.int32), dv.empty((A,), torch.int32)
        _lib.check(lib.gd_fill_states(dv.ptr(d[f0:f1]), f1 - f0, A, direction, dv.ptr(chained), dv.ptr(c_out),
                                      dv.ptr(w2), dv.ptr(out2), s))
        assert_array_equal(out2.cpu().numpy(), want[f0:f1])
        chained = c_out
        carry = np.where(last[r] != -1, last[r], carry).astype(np.int32)
        assert_array_equal(c_out.cpu().numpy(), carry)


# ------------------------------------------------------------------ K4 / K5 at the BASELINE sizes


def test_msd_accumulation_at_a_million_frames():
    """K4 at T = 10^6 (config 4's trajectory length): 64 lags spread over [1, T) for 4 atoms against
    the long-double direct sum of oracle/msd.py -- (i) on the device's own unwrapped coordinates
    (isolates the accumulation of k_msd: 1e-12 relative), (ii) end to end from the wrapped positions
    against the oracle's sequential np.cumsum unwrap (1e-10 relative, north_star's tolerance)."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib, synth
    from oracle import msd as omsd

    T = 1_000_000
    rec = synth.make_recipe(synth.garnet_like_sites((3, 3, 2)), 4, T, seed=77, epoch_len=1024, p_hop=0.3)
    pos = rec.device_positions(0, T)
    A = rec.n_atoms
    lib, s = _lib.lib, dv.stream_ptr()
    lat = np.ascontiguousarray(rec.lattice, dtype=np.float64)
    work = dv.empty((int(lib.gd_unwrap_chunks(T, A)) * A * 3,), torch.int32)
    cart = dv.empty((T, A, 3), torch.float64)
    wind = dv.empty((A, 3), torch.int32)
    _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None, dv.ptr(work), None,
                             dv.ptr(cart), None, None, dv.ptr(wind), s))
    soa = dv.empty((A, 3, T), torch.float64)
    _lib.check(lib.gd_transpose_soa(dv.ptr(cart), T, A, dv.ptr(soa), s))
    lags = np.unique(np.concatenate([[1, 2, 3, 7, 2047, 2048, 2049, T - 2, T - 1],
                                     np.geomspace(10, T - 10, 55).astype(np.int64)]))
    got = np.zeros((A, len(lags)))
    one = dv.empty((A, 1), torch.float64)
    for j, k in enumerate(lags):
        _lib.check(lib.gd_msd_direct(dv.ptr(soa), T, A, int(k), 1, dv.ptr(one), s))
        got[:, j] = one.cpu().numpy()[:, 0]
    r = cart.cpu().numpy().transpose(1, 0, 2)  # [A, T, 3], the device's own unwrap
    own = np.zeros_like(got)
    for j, k in enumerate(lags):
        diff = (r[:, k:, :] - r[:, : T - k, :]).astype(np.longdouble)
        own[:, j] = (np.sum(diff * diff, axis=(1, 2), dtype=np.longdouble) / (T - k)).astype(np.float64)
    assert_allclose(got, own, rtol=1e-12, atol=0)
    want = omsd.msd_direct(rec.host_positions(), rec.lattice, lags=lags)
    assert_allclose(got, want, rtol=1e-10, atol=0)


def test_density_at_a_tenth_of_an_angstrom_on_a_c3_window():
    """K5 at BASELINE config 5's resolution (0.1 A, grid 389 x 389 x 259 on the C3 cell) on a
    2000-frame window of the C3 trajectory: bit-exact against np.digitize + np.unique
    (volume.py:471-503)."""
    from gemdat_b200 import synth, trajectory_to_volume
    from oracle import volume as ovolume

    rec = synth.config_c3(n_frames=1_000_000)
    frame0, n = 654_321, 2000
    traj = rec.trajectory(device=True, frame0=frame0, n_frames=n)
    vol = trajectory_to_volume(traj, resolution=0.1)
    want = ovolume.trajectory_to_volume(rec.host_positions(frame0, n), rec.lattice, 0.1)
    assert vol.data.shape == want.shape == (389, 389, 259)
    assert int(vol.data.sum()) == n * rec.n_atoms
    assert_array_equal(vol.data, want)


# ------------------------------------------------------------------ smaller kernels against numpy


@pytest.mark.parametrize('n_frames,n_atoms,cell,res', [
    (64, 5, (4.0, 5.0, 6.0), 0.5),        # 8 x 10 x 12 voxels: CTA-private grid in shared memory
    (700, 33, (9.0, 10.5, 12.25), 0.5),   # partial strip, shared-memory route
    (300, 64, (19.8, 9.9, 9.9), 0.1),     # 1.9e6 voxels: runs go to the global grid
    (1, 1, (3.0, 3.0, 3.0), 0.2),
])
def test_density_routes_agree_and_match_numpy(n_frames, n_atoms, cell, res):
    """gd_density_frames (run-length strip walker; shared-memory grid when it fits) and gd_density
    (flat point list, one RED per point) against np.digitize + np.unique (volume.py:471-503) on
    slowly moving atoms (long runs in one voxel) mixed with jumps across the cell."""
    from gemdat_b200 import _device as dv
    from gemdat_b200.lattice import Lattice
    from gemdat_b200.volume import density_device, grid_edges
    from oracle import volume as ovolume

    rng = np.random.default_rng(n_frames * 7 + n_atoms)
    pos = np.mod(rng.random((1, n_atoms, 3)) + np.cumsum(rng.normal(scale=0.004, size=(n_frames, n_atoms, 3)), axis=0), 1)
    jump = rng.random((n_frames, n_atoms)) < 0.02
    pos[jump] = rng.random((int(jump.sum()), 3))
    pos[0, 0] = [0.0, 0.0, 0.0]
    pos[-1, -1] = [1.0 - 2**-53] * 3
    lat = np.diag(cell)
    want = ovolume.trajectory_to_volume(pos, lat, res)
    nedges = grid_edges(Lattice(lat), res)
    d_pos = dv.to_device(pos)
    by_frames, flag = density_device(d_pos, nedges)
    assert int(flag.item()) == 0
    assert_array_equal(by_frames.cpu().numpy().view(np.uint32).astype(np.int64), want)
    flat, flag = density_device(d_pos.reshape(-1, 3), nedges)
    assert int(flag.item()) == 0
    assert_array_equal(flat.cpu().numpy().view(np.uint32).astype(np.int64), want)


@pytest.mark.parametrize('n,n_atoms', [(2, 1), (5, 3), (64, 2), (777, 5), (4096, 48)])
def test_meanfreq_kernel_vs_periodogram_restatement(n, n_atoms):
    """gd_meanfreq (weights, both spectral sums, ratio; DC bin skipped) on cuFFT's one-sided
    spectrum against the numpy restatement of scipy.signal.periodogram + utils.py:128-160, for
    even and odd lengths, with a constant offset on every row (only the DC bin sees it)."""
    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from oracle import spectrum

    rng = np.random.default_rng(n + n_atoms)
    x = rng.normal(size=(n_atoms, n)) + rng.normal(size=(n_atoms, 1))
    fs = 5e14
    want = spectrum.meanfreq(x, fs)[:, 0]
    spec = torch.fft.rfft(dv.to_device(x), dim=1)
    out = dv.empty((n_atoms,), torch.float64)
    _lib.check(_lib.lib.gd_meanfreq(dv.ptr(spec), n, n_atoms, fs, dv.ptr(out), dv.stream_ptr()))
    assert_allclose(out.cpu().numpy(), want, rtol=1e-10, equal_nan=True)
