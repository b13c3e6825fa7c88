#!/usr/bin/env python
"""bench.py -- site-assignment + jump throughput (atom-frames/s) of the hot path.

    python bench.py --gpus N --steps K --warmup W            # our arm
    python bench.py --impl reference --gpus N --steps K ...   # CPU reference arm (oracle port)

A step = one pass of the hot path (K1 site states, K2 events, K3 jumps) over the
whole synthetic trajectory, resident in HBM (``value``), and through the public
Python API with pinned HOST buffers incl. host<->device copies (``e2e``).
Workload at N=1: BASELINE config 3 (LLZO-like, 1008 Li x 10^6 frames, 2160 sites);
at N>1 the frame axis of the same trajectory is sharded over the ranks.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = 'site_assignment_plus_jumps_atom_frames_per_s'
UNIT = 'atom-frames/s'
RADIUS, INNER, MIN_RES = 0.6, 0.5, 3


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--frames', type=int, default=1_000_000, help='total frames of the C3 trajectory')
    ap.add_argument('--e2e-frames', type=int, default=100_000, help='frames per e2e step (host buffers)')
    ap.add_argument('--cpu-frames', type=int, default=2_000, help='frames of the CPU baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true', help='skip the MSD / unwrap / density / RDF / collective-jump section')
    ap.add_argument('--phases', action='store_true',
                    help='after the timed region, run 3 more steps synchronising at phase boundaries '
                         'and report wall ms per phase (diagnostics)')
    ap.add_argument('--msd-lags', type=int, default=2048,
                    help='lags 1..n of the MSD section (whole job); 0 = all lags (BASELINE config 4: ~4 min on one GPU)')
    ap.add_argument('--sustain-s', type=float, default=2.0, help='seconds of back-to-back steps for `sustained` (0 = skip)')
    ap.add_argument('--c2-frames', type=int, default=100_000, help='frames of the C2 section')
    ap.add_argument('--no-c2', action='store_true', help='skip the BASELINE config 2 section (extras.c2)')
    return ap.parse_args()


# ------------------------------------------------------------------------------ clocks


class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region: NVML in-process every
    few ms (the timed region of a 13 ms step is too short for repeated nvidia-smi calls),
    nvidia-smi as the fallback."""

    QUERY = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,'
             'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
             'clocks_event_reasons.sw_power_cap')
    NAMES = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']

    def __init__(self, index: int):
        self.index = index
        self.samples = []  # (sm_mhz, [active reason names])
        self.sm_max = None
        self.source = 'nvidia-smi'
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._nvml = None
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(index).uuid)
            uuid = uuid if uuid.startswith('GPU-') else 'GPU-' + uuid
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._bits = [(getattr(pynvml, 'nvmlClocksThrottleReasonHwSlowdown', 0x8), 'hw_slowdown'),
                          (getattr(pynvml, 'nvmlClocksThrottleReasonHwThermalSlowdown', 0x40), 'hw_thermal_slowdown'),
                          (getattr(pynvml, 'nvmlClocksThrottleReasonSwThermalSlowdown', 0x20), 'sw_thermal_slowdown'),
                          (getattr(pynvml, 'nvmlClocksThrottleReasonSwPowerCap', 0x4), 'sw_power_cap')]
            self._nvml, self._h, self.source = pynvml, h, 'nvml'
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        sm = int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM))
        mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        self.samples.append((sm, [name for bit, name in self._bits if mask & bit]))

    def _sample_smi(self):
        out = subprocess.run(['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.QUERY}',
                              '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
        parts = [p.strip() for p in out.strip().split(',')]
        if len(parts) >= 6 and parts[0].isdigit():
            if parts[1].isdigit():
                self.sm_max = int(parts[1])
            self.samples.append((int(parts[0]),
                                 [n for k, n in enumerate(self.NAMES) if parts[2 + k].lower().startswith('active')]))

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.003 if self._nvml is not None else 0.2)

    def __enter__(self):
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.sm_max, 'reasons': ['unavailable']}
        sm = sorted(s[0] for s in self.samples)
        reasons = [n for n in self.NAMES if any(n in s[1] for s in self.samples)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': self.sm_max, 'reasons': reasons,
                'samples': len(self.samples), 'source': self.source}


# ------------------------------------------------------------------------------ CPU arm


def cpu_reference_pass(rec, frame0, n_frames):
    """The reference's CPU path on one window: restated oracle (scipy cKDTree ball query,
    two calls like transitions.py:160-172), events, jumps.  Returns seconds."""
    import warnings

    from oracle import events as oevents
    from oracle import jumps as ojumps
    from oracle import states as ostates

    pos = rec.host_positions(frame0, n_frames)
    kw = dict(site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=pos,
              lattice_matrix=rec.lattice, site_radius={'': RADIUS})
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        so = ostates.calculate_atom_states(**kw, site_inner_fraction={'': 1.0})
        si = ostates.calculate_atom_states(**kw, site_inner_fraction={'': INNER})
    try:  # (the reference's own error paths for degenerate windows: transitions.py:451,470, jumps.py:107)
        ev = oevents.calculate_transition_events(so, si)
        ojumps.transitions_to_jumps(ev, n_frames, MIN_RES)
    except (ValueError, IndexError):
        pass
    return time.perf_counter() - t0, pos.shape[0] * pos.shape[1]


def _cpu_worker(job):
    """One window of the CPU reference path in a worker process (numpy/scipy single-threaded)."""
    frames_total, frame0, n_frames = job
    os.environ.setdefault('OMP_NUM_THREADS', '1')
    from gemdat_b200 import synth  # recipe only (numpy); no GPU work here

    rec = synth.config_c3(n_frames=frames_total)
    return cpu_reference_pass(rec, frame0, n_frames)


def cpu_reference_parallel(args, n_windows_per_core=1, cores=None):
    """The reference path on ALL host cores: independent frame windows of the same trajectory,
    one process per core (the path has no cross-frame dependency below the event tables, so
    this is how a user would parallelise the reference).  Returns (atom-frames/s, cores, units,
    wall seconds)."""
    import multiprocessing as mp

    cores = cores or os.cpu_count() or 1
    span = max(1, args.frames - args.cpu_frames)
    jobs = [(args.frames, (i * args.cpu_frames) % span, args.cpu_frames)
            for i in range(cores * n_windows_per_core)]
    ctx = mp.get_context('spawn')
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_worker, jobs[:cores])  # warm the workers (imports, page-in)
        t0 = time.perf_counter()
        res = pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
    units = sum(u for _, u in res)
    return units / wall, cores, units, wall


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from gemdat_b200 import synth  # recipe only (numpy); no GPU work in this arm

    rec = synth.config_c3(n_frames=args.frames)
    vals, walls = [], []
    for i in range(args.warmup + args.steps):
        v, cores, units, wall = cpu_reference_parallel(args)
        if i >= args.warmup:
            vals.append(v)
            walls.append(wall)
    value = float(np.mean(vals))
    sample = (f'{cores} windows x {args.cpu_frames} frames x {rec.n_atoms} atoms of C3 per step, one '
              f'process per core; oracle port (numpy + scipy cKDTree, two tree passes like the reference)')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * float(np.mean(walls)),
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic', 'config': workload_config(rec, args),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(rec, args):
    return {
        'workload': f'C3 synthetic LLZO-like garnet {rec.n_atoms} Li x {args.frames} frames, '
                    f'{len(rec.sites)} sites, site_radius={RADIUS}, inner_fraction={INNER}, '
                    f'minimal_residence={MIN_RES}',
        'frame_sharding': f'{args.gpus} rank(s), contiguous frame blocks',
        'l2': 'inputs (24 B/atom-frame) exceed the 126 MB L2 by >100x; no explicit flush',
    }


# ------------------------------------------------------------------------------ MSD / K0 / K5


def run_extras(args, rec, world, rank, peaks, jumps=None):
    """The other kernels of the path on the same C3 trajectory, timed with CUDA events (max over
    ranks).  At N > 1 every stage runs on the rank's FRAME SHARD through gemdat_b200.sharded
    (K0 halo + winding scan, fill carry, K5 / N1 all-reduce, K4 frames -> atoms all-to-all +
    per-lag partial-sum all-reduce); the species RDF (N4, bounded window) and the collective
    pairs (N3) stand beside their numpy ports at N = 1."""
    import ctypes as C

    import torch
    import torch.distributed as dist

    from gemdat_b200 import _device as dv
    from gemdat_b200 import _lib
    from gemdat_b200.lattice import Lattice
    from gemdat_b200.sharded import FrameShards
    from gemdat_b200.volume import grid_edges

    lib = _lib.lib
    s = dv.stream_ptr()
    T, A = args.frames, rec.n_atoms
    out = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, reps=3, warm=True):
        if warm:
            fn()
        best = 1e30
        for _ in range(reps):
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        if world > 1:
            t = torch.tensor([best], device='cuda', dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = float(t.item())
        return best

    def hbm_line(ms, bytes_per_af, **more):
        gbs = bytes_per_af * T * A / (ms * 1e-3) / 1e9
        # N ranks stream their shards at the same time: the roofline is N x one GPU's bandwidth
        return {'ms': ms, 'atom_frames_per_s': T * A / (ms * 1e-3), 'algorithmic_GBps': gbs,
                'frac_of_hbm': gbs / (hbm * world), 'bytes_per_atom_frame': bytes_per_af,
                'frame_shards': world, **more}

    hbm = float(peaks.get('hbm_gbs', 6650.0))
    fs = FrameShards(lattice=rec.lattice, n_atoms=A, n_frames_total=T, world=world, rank=rank)
    f0, f1 = fs.frame_bounds[rank]
    pos = rec.device_positions(f0, f1 - f0)
    n_own = f1 - f0

    # ---- K0: wrapped positions -> cumulative Cartesian displacement.  SURVEY 8(d): 24 B read +
    # 24 B written per atom-frame (the kernel reads the input twice: winding pass + apply pass)
    got = {}

    def k0():
        got['k0'] = fs.unwrap(pos, want_cart=True)

    ms = timed(k0)
    out['unwrap'] = hbm_line(ms, 48, note='algorithmic 24 B in + 24 B out; the winding pass re-reads the input '
                                          '(72 B moved)')
    got.clear()

    # ---- drift correction and centre of mass (callers on the input side): gd_drift (24 B read) ->
    # summed drift -> K0 cumulative displacements (48 B) -> gd_drift_correct in place (48 B), and
    # gd_frame_mean over the same array (24 B).  One shard only (the summed drift is not carried
    # across ranks here).
    if world == 1:
        def drift_correction():
            d = dv.empty((n_own, 3), torch.float64)
            _lib.check(lib.gd_drift(dv.ptr(pos), n_own, A, None, A, None, dv.ptr(d), s))
            summed = dv.empty((n_own, 3), torch.float64)
            scan_work = dv.empty((int(lib.gd_cumsum_frames3_work_bytes(n_own)),), torch.uint8)
            _lib.check(lib.gd_cumsum_frames3(dv.ptr(d), n_own, dv.ptr(scan_work), dv.ptr(summed), s))
            cum = fs.unwrap(pos, want_frac=True)['frac']
            base = pos[0].clone()
            _lib.check(lib.gd_drift_correct(dv.ptr(cum), dv.ptr(base), dv.ptr(summed), n_own, A, dv.ptr(cum), s))
            got['corrected'] = (cum, base)

        ms = timed(drift_correction, reps=2)
        cum, base = got['corrected']
        out['drift_correction'] = hbm_line(ms, 24 + 48 + 48, in_unit_cell=bool(
            float(cum.min().item()) >= 0.0 and float(cum.max().item()) < 1.0),
            note='Trajectory.apply_drift_correction on device-resident positions: drift 24 B + K0 48 B + '
                 'correct-in-place 48 B per atom-frame')
        weights = torch.ones((A,), dtype=torch.float64, device=pos.device)
        com = dv.empty((n_own, 3), torch.float64)

        def frame_mean():
            _lib.check(lib.gd_frame_mean(dv.ptr(cum), dv.ptr(base), dv.ptr(weights), n_own, A, dv.ptr(com), s))

        ms = timed(frame_mean)
        out['center_of_mass'] = hbm_line(ms, 24, note='gd_frame_mean: weighted mean over the atoms of every frame')
        del cum, base, com
        got.clear()

    # ---- a4: speed (K0) -> amplitudes -> std, the site_radius=None route (nothing O(T) downloaded)
    def a4():
        sp = fs.unwrap(pos, want_speed=True)['speed']
        work = dv.empty((int(lib.gd_amplitudes_work_bytes(n_own, A)),), torch.uint8)
        off = dv.empty((A + 1,), torch.int64)
        n_amp = C.c_int64(0)
        _lib.check(lib.gd_amplitudes_count(dv.ptr(sp), n_own, A, dv.ptr(work), dv.ptr(off), C.byref(n_amp), s))
        amp = dv.empty((n_amp.value,), torch.float64)
        _lib.check(lib.gd_amplitudes_fill(dv.ptr(sp), n_own, A, dv.ptr(work), dv.ptr(off), dv.ptr(amp), s))
        mw = dv.empty((int(lib.gd_mean_std_work_bytes()),), torch.uint8)
        mean, std = C.c_double(0), C.c_double(0)
        _lib.check(lib.gd_mean_std(dv.ptr(amp), n_amp.value, dv.ptr(mw), C.byref(mean), C.byref(std), s))
        got['a4'] = (n_amp.value, std.value)

    ms = timed(a4, reps=2)
    out['vibration_amplitude'] = hbm_line(ms, 24 + 8 + 16, amplitudes_rank0=got['a4'][0],
                                          std_rank0_A=got['a4'][1],
                                          note='K0 speed + amplitude count/fill + mean/std; per-rank rows '
                                               '(the amplitude split is not exchanged across shards)')
    got.clear()

    # ---- K5: density histogram at 0.1 A (uint32 grid; one RED per run of frames in one voxel)
    nedges = grid_edges(Lattice(rec.lattice), 0.1)

    def k5():
        got['grid'] = fs.density(pos, nedges)

    ms = timed(k5)
    out['density'] = hbm_line(ms, 28, grid=[int(n) - 1 for n in nedges], resolution_A=0.1,
                              count_ok=bool(int(got['grid'].view(torch.int32).to(torch.int64).sum().item()) == T * A),
                              includes='grid zeroing, kernel, all-reduce of the grid at N > 1')
    got.clear()

    # ---- K1 states for the fills / N1 (not timed here: it is the headline kernel)
    from gemdat_b200.lattice import site_arrays
    from gemdat_b200.transitions import SiteGrid, _radius_to_dict

    frac, labels, _ = site_arrays(rec.sites)
    r_, f_ = _radius_to_dict(RADIUS, INNER)
    sg = SiteGrid(lattice_matrix=rec.lattice, site_frac=frac, site_labels=labels, site_radius=r_,
                  site_inner_fraction=f_)
    outer = dv.empty((n_own, A), torch.int32)
    inner = dv.empty((n_own, A), torch.int32)
    _lib.check(lib.gd_site_states(dv.ptr(pos), n_own, A, None, dv.ptr(sg.d_blob), sg.h_blob.ctypes.data,
                                  dv.ptr(outer), dv.ptr(inner), None, None, None, None, s))
    del inner

    def fill():
        got['prev'] = fs.fill(outer, +1)

    ms = timed(fill)
    out['ffill'] = hbm_line(ms, 12, note='states_prev: 4 B summary pass + 4 B read + 4 B written')
    got.clear()

    def n1():
        got['counts'] = fs.site_counts(outer, len(frac))

    ms = timed(n1)
    out['occupancy'] = hbm_line(ms, 4, count_ok=bool(int(got['counts'].sum().item()) == T * A))
    got.clear()
    del outer

    # ---- K4: direct MSD.  Frame shards are re-cut by atom once (all-to-all), every rank sums all
    # requested lags for its atom block, the atom-averaged curve is the all-reduce of the per-lag
    # partial sums.  --msd-lags 0 = every lag 1 .. T-1 (BASELINE config 4).
    n_lags = (T - 1) if args.msd_lags <= 0 else min(args.msd_lags, T - 1)

    def regroup():
        cart = fs.unwrap(pos, want_cart=True)['cart']
        got['soa'] = fs.regroup_by_atom(cart)

    ms_regroup = timed(regroup, reps=1)
    soa = got.pop('soa')
    a_lo, a_hi = fs.atoms_of_rank()
    msd = dv.empty((a_hi - a_lo, n_lags), torch.float64)

    def k4():
        _lib.check(lib.gd_msd_direct(dv.ptr(soa), T, a_hi - a_lo, 1, n_lags, dv.ptr(msd), s))
        curve = msd.sum(dim=0)
        fs.comm.all_reduce_sum(curve)
        got['curve'] = curve / A

    ms = timed(k4, reps=1 if n_lags > 4096 else 2, warm=n_lags <= 100_000)  # (all lags: ~4 min on one GPU)
    peak = C.c_double(0.0)
    _lib.check(lib.gd_fp64_peak(1 << 14, C.byref(peak), s))
    pairs = (float(n_lags) * T - float(n_lags) * (n_lags + 1) / 2) * A
    out['msd'] = {
        'ms': ms, 'lags': n_lags, 'lags_per_s': n_lags / (ms * 1e-3),
        'pair_terms_per_s': pairs / (ms * 1e-3), 'fp64_inst_per_pair': 6,
        'fp64_peak_dfma_per_s_measured': peak.value * world,
        'frac_of_fp64_pipe': (6.0 * pairs / (ms * 1e-3)) / (peak.value * world),
        'atom_blocks': world, 'regroup_ms': ms_regroup,
        'regroup': 'K0 unwrap + transpose + frames->atoms all-to-all (24 B per atom-frame), once',
        'msd_at_last_lag_A2': float(got['curve'][-1].item()),
        'note': 'unit lag = one lag over all atoms and all frame pairs of the trajectory; the curve is the '
                'all-reduce of per-lag partial sums over the ranks\' atom blocks',
    }
    del soa, msd
    got.clear()

    # ---- N4: species-species distance histogram (rdf.py:193-268) over a bounded frame window
    from gemdat_b200.rdf import _hist

    rdf_own = min(n_own, max(1, 2000 // world))  # a bounded window of every rank's shard
    bins = np.arange(0, 5.0 + 0.1, 0.1)
    atoms = np.arange(A)

    def n4():
        got['h'] = fs.species_rdf_counts(pos[:rdf_own], atoms, atoms, bins, 0.1, True)

    ms = timed(n4, reps=2)
    n_pairs = float(rdf_own) * world * A * (A - 1)
    out['rdf'] = {'ms': ms, 'frames': rdf_own * world, 'pairs_per_s': n_pairs / (ms * 1e-3),
                  'counted': int(got['h'].sum().item()), 'bins': len(bins) - 1, 'max_dist_A': 5.0,
                  'frame_shards': world,
                  'includes': 'index/edge upload, kernel, all-reduce of the [class, group, bin] table',
                  'note': 'axis-aligned cell: one candidate per axis is squared (exact, see gd_pbc.cuh)'}
    if world == 1:
        if not args.no_cpu_baseline:
            from oracle import rdf as ordf  # checker / CPU baseline leg only

            host = pos[:2].cpu().numpy()
            t0 = time.perf_counter()
            _, y = ordf.between_species(coords_1=host, coords_2=host, same=True, lattice=rec.lattice)
            dt = time.perf_counter() - t0
            mine = _hist(pos[:2], rec.lattice, atoms, None, 1, atoms, None, 1, bins, 0.1, 1, True)[0, 0, 1:len(bins)]
            shell = (bins + 0.1) ** 3 - bins**3
            norm = ((A * A / float(abs(np.linalg.det(np.asarray(rec.lattice, dtype=np.float64))))) * (4 / 3) * np.pi * shell)[:-1]
            out['rdf']['cpu_port'] = {'frames': 2, 's': dt, 'pairs_per_s': 2.0 * A * (A - 1) / dt,
                                      'equal': bool(np.allclose(mine / norm / 2, y, rtol=1e-12, atol=0))}

    # ---- N3: collective-jump pair search on the step's own jump table (rank-local, N=1 only)
    if jumps is not None and world == 1 and int(jumps.shape[0]) > 1:
        from gemdat_b200.collective import _close_matrix, collective_pairs_device

        frac = np.ascontiguousarray(rec.sites.frac_coords, dtype=np.float64)
        max_steps, max_dist = 200, 4.0
        close = _close_matrix(rec.lattice, frac, max_dist)
        found = {}

        def n3():
            found['ev'], found['pairs'], found['solo'] = collective_pairs_device(jumps, close, max_steps)

        ms = timed(n3)
        J, P = int(jumps.shape[0]), int(found['pairs'].shape[0])
        out['collective'] = {'ms': ms, 'jumps': J, 'pairs': P, 'n_solo_jumps': int(found['solo']),
                             'max_steps': max_steps, 'max_dist_A': max_dist, 'jumps_per_s': J / (ms * 1e-3),
                             'includes': 'sort + count + scan + fill, two host reads'}
        if not args.no_cpu_baseline:
            from oracle import collective as ocoll  # checker / CPU baseline leg only

            n0 = min(J, 4000)
            head = found['ev'][:n0].cpu().numpy()
            host_close = ocoll.close_matrix(frac, rec.lattice, max_dist)  # untimed on both sides
            t0 = time.perf_counter()
            _, ref_pairs = ocoll.collective_pairs(head, frac, rec.lattice, max_steps, max_dist, host_close)
            dt = time.perf_counter() - t0
            mine = found['pairs'].cpu().numpy()
            mine = mine[(mine[:, 0] < n0) & (mine[:, 1] < n0)]
            # a time prefix of the sorted table: every i < n0 breaks at the same j or at the cut
            out['collective']['cpu_port'] = {'jumps': n0, 's': dt, 'jumps_per_s': n0 / dt,
                                             'pairs_equal': bool(np.array_equal(mine, ref_pairs))}
    return out


def run_c2(args):
    """BASELINE config 2 (synthetic argyrodite 2x1x1: 48 Li, 96 sites, 10^5 frames) through the
    public API on a device-resident trajectory, with `site_radius=0.9` and with the default radius
    (`site_radius=None`: vibration amplitude -> radius on the device); every table is compared
    with the oracle in the same run."""
    import warnings

    import torch

    from gemdat_b200 import synth

    T = args.c2_frames
    rec = synth.config_c2(n_frames=T)
    pos = rec.device_positions(0, T)
    out = {'workload': f'C2 synthetic argyrodite 2x1x1, {rec.n_atoms} Li x {T} frames, {len(rec.sites)} sites'}

    def one(radius, inner, mr):
        from gemdat_b200 import Trajectory

        traj = Trajectory(species=['Li'] * rec.n_atoms, coords=pos, lattice=rec.lattice, time_step=2e-15)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=radius, site_inner_fraction=inner)
            jumps = tr.jumps(minimal_residence=mr)
        return tr, jumps

    for tag, radius, inner, mr in (('r09_inner05', 0.9, 0.5, 3), ('default_radius', None, 1.0, 0)):
        one(radius, inner, mr)
        times = []
        for _ in range(3):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            tr, jumps = one(radius, inner, mr)
            b.record()
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b))
        ms = float(np.median(times))
        line = {'ms': ms, 'atom_frames_per_s': T * rec.n_atoms / (ms * 1e-3), 'n_events': tr.n_events,
                'n_jumps': jumps.n_jumps, 'site_radius': dict(tr.site_radius.radius),
                'what': 'Trajectory.transitions_between_sites + Transitions.jumps, trajectory resident in HBM, '
                        'tables left on the device (CUDA events, median of 3)'}
        if not args.no_cpu_baseline:
            from oracle import cstates, displ
            from oracle import events as oevents
            from oracle import jumps as ojumps
            from oracle import radius as oradius

            host = rec.host_positions(0, T)
            t0 = time.perf_counter()
            r_used = float(tr.site_radius.radius[''])
            checks = {}
            if radius is None:
                # the radius follows from the vibration amplitude (a std over ~10^6 amplitudes: the
                # device reduction and numpy's pairwise sums agree to ~1e-15); the tables are then
                # compared for the radius the product used
                r = oradius.radius_from_vibration_amplitude(rec.sites.frac_coords, rec.lattice,
                                                            displ.vibration_amplitude(host, rec.lattice))
                line['oracle_radius'] = r
                line['radius_rel_diff'] = abs(r_used - r) / r
                checks['radius_1e-10'] = line['radius_rel_diff'] < 1e-10
            with warnings.catch_warnings():
                warnings.simplefilter('ignore')
                so, si = cstates.calculate_atom_states_pair(
                    site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=host,
                    lattice_matrix=rec.lattice, site_radius={'': r_used}, site_inner_fraction={'': inner})
            ev = oevents.calculate_transition_events(so, si)
            oj = ojumps.transitions_to_jumps(ev, T, mr)
            line['oracle_s'] = time.perf_counter() - t0
            checks.update(states=bool(np.array_equal(tr.states, so)),
                          inner_states=bool(np.array_equal(tr.inner_states, si)),
                          events=bool(np.array_equal(tr.events.to_numpy(), ev)),
                          jumps=bool(np.array_equal(jumps.data.to_numpy(), oj)))
            line['parity_checks'] = checks
            line['parity'] = all(checks.values())
            assert line['parity'], f'C2 {tag}: differs from the oracle: {checks}'
        out[tag] = line
    return out


# ------------------------------------------------------------------------------ GPU arm


def pin_to_gpu_numa_node(local_rank: int):
    """One process per GPU: run on the CPU cores NVML lists as local to that GPU, so that the
    pinned host buffers of the e2e leg are first touched on the GPU's own NUMA node (8 ranks
    move ~35 GB per e2e step between host memory and their GPUs).  Best effort; returns the
    number of cores, or None when NVML / the scheduler call is unavailable."""
    try:
        import pynvml
        import torch

        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local_rank).uuid)
        uuid = uuid if uuid.startswith('GPU-') else 'GPU-' + uuid
        h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        allowed = os.sched_getaffinity(0)
        cpus = {64 * w + b for w, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1} & allowed
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None




def run_ours(args):
    import torch
    import torch.distributed as dist

    from gemdat_b200 import _lib, synth
    from gemdat_b200.pipeline import HotPath

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    affinity = pin_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    rec = synth.config_c3(n_frames=args.frames)
    n_atoms = rec.n_atoms
    # contiguous frame shard of this rank
    f0 = args.frames * rank // world
    f1 = args.frames * (rank + 1) // world
    pos = rec.device_positions(f0, f1 - f0)
    # the frame after this shard (every rank reads it from its own input; no halo exchange)
    next_row = rec.device_positions(f1, 1)[0].contiguous() if (world > 1 and rank + 1 < world) else None
    halo = dict(next_pos_row=next_row) if world > 1 else {}
    if world > 1 and next_row is None:
        halo = dict(next_pos_row=pos[0])  # last rank: says "I bring my own halo" (the row is not used)
    hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=n_atoms, site_radius=RADIUS,
                 site_inner_fraction=INNER, minimal_residence=MIN_RES, n_frames_total=args.frames,
                 frame0=f0, world=world, rank=rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        res = hp.run(pos, **halo)
    barrier()
    k1_ms = []
    sampler = ClockSampler(local_rank)
    with sampler:
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            res = hp.run(pos, **halo)
        e1.record()
        barrier()
        launch_mode, pass_route = hp.last_launch_mode, hp.last_pass
        # K1's own duration: CUDA events around the launch.  The timed steps above replay one CUDA
        # graph per step (timing events cannot sit inside a captured graph), so the same K steps run
        # once more launch by launch, still under the clock sampler, for the per-kernel number.
        use_graph, hp.use_graph = hp.use_graph, False
        for _ in range(args.steps):
            hp.run(pos, **halo)
            k1_ms.append(hp.last_k1_events)
        hp.use_graph = use_graph
        barrier()
    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device='cuda', dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    total_ms = max_over_ranks(e0.elapsed_time(e1))

    # ---- what a user analysing ONE trajectory gets: the checked pass of a fresh HotPath (sizes
    # unknown, every count read back where it is needed; no graph), kernels already loaded
    first_ms = []
    for _ in range(3):
        hp1 = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=n_atoms, site_radius=RADIUS,
                      site_inner_fraction=INNER, minimal_residence=MIN_RES, n_frames_total=args.frames,
                      frame0=f0, world=world, rank=rank)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        r1 = hp1.run(pos, **halo)
        b.record()
        barrier()
        first_ms.append(max_over_ranks(a.elapsed_time(b)))
        assert r1['n_jumps'] == res['n_jumps']
        first_pass = hp1.last_pass
        del hp1, r1
    first_call_ms = float(np.median(first_ms))

    # ---- sustained: back-to-back steps for >= 2 s under the clock sampler (power / thermal state
    # of a long job, next to the driver's K steps)
    sustained = None
    if args.sustain_s > 0:
        n_sus = max(args.steps, int(args.sustain_s * 1e3 / max(total_ms / args.steps, 1e-3)) + 1)
        sus_sampler = ClockSampler(local_rank)
        with sus_sampler:
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n_sus):
                hp.run(pos, **halo)
            b.record()
            barrier()
        sus_ms = max_over_ranks(a.elapsed_time(b)) / n_sus
        sustained = {'steps': n_sus, 'ms_per_step': sus_ms, 'value': args.frames * n_atoms / (sus_ms * 1e-3),
                     'seconds': sus_ms * n_sus * 1e-3, 'clocks': sus_sampler.summary()}
    k1_avg_ms = float(np.mean([a.elapsed_time(b) for a, b in k1_ms]))
    ms_per_step = total_ms / args.steps
    n_af_total = args.frames * n_atoms
    value = n_af_total / (ms_per_step * 1e-3)

    line = None
    peaks = {}
    try:
        peaks = json.loads((ROOT / 'MEASURED_PEAKS.json').read_text())
    except Exception:
        pass
    if rank == 0:
        peak = float(peaks.get('hbm_gbs', 6650.0))
        # K1 reads 24 B of positions and writes 2 x 4 B of states per atom-frame; on the staged
        # route it also writes the 24 B event rows (K2 then only compacts them)
        staged = hp.last_events_route == 'staged'
        k1_bytes = 32.0 * (f1 - f0) * n_atoms + (24.0 * hp.last_n_local_events if staged else 0.0)
        achieved = k1_bytes / (k1_avg_ms * 1e-3) / 1e9
        traffic, traffic_src = None, None
        try:  # DRAM bytes per launch from the committed ncu --set full capture of this kernel
            tr = json.loads((ROOT / 'profiles' / 'k1_traffic.json').read_text())
            traffic = float(tr['dram_bytes_per_atom_frame']) * (f1 - f0) * n_atoms
            traffic_src = tr['source']
        except Exception:
            pass
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_per_step, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(rec, args),
            'roofline': {'bound': 'hbm', 'kernel': 'k_site_states', 'achieved': achieved, 'peak': peak,
                         'unit': 'GB/s', 'frac': achieved / peak, 'traffic': traffic,
                         'traffic_unit': 'bytes per launch (dram read + write)', 'traffic_source': traffic_src,
                         'algorithmic_bytes_per_launch': k1_bytes,
                         'peak_source': 'MEASURED_PEAKS.json hbm_gbs (of measured)' if peaks else 'fallback',
                         'algorithmic_bytes_per_atom_frame': k1_bytes / ((f1 - f0) * n_atoms),
                         'events_route': hp.last_events_route, 'pass': pass_route, 'step_launch_mode': launch_mode,
                         'launch_ms': k1_avg_ms,
                         'launch_ms_how': 'CUDA events around the K1 launch, K eager steps right after the timed region'},
            'gpu_launches': hp.launches_per_run * args.steps,
            'clocks': sampler.summary(),
            'results': {'n_events': int(res['n_events']), 'n_jumps': int(res['n_jumps'])},
            'first_call': {'ms': first_call_ms, 'value': n_af_total / (first_call_ms * 1e-3), 'unit': UNIT,
                           'pass': first_pass,
                           'what': 'first call of a fresh HotPath (no remembered table sizes, no CUDA graph): the '
                                   'one-synchronisation pass with upper-bound table sizes, or the checked pass if '
                                   'that was rejected; median of 3, CUDA events, max over ranks'},
            'sustained': sustained,
        }

    if args.phases:
        hp.profile_phases = True
        acc = {}
        for _ in range(3):
            barrier()
            hp.run(pos, **halo)
            for k, v in hp.last_phases.items():
                acc[k] = acc.get(k, 0.0) + v / 3
        hp.profile_phases = False
        if rank == 0:
            line['phases_ms'] = {k: round(v, 3) for k, v in acc.items()}

    extras = None
    if not args.no_extras:
        step_jumps = res['jumps'].clone() if world == 1 else None
        del res
        hp._buf.clear()
        extras = run_extras(args, rec, world, rank, peaks, step_jumps)
        if not args.no_c2 and world == 1:
            extras['c2'] = run_c2(args)

    # ---- e2e through the public API with pinned host buffers (rank-local window)
    from gemdat_b200 import Trajectory

    e2e_frames = min(args.e2e_frames, f1 - f0)
    host = torch.empty((e2e_frames, n_atoms, 3), dtype=torch.float64, pin_memory=True)
    host.copy_(pos[:e2e_frames])
    torch.cuda.synchronize()
    host_np = host.numpy()
    e2e_times, e2e_marks = [], []
    for i in range(args.warmup + args.steps):
        barrier()
        t0 = time.perf_counter()
        traj = Trajectory(species=['Li'] * n_atoms, coords=host_np, lattice=rec.lattice, time_step=2e-15)
        tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=RADIUS, site_inner_fraction=INNER)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        jumps = tr.jumps(minimal_residence=MIN_RES)
        n_j = jumps.n_jumps
        t2 = time.perf_counter()
        # device -> host: int64 arrays and DataFrames like the reference returns them
        states, inner_states = tr.states, tr.inner_states
        t3 = time.perf_counter()
        events_df = tr.events
        torch.cuda.synchronize()
        t4 = time.perf_counter()
        if i >= args.warmup:
            e2e_times.append(t4 - t0)
            e2e_marks.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
    e2e_dt = max_over_ranks(float(np.mean(e2e_times)))
    marks = torch.tensor(1e3 * np.mean(np.array(e2e_marks), axis=0), device='cuda', dtype=torch.float64)
    if world > 1:
        allm = [torch.empty_like(marks) for _ in range(world)]
        dist.all_gather(allm, marks)
        marks_all = torch.stack(allm).cpu().numpy()
    else:
        marks_all = marks.cpu().numpy()[None]

    # ---- the same call from a PAGEABLE numpy array (what `np.load` hands a user), pinned-host
    # pool emptied first: the cold call pays for cudaHostAlloc of the result buffers and for the
    # staged (unpinned) upload
    pageable = np.array(host_np, copy=True)
    try:
        torch._C._host_emptyCache()
        cold_pool = True
    except Exception:
        cold_pool = False
    page_times = []
    for i in range(4):
        barrier()
        t0 = time.perf_counter()
        traj = Trajectory(species=['Li'] * n_atoms, coords=pageable, lattice=rec.lattice, time_step=2e-15)
        tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=RADIUS, site_inner_fraction=INNER)
        jumps = tr.jumps(minimal_residence=MIN_RES)
        _ = (tr.states, tr.inner_states, tr.events, jumps.data)
        torch.cuda.synchronize()
        page_times.append(max_over_ranks(time.perf_counter() - t0))
        del traj, tr, _
    if rank == 0:
        if extras is not None:
            line['extras'] = extras
        line['e2e'] = {
            'value': world * e2e_frames * n_atoms / e2e_dt, 'unit': UNIT,
            'h2d_bytes_per_step': int(host_np.nbytes),
            'd2h_bytes_per_step': int(states.nbytes + inner_states.nbytes + events_df.to_numpy().nbytes
                                      + jumps.data.to_numpy().nbytes),
            'breakdown_ms': dict(zip(['upload+transitions', 'jumps', 'states_d2h', 'events_d2h'],
                                     marks_all[0].round(2).tolist())),
            'breakdown_ms_per_rank': {'columns': ['upload+transitions', 'jumps', 'states_d2h', 'events_d2h'],
                                      'rows': marks_all.round(2).tolist()},
            'pageable': {'first_call_value': world * e2e_frames * n_atoms / page_times[0],
                         'warm_value': world * e2e_frames * n_atoms / page_times[-1],
                         'ms': [round(1e3 * t, 2) for t in page_times], 'pinned_pool_emptied_first': cold_pool,
                         'what': 'plain (pageable) np.ndarray input, staged through two page-locked buffers by host '
                                 'threads; first call with an empty pinned-host pool (the pool fills over '
                                 'the first two calls), warm_value = the last call'},
            'frames_per_step': e2e_frames, 'ms_per_step': 1e3 * e2e_dt, 'cpu_affinity_cores': affinity,
            'api': 'Trajectory.transitions_between_sites + Transitions.jumps (numpy in, numpy/pandas out)',
        }
        if not args.no_cpu_baseline and world == 1:  # rank 0 at N=1 only
            v, cores, units, wall = cpu_reference_parallel(args)
            line['cpu_baseline'] = {
                'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                'sample': f'{cores} windows x {args.cpu_frames} frames x {n_atoms} atoms of C3, one process '
                          f'per core ({wall:.1f} s wall; numpy + scipy cKDTree oracle, two tree passes)',
            }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
