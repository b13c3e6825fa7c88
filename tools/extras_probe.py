"""One launch of every secondary kernel of the path on a C3-shaped trajectory (for ncu captures
and quick A/B timing).  usage: python tools/extras_probe.py [frames] [msd_lags]

Prints CUDA-event times per call; run it plain first, then under
`ncu --set full -k regex:k_density|k_unwrap|k_msd|k_fill|k_occupancy|k_residence ...`.
"""
import ctypes as C
import os
import sys

sys.path.insert(0, '.')
import numpy as np
import torch

from gemdat_b200 import _device as dv
from gemdat_b200 import _lib, synth
from gemdat_b200.lattice import Lattice
from gemdat_b200.volume import density_device, grid_edges

T = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
LAGS = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
lib = _lib.lib
rec = synth.config_c3(n_frames=T)
A = rec.n_atoms
pos = rec.device_positions(0, T)
s = dv.stream_ptr()
lat = np.ascontiguousarray(rec.lattice, dtype=np.float64)


REPS = int(os.environ.get('PROBE_REPS', '2'))  # 0 under ncu: one launch per kernel


def timed(name, fn, unit_bytes=None, reps=2):
    reps = min(reps, REPS)
    fn()
    torch.cuda.synchronize()
    best = 1e30
    if reps == 0:
        print(f'{name}: launched once', flush=True)
        return 1.0
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    msg = f'{name}: {best:.3f} ms'
    if unit_bytes:
        msg += f'  {unit_bytes * T * A / best / 1e6:.0f} GB/s algorithmic ({unit_bytes} B/atom-frame)'
    print(msg, flush=True)
    return best


# K0
n_chunks = int(lib.gd_unwrap_chunks(T, A))
work = dv.empty((n_chunks * A * 3,), torch.int32)
cart = dv.empty((T, A, 3), torch.float64)
dist = dv.empty((A, T), torch.float64)
wind = dv.empty((A, 3), torch.int32)
timed('K0 unwrap -> cart', lambda: _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None,
      dv.ptr(work), None, dv.ptr(cart), None, None, dv.ptr(wind), s)), 48)
timed('K0 unwrap -> dist', lambda: _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None,
      dv.ptr(work), None, None, dv.ptr(dist), None, dv.ptr(wind), s)), 32)
del dist
speed = dv.empty((T, A), torch.float64)
timed('K0 unwrap -> speed', lambda: _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None,
      dv.ptr(work), None, None, None, dv.ptr(speed), dv.ptr(wind), s)), 32)
awork = dv.empty((int(lib.gd_amplitudes_work_bytes(T, A)),), torch.uint8)
aoff = dv.empty((A + 1,), torch.int64)
n_amp = C.c_int64(0)
timed('a4 amplitudes count', lambda: _lib.check(lib.gd_amplitudes_count(dv.ptr(speed), T, A, dv.ptr(awork), dv.ptr(aoff),
      C.byref(n_amp), s)), 8)
amp = dv.empty((n_amp.value,), torch.float64)
timed('a4 amplitudes fill', lambda: _lib.check(lib.gd_amplitudes_fill(dv.ptr(speed), T, A, dv.ptr(awork), dv.ptr(aoff),
      dv.ptr(amp), s)), 8 + 8 * n_amp.value / (T * A))
print(f'  amplitudes: {n_amp.value} ({n_amp.value / (T * A):.3f} per atom-frame)', flush=True)
del speed, amp, awork

# drift correction / centre of mass
if hasattr(lib, 'gd_drift'):
    d_drift = dv.empty((T, 3), torch.float64)
    d_sum = dv.empty((T, 3), torch.float64)
    frac = dv.empty((T, A, 3), torch.float64)
    base = pos[0].clone()
    com = dv.empty((T, 3), torch.float64)
    timed('gd_drift (all atoms)', lambda: _lib.check(lib.gd_drift(dv.ptr(pos), T, A, None, A, None, dv.ptr(d_drift), s)), 24)
    d_scan = dv.empty((int(lib.gd_cumsum_frames3_work_bytes(T)),), torch.uint8)
    timed('gd_cumsum_frames3', lambda: _lib.check(lib.gd_cumsum_frames3(dv.ptr(d_drift), T, dv.ptr(d_scan), dv.ptr(d_sum), s)))
    timed('K0 unwrap -> frac', lambda: _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None,
          dv.ptr(work), dv.ptr(frac), None, None, None, dv.ptr(wind), s)), 48)
    timed('gd_frame_mean', lambda: _lib.check(lib.gd_frame_mean(dv.ptr(frac), dv.ptr(base), None, T, A, dv.ptr(com), s)), 24)
    timed('gd_drift_correct (in place)', lambda: _lib.check(lib.gd_drift_correct(dv.ptr(frac), dv.ptr(base), dv.ptr(d_sum),
          T, A, dv.ptr(frac), s)), 48)
    del frac, com

# K5
nedges = grid_edges(Lattice(rec.lattice), 0.1)
grid, flag = density_device(pos, nedges)


def k5():
    grid.zero_()
    density_device(pos, nedges, grid)


timed('K5 density 0.1 A', k5, 28)
assert int(grid.view(torch.int32).to(torch.int64).sum().item()) == T * A
del grid

# K4
soa = dv.empty((A, 3, T), torch.float64)
timed('K4 transpose', lambda: _lib.check(lib.gd_transpose_soa(dv.ptr(cart), T, A, dv.ptr(soa), s)), 48)
del cart
msd = dv.empty((A, LAGS), torch.float64)
ms = timed('K4 msd', lambda: _lib.check(lib.gd_msd_direct(dv.ptr(soa), T, A, 1, LAGS, dv.ptr(msd), s)), reps=1)
pairs = float(sum(T - k for k in range(1, LAGS + 1))) * A
peak = C.c_double(0.0)
_lib.check(lib.gd_fp64_peak(1 << 14, C.byref(peak), s))
print(f'  msd: {pairs / ms / 1e-3:.3e} pair-terms/s, {6 * pairs / (ms * 1e-3) / peak.value:.3f} of the fp64 pipe '
      f'({peak.value:.3e} DFMA/s measured)', flush=True)
del soa, msd

# states for the fills / N1 (K1 without events)
from gemdat_b200.transitions import SiteGrid, _radius_to_dict
from gemdat_b200.lattice import site_arrays

frac, labels, _ = site_arrays(rec.sites)
r, f = _radius_to_dict(0.6, 0.5)
gridS = SiteGrid(lattice_matrix=rec.lattice, site_frac=frac, site_labels=labels, site_radius=r, site_inner_fraction=f)
outer = dv.empty((T, A), torch.int32)
inner = dv.empty((T, A), torch.int32)
timed('K1 states only', lambda: _lib.check(lib.gd_site_states(dv.ptr(pos), T, A, None, dv.ptr(gridS.d_blob),
      gridS.h_blob.ctypes.data, dv.ptr(outer), dv.ptr(inner), None, None, None, None, s)), 32)
fwork = dv.empty((int(lib.gd_events_work_elems(T, A)),), torch.int32)
filled = dv.empty((T, A), torch.int32)
timed('ffill', lambda: _lib.check(lib.gd_fill_states(dv.ptr(outer), T, A, 1, None, None, dv.ptr(fwork),
      dv.ptr(filled), s)), 12)
if hasattr(lib, 'gd_occupancy'):
    occ = dv.zeros((len(frac) + 1,), torch.int64)
    timed('N1 occupancy', lambda: _lib.check(lib.gd_occupancy(dv.ptr(outer), T * A, len(frac), dv.ptr(occ), s)), 4)

# N2: attempt frequency (K0 speed -> cuFFT spectrum -> mean frequency), against the K0 pass that feeds it
from gemdat_b200 import Trajectory, TrajectoryMetrics

del outer, inner, filled
traj = Trajectory(species=['Li'] * A, coords=pos, lattice=rec.lattice, time_step=2e-15)
traj._unit_wrapped = True


def n2():
    TrajectoryMetrics(traj)._meanfreq()


timed('N2 attempt frequency (K0 speed + cuFFT + reductions)', n2, reps=2)
