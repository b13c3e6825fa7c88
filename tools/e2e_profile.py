"""e2e step timing with and without the upload/K1/download pipeline."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
import gemdat_b200.transitions as gt
from gemdat_b200 import Trajectory, synth
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
rec = synth.config_c3(n_frames=frames)
pos = rec.device_positions(0, frames)
host = torch.empty(tuple(pos.shape), dtype=torch.float64, pin_memory=True); host.copy_(pos); torch.cuda.synchronize()
host_np = host.numpy(); del pos
import os
sites = rec.sites
if os.environ.get('E2E_SITES'):  # fewer sites: the grid build is fast, the table upload meets part copies in flight
    from gemdat_b200 import Sites
    from gemdat_b200.lattice import site_arrays
    f, l, _ = site_arrays(rec.sites); n = int(os.environ['E2E_SITES'])
    sites = Sites(rec.lattice, 'Li', f[:n], labels=list(l[:n]))
def step():
    t0 = time.perf_counter()
    traj = Trajectory(species=['Li'] * rec.n_atoms, coords=host_np, lattice=rec.lattice, time_step=2e-15)
    tr = traj.transitions_between_sites(sites, 'Li', site_radius=0.6, site_inner_fraction=0.5)
    t1 = time.perf_counter()
    j = tr.jumps(minimal_residence=3); n = j.n_jumps
    s, i = tr.states, tr.inner_states
    t2 = time.perf_counter()
    e = tr.events
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    return np.array([t1 - t0, t2 - t1, t3 - t2, t3 - t0]) * 1e3
for parts, ahead in ((32, 1), (32, 2), (32, 4), (32, 6), (32, 10), (32, 32), (16, 3)):
    gt.PIPELINE_MIN_BYTES = 0
    gt.PIPELINE_PARTS = parts
    gt.PIPELINE_AHEAD = ahead
    for _ in range(3): step()
    rs = np.array([step() for _ in range(6)])
    r = rs.mean(axis=0)
    print(f'parts={parts} ahead={ahead}: transitions {r[0]:.1f} ms (min {rs[:, 0].min():.1f}, max {rs[:, 0].max():.1f}), '
          f'jumps+states {r[1]:.1f} ms, events {r[2]:.1f} ms, total {r[3]:.1f} ms', flush=True)
