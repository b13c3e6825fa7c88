"""Warm public-API call from a plain (pageable) numpy array against the number of host threads that
stage the parts into page-locked buffers (transitions.COPY_THREADS)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
import gemdat_b200.transitions as gt
from gemdat_b200 import Trajectory, synth
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
rec = synth.config_c3(n_frames=frames)
host_np = rec.device_positions(0, frames).cpu().numpy().copy()


def step():
    t0 = time.perf_counter()
    traj = Trajectory(species=['Li'] * rec.n_atoms, coords=host_np, lattice=rec.lattice, time_step=2e-15)
    tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=0.6, site_inner_fraction=0.5)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


import os
print('host cores', os.cpu_count(), 'affinity', len(os.sched_getaffinity(0)))
for threads in (8, 4, 12, 16, 24, 32):
    gt.COPY_THREADS = threads
    gt._COPY_POOL = None
    for _ in range(3): step()
    r = [step() for _ in range(5)]
    print(f'COPY_THREADS={threads}: {np.mean(r):.1f} ms (min {min(r):.1f})', flush=True)
