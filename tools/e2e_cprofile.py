"""cProfile of one warm public-API call on a C3 window (host numpy in, numpy / pandas out)."""
import cProfile, pstats, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from gemdat_b200 import Trajectory, synth
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
rec = synth.config_c3(n_frames=frames)
pos = rec.device_positions(0, frames)
host = torch.empty(tuple(pos.shape), dtype=torch.float64, pin_memory=True); host.copy_(pos); torch.cuda.synchronize()
host_np = host.numpy(); del pos


def step():
    traj = Trajectory(species=['Li'] * rec.n_atoms, coords=host_np, lattice=rec.lattice, time_step=2e-15)
    tr = traj.transitions_between_sites(rec.sites, 'Li', site_radius=0.6, site_inner_fraction=0.5)
    torch.cuda.synchronize()
    return tr


for _ in range(3):
    step()
t0 = time.perf_counter(); step(); print('one call', (time.perf_counter() - t0) * 1e3, 'ms')
pr = cProfile.Profile(); pr.enable(); step(); pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(28)
