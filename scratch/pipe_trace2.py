import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import gemdat_b200.transitions as gt
from gemdat_b200 import Trajectory, synth
from gemdat_b200.lattice import site_arrays
frames = 100_000
rec = synth.config_c3(n_frames=frames)
pos = rec.device_positions(0, frames)
host = torch.empty(tuple(pos.shape), dtype=torch.float64, pin_memory=True); host.copy_(pos); torch.cuda.synchronize()
host_np = host.numpy(); del pos
frac, labels, _ = site_arrays(rec.sites)
grid = gt.SiteGrid(lattice_matrix=rec.lattice, site_frac=frac, site_labels=labels, site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
for rep in range(4):
    if rep >= 2:
        grid = gt.SiteGrid(lattice_matrix=rec.lattice, site_frac=frac, site_labels=labels, site_radius={'': 0.6}, site_inner_fraction={'': 0.5})
    traj = Trajectory(species=['Li'] * rec.n_atoms, coords=host_np, lattice=rec.lattice, time_step=2e-15)
    trace = []
    torch.cuda.synchronize(); w0 = time.perf_counter()
    t0 = torch.cuda.Event(enable_timing=True); t0.record()
    out = gt._pipelined_states(traj, grid, trace)
    w1 = time.perf_counter()
    torch.cuda.synchronize(); w2 = time.perf_counter()
    for k, (a, c, d) in enumerate(trace[:3]):
        print(f'  part {k}: upload done {t0.elapsed_time(a):6.1f}  K1+widen done {t0.elapsed_time(c):6.1f}  download done {t0.elapsed_time(d):6.1f}')
    print(f'rep {rep}: host returns after {1e3*(w1-w0):.1f} ms, all done after {1e3*(w2-w0):.1f} ms')
for k, (a, c, d) in enumerate(trace):
    print(f'  part {k}: upload done {t0.elapsed_time(a):6.1f}  K1+widen done {t0.elapsed_time(c):6.1f}  download done {t0.elapsed_time(d):6.1f}')
