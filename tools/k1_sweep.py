"""Times gd_site_states alone (CUDA events) on a C3 window for tuning knobs given in the env.
usage: python tools/k1_sweep.py [frames]   (GD_GRID_K / GD_K1_SPG are read by the library)"""
import json, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from gemdat_b200 import synth
from gemdat_b200.pipeline import HotPath

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
rec = synth.config_c3(n_frames=frames)
pos = rec.device_positions(0, frames)
hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=rec.n_atoms, site_radius=0.6,
             site_inner_fraction=0.5, minimal_residence=3, n_frames_total=frames)
ms = []
for i in range(8):
    hp.run(pos)
    torch.cuda.synchronize()
    a, b = hp.last_k1_events
    if i >= 3:
        ms.append(a.elapsed_time(b))
t = sum(ms) / len(ms)
hp.profile_phases = True
acc = {}
for i in range(4):
    hp.run(pos)
    for k, v in hp.last_phases.items():
        acc[k] = acc.get(k, 0.0) + v / 4
phases = {k: round(v, 3) for k, v in acc.items() if k in ('k2_events', 'k3_jumps')}
print(json.dumps({'grid_k': os.environ.get('GD_GRID_K'), 'spg': os.environ.get('GD_K1_SPG'), 'frames': frames,
                  'k1_ms': round(t, 4), 'GBps': round(32 * frames * rec.n_atoms / t / 1e6, 1),
                  'phases_ms': phases}))
