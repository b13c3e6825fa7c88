"""A/B of the C3 step under environment knobs (phases in ms, eager launches with a
synchronisation at every phase boundary).  usage: python tools/step_ab.py [frames]"""
import os
import sys
import warnings

sys.path.insert(0, '.')
import torch

from gemdat_b200 import synth
from gemdat_b200.pipeline import HotPath

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rec = synth.config_c3(n_frames=T)
pos = rec.device_positions(0, T)


def phases(label):
    hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=rec.n_atoms, site_radius=0.6,
                 site_inner_fraction=0.5, minimal_residence=3, n_frames_total=T)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        for _ in range(3):
            hp.run(pos)
        hp.profile_phases = True
        acc = {}
        for _ in range(3):
            hp.run(pos)
            for k, v in hp.last_phases.items():
                acc[k] = acc.get(k, 0.0) + v / 3
        hp.profile_phases = False
        # whole step, graph replay
        for _ in range(2):
            hp.run(pos)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            hp.run(pos)
        b.record()
        torch.cuda.synchronize()
    print(f'{label}: step {a.elapsed_time(b) / 5:.3f} ms ({hp.last_launch_mode})  ' +
          '  '.join(f'{k} {v:.3f}' for k, v in acc.items()), flush=True)


for per_sm in (os.environ.get("GD_COMPACT_CTAS_PER_SM", "100000"),):
    os.environ['GD_COMPACT_CTAS_PER_SM'] = per_sm
    phases(f'compact CTAs/SM {per_sm} warp-compaction {os.environ.get("GD_COMPACT_WARP", "default")} tiled-spec {os.environ.get("GD_K3_SPEC_TILED", "default")}')
