"""Timeline of the upload / K1 / download pipeline (events on each stream)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import gemdat_b200.transitions as gt
from gemdat_b200 import Trajectory, synth
frames = 100_000
rec = synth.config_c3(n_frames=frames)
pos = rec.device_positions(0, frames)
host = torch.empty(tuple(pos.shape), dtype=torch.float64, pin_memory=True); host.copy_(pos); torch.cuda.synchronize()
src = torch.from_numpy(host.numpy())
print('src pinned:', src.is_pinned(), 'slice pinned:', src[10:20].is_pinned())
n_atoms = rec.n_atoms
P = 8
b = [frames * k // P for k in range(P + 1)]
d_pos = torch.empty_like(pos)
stage = [torch.empty((frames // P, n_atoms), dtype=torch.int64, device='cuda') for _ in range(2)]
h_out = torch.empty((frames, n_atoms), dtype=torch.int64, pin_memory=True)
main = torch.cuda.current_stream(); up = torch.cuda.Stream(); down = torch.cuda.Stream()
for variant in ('from_numpy', 'torch_pinned'):
    s = src if variant == 'from_numpy' else host
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True); t0.record(main)
        up.wait_stream(main); down.wait_stream(main)
        ev = []
        for k in range(P):
            with torch.cuda.stream(up):
                d_pos[b[k]:b[k+1]].copy_(s[b[k]:b[k+1]], non_blocking=True)
                e_up = torch.cuda.Event(enable_timing=True); e_up.record(up)
            main.wait_event(e_up)
            stage[k % 2].fill_(k)  # stands for K1 + widen
            e_k = torch.cuda.Event(enable_timing=True); e_k.record(main)
            with torch.cuda.stream(down):
                down.wait_event(e_k)
                h_out[b[k]:b[k+1]].copy_(stage[k % 2], non_blocking=True)
                e_dn = torch.cuda.Event(enable_timing=True); e_dn.record(down)
            ev.append((e_up, e_k, e_dn))
        torch.cuda.synchronize()
    print(variant)
    for k, (a, c, d) in enumerate(ev):
        print(f'  part {k}: upload done {t0.elapsed_time(a):6.1f}  kernel done {t0.elapsed_time(c):6.1f}  download done {t0.elapsed_time(d):6.1f}')
