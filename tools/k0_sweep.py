"""K0 fused kernel: resident CTAs per SM (L2 footprint of the tiles between the two passes) and
K2 compaction grid size, timed with CUDA events.  usage: python tools/k0_sweep.py [frames]"""
import os
import sys

sys.path.insert(0, '.')
import numpy as np
import torch

from gemdat_b200 import _device as dv
from gemdat_b200 import _lib, synth

T = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
lib = _lib.lib
rec = synth.config_c3(n_frames=T)
A = rec.n_atoms
pos = rec.device_positions(0, T)
s = dv.stream_ptr()
lat = np.ascontiguousarray(rec.lattice, dtype=np.float64)
work = dv.empty((int(lib.gd_unwrap_chunks(T, A)) * A * 3,), torch.int32)
cart = dv.empty((T, A, 3), torch.float64)
wind = dv.empty((A, 3), torch.int32)


def run():
    _lib.check(lib.gd_unwrap(dv.ptr(pos), T, A, lat.ctypes.data, None, None, None, dv.ptr(work), None,
                             dv.ptr(cart), None, None, dv.ptr(wind), s))


def best(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    out = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out = min(out, a.elapsed_time(b))
    return out


os.environ['GD_UNWRAP_FUSED'] = '0'
ms = best(run)
ref = cart.clone()
print(f'three-pass: {ms:.3f} ms  {48 * T * A / ms / 1e6:.0f} GB/s (48 B/atom-frame)', flush=True)
os.environ['GD_UNWRAP_FUSED'] = '1'
for prefetch in ('0', '1'):
    os.environ['GD_UNWRAP_PREFETCH'] = prefetch
    for per_sm in (3, 4, 5, 6, 8, 12):
        os.environ['GD_UNWRAP_CTAS_PER_SM'] = str(per_sm)
        ms = best(run)
        print(f'fused, L2 prefetch {prefetch}, {per_sm} CTAs/SM: {ms:.3f} ms  {48 * T * A / ms / 1e6:.0f} GB/s  '
              f'equal={bool(torch.equal(cart, ref))}', flush=True)
