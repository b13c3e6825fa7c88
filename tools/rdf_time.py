import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from gemdat_b200.rdf import _hist
torch.manual_seed(0)
T, A = 2000, 1008
pos = torch.rand((T, A, 3), dtype=torch.float64, device='cuda')
bins = np.arange(0, 5.0 + 0.1, 0.1)
atoms = np.arange(A)
for name, lat in [('diag', np.diag([38.9, 38.9, 26.0])), ('tri', np.array([[38.9, 0, 0], [3.0, 38.9, 0], [1.0, -2.0, 26.0]]))]:
    for mode in (1, 0):
        _hist(pos[:50], lat, atoms, None, 1, atoms, None, 1, bins, 0.1, mode, True)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        h = _hist(pos, lat, atoms, None, 1, atoms, None, 1, bins, 0.1, mode, True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print(name, 'mode', mode, 'ms', round(dt * 1e3, 2), 'pairs/s %.3e' % (T * A * (A - 1) / dt), 'sum', int(h.sum()), flush=True)

# one lattice per frame (gd_rdf_hist_frames): the same cells, repeated per frame
from gemdat_b200.rdf import _hist_frames
for name, lat in [('diag per frame', np.diag([38.9, 38.9, 26.0])),
                  ('tri per frame', np.array([[38.9, 0, 0], [3.0, 38.9, 0], [1.0, -2.0, 26.0]]))]:
    lattices = np.ascontiguousarray(np.broadcast_to(lat.reshape(1, 9), (T, 9)), dtype=np.float64)
    _hist_frames(pos[:50], lattices[:50].copy(), atoms, atoms, bins, 0.1, True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    h = _hist_frames(pos, lattices, atoms, atoms, bins, 0.1, True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(name, 'ms', round(dt * 1e3, 2), 'pairs/s %.3e' % (T * A * (A - 1) / dt), 'sum', int(h.sum()), flush=True)
