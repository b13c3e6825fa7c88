import sys, warnings
sys.path.insert(0, '.')
import numpy as np, torch
from gemdat_b200 import synth
from gemdat_b200.transitions import _site_states_device, widen_to_host, SiteGrid
from oracle import cstates

rec = synth.config_c3(n_frames=1500)
traj = rec.trajectory(device=True)
radius = {'24d': 0.6, '96h': 0.5}
inner = {'24d': 0.5, '96h': 0.8}
pos = traj.positions.cpu().numpy()
print('device==host positions', np.array_equal(pos, rec.host_positions()))
with warnings.catch_warnings():
    warnings.simplefilter('ignore')
    eo, ei = cstates.calculate_atom_states_pair(site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels,
        positions=pos, lattice_matrix=rec.lattice, site_radius=radius, site_inner_fraction=inner)
    for trial in range(3):
        o, i, info = _site_states_device(sites=rec.sites, trajectory=traj, site_radius=radius, site_inner_fraction=inner)
        po, pi = widen_to_host(o), widen_to_host(i)
        bad = np.argwhere(po != eo)
        print('trial', trial, 'outer mismatches', len(bad), 'inner mismatches', int((pi != ei).sum()))
        print(' visited outer', info['visited_outer'].sum(), 'inner', info['visited_inner'].sum())
        for (t, a) in bad[:40]:
            print('  t', t, 'a', a, 'product', po[t, a], 'oracle', eo[t, a], 'pos', pos[t, a])
# raw (no remap) via direct lib call
from gemdat_b200 import _lib, _device as dv
frac = rec.sites.frac_coords
grid = SiteGrid(lattice_matrix=rec.lattice, site_frac=frac, site_labels=rec.sites.labels, site_radius=radius, site_inner_fraction=inner)
p = traj._device_positions()
outer = dv.empty(p.shape[:2], torch.int32); inn = dv.empty(p.shape[:2], torch.int32)
vis = dv.zeros((2*len(frac),), torch.uint8); band = dv.zeros((4,), torch.int64)
_lib.check(_lib.lib.gd_site_states(dv.ptr(p), p.shape[0]*p.shape[1], dv.ptr(grid.d_blob), grid.h_blob.ctypes.data, dv.ptr(outer), dv.ptr(inn), dv.ptr(vis), dv.ptr(band), dv.stream_ptr()))
raw = outer.cpu().numpy().astype(np.int64)
v = vis.cpu().numpy().astype(bool).reshape(2, -1)
print('raw visited outer', v[0].sum(), ' sites appearing in raw output', np.unique(raw[raw>=0]).size)
appear = np.zeros(len(frac), bool); appear[np.unique(raw[raw>=0])] = True
print('visited but not appearing', np.nonzero(v[0] & ~appear)[0][:20], 'appearing but not visited', np.nonzero(~v[0] & appear)[0][:20])
lut = grid.remap_lut(v[0])
exp = raw.copy(); m = raw >= 0; exp[m] = lut[raw[m]] if lut is not None else raw[m]
print('raw+host-lut vs oracle mismatches', int((exp != eo).sum()))
lut_a = grid.remap_lut(appear)
exp2 = raw.copy(); exp2[m] = lut_a[raw[m]] if lut_a is not None else raw[m]
print('raw+appear-lut vs oracle mismatches', int((exp2 != eo).sum()))
