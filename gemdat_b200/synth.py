"""Synthetic Li trajectories of the BASELINE shapes (bench + tests).

The reference's only real-data fixture (tests/data/short_simulation/vasprun.xml) is a
missing blob, and the 10^6 x 10^3 trajectory (24 GB) cannot be generated or held on
the host in reasonable time, so trajectories are *specified* by a small recipe that
both the CUDA generator (``gd_synth_positions``) and numpy (``host_positions``) replay
bit-identically for any frame window -- only integer hashing and individually
rounded IEEE operations are used:

* sites: C2 = the 48h orbit of (0.183, 0.183, 0.024) under F-43m (argyrodite,
  a = 9.924 A, cf. reference src/gemdat/data/argyrodite.cif) in a 2x1x1 supercell
  (96 sites, 48 Li); C3 = an LLZO-like cubic cell (a = 12.97 A, 120 sites: 24 labelled
  '24d', 96 labelled '96h', >= 1.7 A apart) in a 3x3x2 supercell (2160 sites, 1008 Li);
* every atom sits on a site for an epoch of ``epoch_len`` frames; with probability
  ``p_hop`` it then moves to a free neighbouring site, travelling linearly along the
  minimum-image vector during the last 7 frames of the epoch (weights k/8);
* thermal motion = piecewise-linear interpolation (knots every ``knot_len`` frames)
  of Irwin-Hall(4) integers from a splitmix64 hash of (seed, atom, component, knot),
  scaled to ``sigma`` Angstrom;  x = wrap(centre + w*delta + n*scale).
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _device as dv
from . import _lib
from .lattice import Lattice, Sites
from .trajectory import Trajectory

_IH_STD = np.sqrt(4 * (65536.0**2 - 1) / 12.0)  # std of the sum of four uniform uint16


def argyrodite_sites(supercell=(2, 1, 1)) -> Sites:
    a = 9.924
    x, z = 0.183, 0.024
    pts = []
    for arr in ((x, x, z), (x, z, x), (z, x, x)):
        for sg in ((1, 1, 1), (1, -1, -1), (-1, 1, -1), (-1, -1, 1)):
            for cen in ((0, 0, 0), (0, 0.5, 0.5), (0.5, 0, 0.5), (0.5, 0.5, 0)):
                pts.append([(s * v + c) % 1.0 for s, v, c in zip(sg, arr, cen)])
    sites = Sites(Lattice(np.eye(3) * a), 'Li', np.array(pts), labels=['48h'] * len(pts))
    return sites.make_supercell(supercell)


def garnet_like_sites(supercell=(3, 3, 2), seed: int = 7, min_dist: float = 1.7) -> Sites:
    """LLZO-like site set: 120 sites per cubic cell (no LLZO CIF ships with the reference)."""
    a = 12.97
    rng = np.random.default_rng(seed)
    pts: list[np.ndarray] = []
    while len(pts) < 120:
        p = rng.random(3)
        ok = True
        for q in pts:
            d = p - q
            d -= np.round(d)
            if np.linalg.norm(d) * a < min_dist:
                ok = False
                break
        if ok:
            pts.append(p)
    labels = ['24d'] * 24 + ['96h'] * 96
    sites = Sites(Lattice(np.eye(3) * a), 'Li', np.array(pts), labels=labels)
    return sites.make_supercell(supercell)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)).astype(np.uint64)
    x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)).astype(np.uint64)
    return x ^ (x >> np.uint64(31))


def _noise_knot(seed: int, atom: np.ndarray, comp: np.ndarray, knot: np.ndarray) -> np.ndarray:
    with np.errstate(over='ignore'):
        key = (atom.astype(np.uint64) * np.uint64(3) + comp.astype(np.uint64)) * np.uint64(
            0xD1B54A32D192ED03
        ) + knot.astype(np.uint64) * np.uint64(0x8CB92BA72F3D8DD7)
        h = _splitmix64(np.uint64(seed) ^ key)
    s = (
        (h & np.uint64(0xFFFF)).astype(np.int64)
        + ((h >> np.uint64(16)) & np.uint64(0xFFFF)).astype(np.int64)
        + ((h >> np.uint64(32)) & np.uint64(0xFFFF)).astype(np.int64)
        + ((h >> np.uint64(48)) & np.uint64(0xFFFF)).astype(np.int64)
    )
    return s - 131070


@dataclass
class SynthRecipe:
    sites: Sites
    n_atoms: int
    n_frames: int
    centre: np.ndarray  # [A, n_epochs, 3] fractional site centre per epoch
    delta: np.ndarray   # [A, n_epochs, 3] min-image vector to the next epoch's site
    site_of: np.ndarray  # [A, n_epochs] site index per epoch
    epoch_len: int
    knot_len: int
    noise_scale: np.ndarray  # [3] fractional units per integer noise unit
    seed: int
    time_step: float = 2e-15

    @property
    def lattice(self) -> np.ndarray:
        return self.sites.lattice.matrix

    @property
    def n_epochs(self) -> int:
        return self.centre.shape[1]

    # ---------------------------------------------------------------- host replay
    def host_positions(self, frame0: int = 0, n_frames: int | None = None) -> np.ndarray:
        """numpy replay of frames [frame0, frame0 + n_frames): float64 [n, A, 3]."""
        n = self.n_frames - frame0 if n_frames is None else n_frames
        t = np.arange(frame0, frame0 + n, dtype=np.int64)
        e = t // self.epoch_len
        phi = t - e * self.epoch_len
        w = np.maximum(0, phi - (self.epoch_len - 8)).astype(np.float64) * 0.125
        k = t // self.knot_len
        q = t - k * self.knot_len
        atoms = np.arange(self.n_atoms)
        out = np.empty((n, self.n_atoms, 3))
        for c in range(3):
            comp = np.full((1, 1), c)
            g0 = _noise_knot(self.seed, atoms[None, :], comp, k[:, None])
            g1 = _noise_knot(self.seed, atoms[None, :], comp, k[:, None] + 1)
            nint = g0 * self.knot_len + (g1 - g0) * q[:, None]
            x = self.centre[:, e, c].T + w[:, None] * self.delta[:, e, c].T
            x = x + nint.astype(np.float64) * self.noise_scale[c]
            x = x - np.floor(x)
            x[x >= 1.0] = 0.0
            out[:, :, c] = x
        return out

    # ---------------------------------------------------------------- device generator
    def device_positions(self, frame0: int = 0, n_frames: int | None = None) -> torch.Tensor:
        n = self.n_frames - frame0 if n_frames is None else n_frames
        centre = dv.to_device(self.centre)
        delta = dv.to_device(self.delta)
        pos = dv.empty((n, self.n_atoms, 3), torch.float64)
        scale = np.ascontiguousarray(self.noise_scale, dtype=np.float64)
        _lib.check(
            _lib.lib.gd_synth_positions(
                dv.ptr(centre), dv.ptr(delta), self.n_atoms, self.n_epochs, self.epoch_len,
                self.knot_len, scale.ctypes.data, self.seed, frame0, n, dv.ptr(pos), dv.stream_ptr(),
            )
        )
        return pos

    def trajectory(self, *, device: bool = False, frame0: int = 0, n_frames: int | None = None) -> Trajectory:
        coords = self.device_positions(frame0, n_frames) if device else self.host_positions(frame0, n_frames)
        return Trajectory(
            species=['Li'] * self.n_atoms, coords=coords, lattice=self.lattice,
            metadata={'temperature': 650.0}, time_step=self.time_step,
        )


def make_recipe(
    sites: Sites,
    n_atoms: int,
    n_frames: int,
    *,
    seed: int,
    epoch_len: int = 256,
    knot_len: int = 16,
    p_hop: float = 0.2,
    sigma: float = 0.25,
    neighbour_cut: float = 2.6,
) -> SynthRecipe:
    if epoch_len % knot_len or epoch_len < 16:
        raise ValueError('epoch_len must be a multiple of knot_len and >= 16')
    rng = np.random.default_rng(seed)
    frac = sites.frac_coords
    n_sites = len(frac)
    m = sites.lattice.matrix
    # neighbour lists on the host (small; pure numpy so that no GPU is needed to build a recipe)
    d = frac[None, :, :] - frac[:, None, :]
    d -= np.round(d)
    dist = np.linalg.norm(d @ m, axis=-1)
    np.fill_diagonal(dist, np.inf)
    neigh = [np.nonzero(row < neighbour_cut)[0] for row in dist]

    n_epochs = -(-n_frames // epoch_len) + 1
    site_of = np.empty((n_atoms, n_epochs), dtype=np.int64)
    site_of[:, 0] = rng.choice(n_sites, size=n_atoms, replace=False)
    occupied = np.zeros(n_sites, dtype=bool)
    occupied[site_of[:, 0]] = True
    for e in range(1, n_epochs):
        cur = site_of[:, e - 1].copy()
        movers = np.nonzero(rng.random(n_atoms) < p_hop)[0]
        for a in movers:
            cand = neigh[cur[a]]
            cand = cand[~occupied[cand]]
            if len(cand):
                new = cand[rng.integers(len(cand))]
                occupied[cur[a]] = False
                occupied[new] = True
                cur[a] = new
        site_of[:, e] = cur
    centre = frac[site_of]  # [A, E, 3]
    nxt = np.concatenate([site_of[:, 1:], site_of[:, -1:]], axis=1)
    delta = frac[nxt] - centre
    delta -= np.round(delta)
    lengths = np.array(sites.lattice.lengths)
    noise_scale = sigma / lengths / _IH_STD / knot_len
    return SynthRecipe(
        sites=sites, n_atoms=n_atoms, n_frames=n_frames,
        centre=np.ascontiguousarray(centre), delta=np.ascontiguousarray(delta), site_of=site_of,
        epoch_len=epoch_len, knot_len=knot_len, noise_scale=noise_scale, seed=seed,
    )


def config_c2(n_frames: int = 100_000, seed: int = 20260101) -> SynthRecipe:
    """BASELINE config 2: argyrodite 2x1x1, 48 Li, 96 sites."""
    return make_recipe(argyrodite_sites((2, 1, 1)), 48, n_frames, seed=seed, epoch_len=256, p_hop=0.2)


def config_c3(n_frames: int = 1_000_000, seed: int = 20260102, supercell=(3, 3, 2)) -> SynthRecipe:
    """BASELINE config 3: LLZO-like garnet supercell, ~10^3 Li, 2160 sites."""
    sites = garnet_like_sites(supercell)
    n_atoms = int(round(len(sites) * 1008 / 2160))
    return make_recipe(sites, n_atoms, n_frames, seed=seed, epoch_len=1024, p_hop=0.1)
