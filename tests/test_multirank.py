"""Frame-sharded path: world_size-2 runs over ``torch.distributed``.

* CPU (gloo, no GPU): the host-side plan -- ``Comm`` collectives, the halo rule for the event
  at a shard's last frame, the global OR of the "state ever changes" flags and ``merge_plan``
  -- with the oracle standing in for the kernels, against the oracle on the whole trajectory.
* GPU (``-m gpu``): two ranks sharing ``cuda:0`` (gloo transport; NCCL needs one GPU per
  rank) run the real ``HotPath`` kernels on their shards and must reproduce the single-rank
  tables bit for bit.
"""

from __future__ import annotations

import os
import socket
import sys
import warnings
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _init(rank: int, world: int, port: int):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)


def _small_case(n_frames=1200):
    from gemdat_b200 import synth

    rec = synth.config_c2(n_frames=n_frames, seed=11)
    return rec, rec.host_positions(0, n_frames)


def _oracle_states(rec, pos):
    from oracle import cstates

    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return cstates.calculate_atom_states_pair(
            site_frac=rec.sites.frac_coords, site_labels=rec.sites.labels, positions=pos,
            lattice_matrix=rec.lattice, site_radius={'': 0.9}, site_inner_fraction={'': 0.5})


def _shard_events(so, si, next_o, next_i, changed_outer, frame0):
    """What gd_events_fill produces for one shard: rows sorted by (atom, time); the event at the
    last frame uses the following shard's first states; atoms whose outer state never changes
    anywhere are skipped (transitions.py:442-443)."""
    if next_o is not None:
        so = np.vstack([so, next_o[None]])
        si = np.vstack([si, next_i[None]])
    rows = []
    for a in range(so.shape[1]):
        if not changed_outer[a]:
            continue
        t = np.nonzero((so[:-1, a] != so[1:, a]) | (si[:-1, a] != si[1:, a]))[0]
        for k in t:
            rows.append((a, so[k, a], so[k + 1, a], si[k, a], si[k + 1, a], frame0 + k))
    return np.array(rows, dtype=np.int64).reshape(-1, 6)


def _cpu_worker(rank: int, world: int, port: int, out_path: str):
    _init(rank, world, port)
    try:
        from gemdat_b200.pipeline import Comm, merge_plan
        from oracle import events as oevents

        rec, pos = _small_case()
        n_frames, n_atoms = pos.shape[:2]
        f0, f1 = n_frames * rank // world, n_frames * (rank + 1) // world
        so, si = _oracle_states(rec, pos[f0:f1])
        comm = Comm(world)

        # halo: first state rows of every shard; rank r uses those of rank r + 1
        first = comm.all_gather(torch.from_numpy(np.stack([so[0], si[0]])))
        assert first.shape == (world, 2, n_atoms)
        nxt = first[rank + 1].numpy() if rank + 1 < world else None
        # "outer state ever changes" is a property of the whole trajectory
        ext = so if nxt is None else np.vstack([so, nxt[0][None]])
        changed = torch.from_numpy((ext[1:] != ext[:-1]).any(axis=0).astype(np.uint8))
        comm.all_reduce_max(changed)
        rows = _shard_events(so, si, None if nxt is None else nxt[0], None if nxt is None else nxt[1],
                             changed.numpy().astype(bool), f0)

        counts = np.bincount(rows[:, 0], minlength=n_atoms) if len(rows) else np.zeros(n_atoms, np.int64)
        all_counts = comm.all_gather(torch.from_numpy(counts.astype(np.int64))).numpy()
        cap = int(all_counts.sum(axis=1).max())
        padded = np.zeros((max(cap, 1), 6), dtype=np.int64)
        padded[: len(rows)] = rows
        gathered = comm.all_gather(torch.from_numpy(padded)).numpy()
        atom_offsets, dst_start = merge_plan(all_counts)
        merged = np.zeros((int(atom_offsets[-1]), 6), dtype=np.int64)
        for r in range(world):
            src_off = np.concatenate([[0], np.cumsum(all_counts[r])])
            for a in range(n_atoms):
                n = all_counts[r, a]
                merged[dst_start[r, a]: dst_start[r, a] + n] = gathered[r, src_off[a]: src_off[a] + n]

        full_o, full_i = _oracle_states(rec, pos)
        expected = oevents.calculate_transition_events(full_o, full_i)
        np.testing.assert_array_equal(merged, expected)

        # the product's route: all-to-all by atom block, per-atom merge of the pieces, and the
        # ragged gather that reassembles the table
        lo, hi = n_atoms * rank // world, n_atoms * (rank + 1) // world
        off = np.concatenate([[0], np.cumsum(all_counts, axis=1)[rank]])
        blocks = [(n_atoms * r // world, n_atoms * (r + 1) // world) for r in range(world)]
        send = [int(off[b] - off[a]) for a, b in blocks]
        recv = [int(all_counts[r, lo:hi].sum()) for r in range(world)]
        pieces = comm.all_to_all_rows(torch.from_numpy(rows), send, recv)
        b_off, b_dst = merge_plan(all_counts[:, lo:hi])
        mine = np.zeros((int(b_off[-1]), 6), dtype=np.int64)
        for r in range(world):
            piece = pieces[r].numpy()
            src_off = np.concatenate([[0], np.cumsum(all_counts[r, lo:hi])])
            for a in range(hi - lo):
                n = all_counts[r, lo + a]
                mine[b_dst[r, a]: b_dst[r, a] + n] = piece[src_off[a]: src_off[a] + n]
        np.testing.assert_array_equal(mine, expected[(expected[:, 0] >= lo) & (expected[:, 0] < hi)])
        np.testing.assert_array_equal(comm.all_gather_rows(torch.from_numpy(mine)).numpy(), expected)
        if rank == 0:
            Path(out_path).write_text(f'ok {len(expected)}')
    finally:
        dist.destroy_process_group()


def test_sharded_plan_world2_gloo_cpu(tmp_path):
    out = tmp_path / 'result.txt'
    mp.spawn(_cpu_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    assert out.read_text().startswith('ok ')
    assert int(out.read_text().split()[1]) > 50


def test_merge_plan_places_rows_by_atom_then_shard():
    from gemdat_b200.pipeline import merge_plan

    counts = np.array([[2, 0, 3], [1, 4, 0]])
    atom_offsets, dst = merge_plan(counts)
    np.testing.assert_array_equal(atom_offsets, [0, 3, 7, 10])
    np.testing.assert_array_equal(dst, [[0, 3, 7], [2, 3, 10]])


def _jump_protocol_worker(rank: int, world: int, port: int, out_path: str):
    """The frame-sharded K3 protocol (include/gemdat_b200.h) with the oracle's explicit-state
    machine standing in for the kernels: pass 1 from the empty state, all-gather of the per-atom
    records, the resolve rule, extra rounds while some entry is unknown, final pass, per-atom
    merge of the shards' jump tables -- against the sequential oracle on the whole table."""
    _init(rank, world, port)
    try:
        from gemdat_b200.pipeline import merge_plan
        from oracle import events as oevents
        from oracle import jumps as ojumps

        rec, pos = _small_case(n_frames=2400)
        n_frames, n_atoms = pos.shape[:2]
        so, si = _oracle_states(rec, pos)
        events = oevents.calculate_transition_events(so, si)
        # atom 0 keeps only a few inner-only rows in the middle shard(s): events but no departure
        # there, so its exit depends on its entry and the resolve needs more than one round
        f0, f1 = n_frames * rank // world, n_frames * (rank + 1) // world
        mid = (events[:, 5] >= n_frames // world) & (events[:, 5] < n_frames - n_frames // world)
        drop = (events[:, 0] == 0) & mid & (events[:, 1] != events[:, 2])
        events = events[~drop]
        mine = events[(events[:, 5] >= f0) & (events[:, 5] < f1)]
        per_atom = [mine[mine[:, 0] == a] for a in range(n_atoms)]
        mr = 3

        def run(entries):
            out = [ojumps.jump_machine_slice(per_atom[a], mr, entries[a][0]) for a in range(n_atoms)]
            recs = [(st, len(per_atom[a]) > 0, dep, bool(dep or entries[a][1]))
                    for a, (_, st, dep) in enumerate(out)]  # (state, has events, synced, known)
            return [j for j, _, _ in out], recs

        def resolve(allrecs, q):
            entry, unknown = [], 0
            for a in range(n_atoms):
                e = (ojumps.EMPTY_STATE, True)
                for p in range(q - 1, -1, -1):
                    st, has, _, known = allrecs[p][a]
                    if not has:
                        continue
                    e = (st, known)
                    break
                unknown += 0 if e[1] else 1
                entry.append(e)
            return entry, unknown

        entries = [(ojumps.EMPTY_STATE, rank == 0)] * n_atoms
        rounds = 0
        while True:
            jumps, recs = run(entries)
            allrecs = [None] * world
            dist.all_gather_object(allrecs, recs)
            entries, _ = resolve(allrecs, rank)
            unresolved = sum(resolve(allrecs, q)[1] for q in range(world))  # same on every rank
            rounds += 1
            if unresolved == 0:
                break
            assert rounds < world
        jumps, recs = run(entries)
        if rank == world - 1:
            for a in range(n_atoms):
                jumps[a] = jumps[a] + ojumps.close_pending(recs[a][0], n_frames, mr)
        tables = [None] * world
        dist.all_gather_object(tables, jumps)
        counts = np.array([[len(t[a]) for a in range(n_atoms)] for t in tables])
        total_off, dst = merge_plan(counts)
        merged = np.zeros((int(total_off[-1]), 6), dtype=np.int64)
        for r in range(world):
            for a in range(n_atoms):
                if counts[r, a]:
                    merged[dst[r, a]: dst[r, a] + counts[r, a]] = np.array(tables[r][a])
        merged = merged[merged[:, 1] != merged[:, 2]]
        expected = ojumps.transitions_to_jumps(events, n_frames, mr)
        np.testing.assert_array_equal(merged, expected)
        if rank == 0:
            Path(out_path).write_text(f'ok {len(expected)} {rounds}')
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_jump_state_protocol_gloo_cpu(tmp_path, world):
    out = tmp_path / 'result.txt'
    mp.spawn(_jump_protocol_worker, args=(world, _free_port(), str(out)), nprocs=world, join=True)
    tag, n_jumps, rounds = out.read_text().split()
    assert tag == 'ok' and int(n_jumps) > 20
    if world == 3:
        assert int(rounds) >= 2, 'the thinned atom should have needed a second resolve round'


# ------------------------------------------------------------------------------------ GPU


def _gpu_worker(rank: int, world: int, port: int, out_path: str, exchange: bool):
    os.environ['GD_K3_EXCHANGE'] = '1' if exchange else '0'
    os.environ['GD_EVENTS_STAGE_CAP'] = '32'  # 32-frame chunks here: staging never overflows
    _init(rank, world, port)
    try:
        torch.cuda.set_device(0)
        from gemdat_b200.pipeline import HotPath

        rec, pos = _small_case(n_frames=3000)
        n_frames, n_atoms = pos.shape[:2]
        f0, f1 = n_frames * rank // world, n_frames * (rank + 1) // world
        hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=n_atoms, site_radius=0.9,
                     site_inner_fraction=0.5, minimal_residence=3, n_frames_total=n_frames,
                     frame0=f0, world=world, rank=rank)
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            d_pos = torch.from_numpy(pos[f0:f1]).cuda()
            first = hp.run(d_pos)['jumps'].cpu().numpy()
            res = hp.run(d_pos)  # second run: queued without read-backs, one gather of offsets + rows
        # (a rejected optimistic pass -- e.g. an atom with events but no departure in a shard -- is
        # redone by the checked pass; with 2 shards of this trajectory it goes through)
        if exchange:
            assert hp.last_pass == 'checked'
        elif world == 2:
            assert hp.last_pass == 'optimistic', hp.last_optimistic_reject
        else:
            assert hp.last_pass == 'optimistic' or 'status' in hp.last_optimistic_reject
        np.testing.assert_array_equal(res['jumps'].cpu().numpy(), first)
        np.savez(f'{out_path}.{rank}.npz', outer=res['outer'].cpu().numpy(), inner=res['inner'].cpu().numpy(),
                 events=hp.gather_events(res).cpu().numpy(), jumps=res['jumps'].cpu().numpy(),
                 n_events=res['n_events'], layout=res['events_layout'],
                 block=np.array(res.get('atom_block', (f0, f1))), local_events=res['events'].cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize('world,exchange', [(2, False), (3, False), (2, True)])
def test_hotpath_frame_shards_equal_one(tmp_path, world, exchange):
    """Ranks sharing cuda:0 (gloo transport) run the real kernels on their frame shards.  Default
    route: events stay on their frame shard, only jump-machine states cross the boundaries.
    ``exchange``: the fallback route (event rows change hands by atom block)."""
    from gemdat_b200.pipeline import HotPath

    out = str(tmp_path / 'rank')
    mp.spawn(_gpu_worker, args=(world, _free_port(), out, exchange), nprocs=world, join=True)
    rec, pos = _small_case(n_frames=3000)
    hp = HotPath(sites=rec.sites, lattice=rec.lattice, n_atoms=pos.shape[1], site_radius=0.9,
                 site_inner_fraction=0.5, minimal_residence=3, n_frames_total=len(pos))
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        one = hp.run(torch.from_numpy(pos).cuda())
    ranks = [np.load(f'{out}.{r}.npz') for r in range(world)]
    np.testing.assert_array_equal(np.vstack([r['outer'] for r in ranks]), one['outer'].cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['inner'] for r in ranks]), one['inner'].cpu().numpy())
    ev = one['events'].cpu().numpy()
    assert len(ev) > 100 and len(one['jumps']) > 0
    for r in ranks:  # every rank holds the whole jump table and can assemble the whole event table
        np.testing.assert_array_equal(r['events'], ev)
        np.testing.assert_array_equal(r['jumps'], one['jumps'].cpu().numpy())
        assert int(r['n_events']) == len(ev)
        lo, hi = r['block']
        if exchange:  # each rank keeps the rows of its atom block
            assert str(r['layout']) == 'atom_block'
            np.testing.assert_array_equal(r['local_events'], ev[(ev[:, 0] >= lo) & (ev[:, 0] < hi)])
        else:         # each rank keeps the rows of its frames, sorted by (atom, time)
            assert str(r['layout']) == 'frame_shard'
            np.testing.assert_array_equal(r['local_events'], ev[(ev[:, 5] >= lo) & (ev[:, 5] < hi)])


# ------------------------------------------------------------------------------------ K0 / fills / K4 / K5 shards


def _walk_case(n_frames=900, n_atoms=13, seed=4):
    rng = np.random.default_rng(seed)
    pos = np.mod(0.3 + np.cumsum(rng.normal(scale=0.2, size=(n_frames, n_atoms, 3)), axis=0), 1)
    lat = np.array([[9.0, 0, 0], [1.0, 10.5, 0], [0.5, -0.7, 12.25]])
    states = rng.integers(0, 7, size=(n_frames, n_atoms))
    states[rng.random(states.shape) < 0.9] = -1
    states[:, 0] = -1
    states[: n_frames // 2 + 50, 1] = -1
    return pos, lat, states


def _shard_rules_worker(rank: int, world: int, port: int, out_path: str):
    """Host-side rules of gemdat_b200/sharded.py with numpy standing in for the kernels."""
    _init(rank, world, port)
    try:
        from gemdat_b200.sharded import FrameShards, shard_base_winding, shard_fill_carry
        from oracle import displ as odispl
        from oracle import fill as ofill

        pos, lat, states = _walk_case()
        T, A = pos.shape[:2]
        fs = FrameShards(lattice=lat, n_atoms=A, n_frames_total=T, world=world, rank=rank)
        f0, f1 = fs.frame_bounds[rank]
        mine = pos[f0:f1]
        # K0: halo row, local integer windings, exclusive sum over shards, apply
        ends = fs.comm.all_gather(torch.from_numpy(np.stack([mine[0], mine[-1]])))
        base_pos = ends[0, 0].numpy()
        prev = ends[rank - 1, 1].numpy() if rank else mine[0]
        steps = np.diff(np.concatenate([prev[None], mine]), axis=0)
        wind = np.cumsum(np.rint(steps).astype(np.int32), axis=0)  # inclusive prefix inside the shard
        totals = fs.comm.all_gather(torch.from_numpy(wind[-1].copy()))
        base = shard_base_winding(totals, rank).numpy()
        cum = (mine - base_pos) - (base + wind)
        want = odispl.cumulative_displacements(pos)[f0:f1]
        np.testing.assert_allclose(cum, want, rtol=0, atol=1e-12)
        # drift correction (FrameShards.drift_correction): the first step of a shard uses the row
        # before it, the summed drift starts from the totals of the shards before (added in shard
        # order), the corrected positions are base + cumulative displacement - summed drift, mod 1
        from oracle import drift as odrift

        sel = np.arange(0, A, 2)
        local_steps = steps - np.rint(steps)
        if rank == 0:
            local_steps[0] = 0.0  # frame 0 of the trajectory has no step
        local_drift = local_steps[:, sel].mean(axis=1)
        summed = np.cumsum(local_drift, axis=0)
        shard_totals = fs.comm.all_gather(torch.from_numpy(summed[-1].copy()))
        for r in range(rank):
            summed = summed + shard_totals[r].numpy()
        corrected = np.mod(base_pos + cum - summed[:, None, :], 1)
        want_pos = odrift.corrected_positions(pos, sel)[f0:f1]
        gap = np.abs(corrected - want_pos)
        assert np.minimum(gap, 1 - gap).max() < 1e-11
        np.testing.assert_allclose(local_drift, odrift.drift(pos, sel)[f0:f1, 0], rtol=1e-12, atol=1e-18)
        # fills: last valid state per shard, nearest valid shard before / after
        for direction, ref in ((1, ofill.ffill), (-1, ofill.bfill)):
            st = states[f0:f1]
            local = ref(st, axis=0)
            last = local[-1] if direction > 0 else local[0]
            all_last = fs.comm.all_gather(torch.from_numpy(last.astype(np.int32)))
            carry = shard_fill_carry(all_last, rank, direction).numpy()
            filled = np.where(local == -1, carry[None, :], local)
            np.testing.assert_array_equal(filled, ref(states, axis=0)[f0:f1])
        # K4: frames -> atoms regrouping (all-to-all of uneven pieces), per-lag partial sums
        cart = torch.from_numpy(pos[f0:f1] @ lat)
        soa = cart.permute(1, 2, 0).contiguous()
        block = fs.exchange_atom_blocks(soa)
        a_lo, a_hi = fs.atoms_of_rank()
        np.testing.assert_array_equal(block.numpy(), np.transpose(pos @ lat, (1, 2, 0))[a_lo:a_hi])
        part = block.sum(dim=(0, 1))[:5].clone()
        fs.comm.all_reduce_sum(part)
        np.testing.assert_allclose(part.numpy(), np.transpose(pos @ lat, (1, 2, 0)).sum(axis=(0, 1))[:5], rtol=1e-12)
        key = torch.tensor([7 + rank, 2**62], dtype=torch.int64)
        fs.comm.all_reduce_min(key)
        assert key.tolist() == [7, 2**62]
        if rank == 0:
            Path(out_path).write_text('ok')
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_shard_rules_gloo_cpu(tmp_path, world):
    out = tmp_path / 'result.txt'
    mp.spawn(_shard_rules_worker, args=(world, _free_port(), str(out)), nprocs=world, join=True)
    assert out.read_text() == 'ok'


def _gpu_shards_worker(rank: int, world: int, port: int, out_path: str):
    _init(rank, world, port)
    try:
        torch.cuda.set_device(0)
        from gemdat_b200.sharded import FrameShards
        from gemdat_b200.volume import grid_edges
        from gemdat_b200.lattice import Lattice

        pos, lat, states = _walk_case()
        T, A = pos.shape[:2]
        fs = FrameShards(lattice=lat, n_atoms=A, n_frames_total=T, world=world, rank=rank)
        f0, f1 = fs.frame_bounds[rank]
        d_pos = torch.from_numpy(pos[f0:f1]).cuda()
        d_states = torch.from_numpy(states[f0:f1].astype(np.int32)).cuda()
        k0 = fs.unwrap(d_pos, want_frac=True, want_cart=True, want_dist=True, want_speed=True)
        msd, (a_lo, a_hi), curve = fs.msd(d_pos)
        idx = np.arange(A)
        bins = np.arange(0, 5.0 + 0.25, 0.25)
        np.savez(f'{out_path}.{rank}.npz',
                 frac=k0['frac'].cpu().numpy(), cart=k0['cart'].cpu().numpy(), dist=k0['dist'].cpu().numpy(),
                 speed=k0['speed'].cpu().numpy(), winding=k0['winding'].cpu().numpy(),
                 prev=fs.fill(d_states, 1).cpu().numpy(), nxt=fs.fill(d_states, -1).cpu().numpy(),
                 grid=fs.density(d_pos, grid_edges(Lattice(lat), 0.5)).cpu().numpy(),
                 counts=fs.site_counts(d_states, 7).cpu().numpy(),
                 msd=msd.cpu().numpy(), block=np.array([a_lo, a_hi]), curve=curve.cpu().numpy(),
                 rdf=fs.species_rdf_counts(d_pos, idx, idx, bins, 0.25, True).cpu().numpy(),
                 corrected=fs.drift_correction(d_pos, np.arange(0, A, 2)).cpu().numpy(),
                 com=fs.center_of_mass(d_pos, np.linspace(1.0, 2.0, A)).cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize('world', [2, 3])
def test_frame_shards_equal_one(tmp_path, world):
    """K0, ffill/bfill, K5, N1 counts, K4 (frames regrouped by atom, per-lag partial sums
    all-reduced), drift correction / centre of mass and the species RDF over frame shards (ranks sharing cuda:0, gloo transport)
    against one shard: bit-identical, the atom-averaged MSD curve to 1e-12."""
    from gemdat_b200.sharded import FrameShards
    from gemdat_b200.volume import grid_edges
    from gemdat_b200.lattice import Lattice

    out = str(tmp_path / 'rank')
    mp.spawn(_gpu_shards_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    pos, lat, states = _walk_case()
    T, A = pos.shape[:2]
    fs = FrameShards(lattice=lat, n_atoms=A, n_frames_total=T)
    d_pos, d_states = torch.from_numpy(pos).cuda(), torch.from_numpy(states.astype(np.int32)).cuda()
    k0 = fs.unwrap(d_pos, want_frac=True, want_cart=True, want_dist=True, want_speed=True)
    msd, _, curve = fs.msd(d_pos)
    idx, bins = np.arange(A), np.arange(0, 5.0 + 0.25, 0.25)
    ranks = [np.load(f'{out}.{r}.npz') for r in range(world)]
    np.testing.assert_array_equal(np.vstack([r['frac'] for r in ranks]), k0['frac'].cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['cart'] for r in ranks]), k0['cart'].cpu().numpy())
    np.testing.assert_array_equal(np.hstack([r['dist'] for r in ranks]), k0['dist'].cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['speed'] for r in ranks]), k0['speed'].cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['prev'] for r in ranks]), fs.fill(d_states, 1).cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['nxt'] for r in ranks]), fs.fill(d_states, -1).cpu().numpy())
    np.testing.assert_array_equal(np.vstack([r['msd'] for r in ranks]), msd.cpu().numpy())
    grid = fs.density(d_pos, grid_edges(Lattice(lat), 0.5)).cpu().numpy()
    counts = fs.site_counts(d_states, 7).cpu().numpy()
    rdf = fs.species_rdf_counts(d_pos, idx, idx, bins, 0.25, True).cpu().numpy()
    for r in ranks:
        np.testing.assert_array_equal(r['winding'], k0['winding'].cpu().numpy())
        np.testing.assert_array_equal(r['grid'], grid)
        np.testing.assert_array_equal(r['counts'], counts)
        np.testing.assert_array_equal(r['rdf'], rdf)
        np.testing.assert_allclose(r['curve'], curve.cpu().numpy(), rtol=1e-12)
    assert int(grid.sum()) == T * A and int(counts.sum()) == T * A
    # drift correction: the summed drift is grouped by shard, so equal up to its rounding (mod 1)
    want = fs.drift_correction(d_pos, np.arange(0, A, 2)).cpu().numpy()
    got = np.vstack([r['corrected'] for r in ranks])
    diff = np.abs(got - want)
    assert np.minimum(diff, 1.0 - diff).max() < 1e-12
    np.testing.assert_array_equal(np.vstack([r['com'] for r in ranks]),
                                  fs.center_of_mass(d_pos, np.linspace(1.0, 2.0, A)).cpu().numpy())
