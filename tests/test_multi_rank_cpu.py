"""CPU, world_size 2 over gloo: the host side of the wavenumber-sharded evaluation.
Every rank takes its contiguous range of segments (``plan.shard_edges``), holds the partial band
sums of that range -- computed here by the oracle's per-bin form, which follows the kernel's rule
for where a band's boundary terms go -- packs net flux and heating rate into one buffer the way
``bench.py`` does, and one all-reduce must give the unsharded result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import load_case, load_table, relmax

from climatemodel_b200.plan import segment_table, shard_edges
import oracle.realgas_oracle as orc


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _bands(c):
    return {'start': c['band_start'], 'len': c['band_len'], 'centre': c['band_centre'], 'delta': c['band_delta'],
            'sw': c['band_sw']}


def _integrand(c):
    tables = [load_table(str(m)) for m in c['molecule_names']]
    P = np.sort(c['P'])
    return orc.layer_integrand(P, c['T_interface_tau'][np.argsort(c['P'])], c['nu'], tables, list(c['q']))


def _worker(rank, world, port, name, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        c = load_case(name)
        bands = _bands(c)
        order = np.argsort(c['P'])                         # oracle grids run top -> surface
        P, T = c['P'][order], c['T'][np.argsort(c['p'])]
        seg_band, is_lw = segment_table(bands['len'], bands['sw'])
        edges = shard_edges(is_lw, world)
        up, dn = orc.get_flux_per_bin(P, T, float(c['T_g']), _integrand(c), c['nu'], bands,
                                      albedo=float(c['albedo']), segment_range=(edges[rank], edges[rank + 1]))
        net = orc.net_flux(up, dn, bands['delta'])
        heat = orc.heating_rate(net, P)
        buf = torch.from_numpy(np.concatenate([net, heat]))            # bench.py's fused output buffer
        dist.all_reduce(buf)
        if rank == 0:
            np.savez(os.path.join(out_dir, 'reduced.npz'), net=buf[:P.size].numpy(), heat=buf[P.size:].numpy(),
                     edges=np.array(edges))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('name', ['two_mol_nz41'])
def test_two_ranks_add_up_to_the_unsharded_result(tmp_path, name):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), name, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / 'reduced.npz')
    c = load_case(name)
    order = np.argsort(c['P'])
    assert relmax(got['net'], c['net'][order]) < 1e-12                 # the reference's own net flux
    assert relmax(got['heat'], c['heating'][np.argsort(c['p'])]) < 1e-9
    edges = got['edges']
    assert edges[0] == 0 and 0 < edges[1] < edges[2]


def test_shard_edges_tile_the_segments_and_balance_cost():
    c = load_case('co2_nz50_earth')
    seg_band, is_lw = segment_table(c['band_len'], c['band_sw'])
    assert seg_band.size == int(np.sum(-(-c['band_len'] // 32)))
    assert np.all(np.diff(seg_band) >= 0)
    for world in (1, 2, 3, 4, 8):
        edges = shard_edges(is_lw, world, lw_cost=2.5)
        assert len(edges) == world + 1 and edges[0] == 0 and edges[-1] == is_lw.size
        assert np.all(np.diff(edges) >= 0)
        cost = np.where(is_lw, 2.5, 1.0)
        shares = [cost[a:b].sum() for a, b in zip(edges[:-1], edges[1:])]
        assert max(shares) - min(shares) <= 2 * 2.5 + 1e-9             # within one segment either side


def test_oracle_partial_sums_follow_the_kernel_rule():
    """Boundary terms of a band travel with its first segment; partial sums tile exactly."""
    c = load_case('co2_nz50_earth')
    bands = _bands(c)
    order = np.argsort(c['P'])
    P, T = c['P'][order], c['T'][np.argsort(c['p'])]
    I = _integrand(c)
    full = orc.get_flux_per_bin(P, T, float(c['T_g']), I, c['nu'], bands, albedo=float(c['albedo']))
    n_seg = segment_table(bands['len'], bands['sw'])[0].size
    cuts = [0, 7, 8, n_seg // 2, n_seg]
    parts = [orc.get_flux_per_bin(P, T, float(c['T_g']), I, c['nu'], bands, albedo=float(c['albedo']),
                                  segment_range=(a, b)) for a, b in zip(cuts[:-1], cuts[1:])]
    for k in (0, 1):
        assert relmax(sum(p[k] for p in parts), full[k]) < 1e-13


def test_wavenumber_shard_count_rule():
    """Ranks sharing one set of columns: as many as leave >= 128 segments each, and a divisor of N."""
    from climatemodel_b200.plan import wavenumber_shards
    assert [wavenumber_shards(332, n) for n in (1, 2, 4, 8)] == [1, 2, 2, 2]        # bench workload, 10.4k bins
    assert [wavenumber_shards(31250, n) for n in (2, 4, 8)] == [2, 4, 8]            # 1M-bin sweep
    assert wavenumber_shards(100, 8) == 1 and wavenumber_shards(400, 6) == 3


# ---- column sharding (independent columns, BASELINE configs[4]): split and gather ------------------------
def _column_worker(rank, world, port, n_all, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from climatemodel_b200.radiation.columns import column_share, gather_columns
        mine = column_share(n_all, rank, world)
        rows = torch.tensor(np.stack([mine * 10.0 + k for k in range(3)], axis=1))      # stand-in for (T, t, steps)
        full = gather_columns(rows, n_all, rank, world)
        np.save(os.path.join(out_dir, 'full_%d.npy' % rank), full.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_all', [7, 8])
def test_column_shares_tile_the_ensemble_and_gather_restores_order(tmp_path, n_all):
    from climatemodel_b200.radiation.columns import column_share
    for world in (1, 2, 3, 8):
        shares = [column_share(n_all, r, world) for r in range(world)]
        assert np.array_equal(np.sort(np.concatenate(shares)), np.arange(n_all))
        assert max(map(len, shares)) - min(map(len, shares)) <= 1
    world = 2
    mp.spawn(_column_worker, args=(world, _free_port(), n_all, str(tmp_path)), nprocs=world, join=True)
    want = np.stack([np.arange(n_all) * 10.0 + k for k in range(3)], axis=1)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / ('full_%d.npy' % r)), want)
