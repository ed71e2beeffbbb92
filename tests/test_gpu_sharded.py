"""GPU: ``FluxPlan.flux_sharded`` -- the chunked host-buffer pipeline (copy in / kernel +
all-reduce / copy out) must give exactly what the one-call host path gives, and across two
ranks (NCCL; skipped on a one-GPU box) the wavenumber shards must add up to it."""
import os
import socket

import numpy as np
import pytest

from util import load_case, relmax
from test_gpu_parity import batch_inputs, make_plan

pytestmark = pytest.mark.gpu


def _inputs(n_col):
    c = load_case('two_mol_nz41')
    T, T_g, T_tau, albedo, q = batch_inputs(c, 64, seed=3)
    rep = -(-n_col // 64)
    tile = lambda a, axis=0: np.ascontiguousarray(np.concatenate([a] * rep, axis=axis).take(range(n_col), axis=axis))
    return c, tile(T), tile(T_g), tile(T_tau), tile(albedo), tile(q, 1)


def test_pipeline_equals_single_call():
    n_col = 3000                                           # chunks of 1184, 1184, 632 columns
    c, T, T_g, T_tau, albedo, q = _inputs(n_col)
    plan = make_plan(c, tables=True)
    ref = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo, outputs=('net', 'heating'))
    out = {'net': np.full((n_col, plan.L), np.nan), 'heating': np.full((n_col, plan.L - 1), np.nan)}
    for _ in range(2):                                     # second call reuses the staging buffers
        plan.flux_sharded(T, T_g, T_tau, q, albedo, out, n_chunks=3)
        assert np.array_equal(out['net'], ref['net'])
        assert np.array_equal(out['heating'], ref['heating'])
    one_q = np.ascontiguousarray(q[:, 0])                  # q shared by all columns: [M, L]
    ref = plan.flux(T, T_g, T_tau=T_tau, q=one_q, albedo=albedo, outputs=('net', 'heating'))
    plan.flux_sharded(T, T_g, T_tau, one_q, albedo, out, n_chunks=2)
    assert np.array_equal(out['net'], ref['net'])


def _rank(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    try:
        import climatemodel_b200.plan as fp
        n_col = 2500
        c, T, T_g, T_tau, albedo, q = _inputs(n_col)
        plan = fp.FluxPlan(c['P'], c['nu'], c['bands'], tables=c['tables'], device=rank, T_star=c['T_star'],
                           R_star=c['R_star'], star_planet_dist=c['dist'])
        plan.shard_by_wavenumber(rank, world)
        out = {'net': np.zeros((n_col, plan.L)), 'heating': np.zeros((n_col, plan.L - 1))}
        plan.flux_sharded(T, T_g, T_tau, q, albedo, out, n_chunks=2)
        np.savez(os.path.join(out_dir, 'rank%d.npz' % rank), **out)
    finally:
        dist.destroy_process_group()


def test_two_ranks_add_up():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_rank, args=(2, port, d), nprocs=2, join=True)
        got = [np.load(os.path.join(d, 'rank%d.npz' % r)) for r in (0, 1)]
    n_col = 2500
    c, T, T_g, T_tau, albedo, q = _inputs(n_col)
    ref = make_plan(c, tables=True).flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo, outputs=('net', 'heating'))
    assert np.array_equal(got[0]['net'], got[1]['net'])
    assert relmax(got[0]['net'], ref['net']) < 1e-12
    scale = np.abs(ref['heating']).max()
    assert np.abs(got[0]['heating'] - ref['heating']).max() < 1e-9 * scale
