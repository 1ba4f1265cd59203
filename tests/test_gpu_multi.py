"""Multi-GPU parity as `-m gpu` tests (they skip on a box with fewer than 2 devices): one process per GPU over NCCL.

  * population sharding: the sharded solve -- one ncclAllGather of candidate costs per CEM iteration, issued
    by libsecem.so -- equals the unsharded solve of the same problem, on the fused small-model path, on the tcgen05
    path, and for a population above the selection kernel's shared-memory cache;
  * instance sharding (BASELINE config C4): independent MPC instances dealt out to the ranks, results gathered on rank
    0, equal to one batched solve of all instances on one GPU.
"""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [(100, 512, 5, 16, 4), (300, 4096, 6, 64, 3), (300, 40960, 3, 640, 2)]  # n_train, N, T, E, I


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from safe_exploration_b200 import dist as sedist, workloads
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = f'cuda:{rank}'
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device(dev))
    out = {}
    try:
        for (n_train, N, T_len, E, I) in CASES:
            w = workloads.pendulum_workload(n_train=n_train, num_rollouts=N, time_horizon=T_len, num_elites=E,
                                            num_iterations=I)
            ssm, mpc = workloads.build_solver(w, device=dev, record='costs', seed=5)
            comm = sedist.attach_population_sharding(mpc)
            p0 = torch.as_tensor(w.p0, device=dev).unsqueeze(0)
            res = mpc.solve(p0, seed=5)
            torch.cuda.synchronize()
            out[(n_train, N)] = {k: res[k].cpu() for k in ('elite_ids', 'means', 'stds', 'actions', 'best_id',
                                                            'objective', 'constraint')}
            mpc._destroy()
            sedist.destroy_nccl_comm(comm)
        # C4 shape: 12 instances x population 512, instance-sharded, gathered on rank 0
        w = workloads.pendulum_workload(n_train=300, num_rollouts=512, time_horizon=4, num_elites=16, num_iterations=3)
        M = 12
        p0_all = 0.05 * torch.randn((M, 2), generator=torch.Generator().manual_seed(3), dtype=torch.float64)
        first, count = sedist.shard_instances(M, rank, world)
        ssm, mpc = workloads.build_solver(w, device=dev, record='none', num_instances=count)
        plans, found = sedist.solve_instances_sharded(mpc, p0_all, seed=50)
        if rank == 0:
            out['c4'] = dict(plans=plans.cpu(), found=found.cpu(), p0=p0_all)
        torch.save(out, os.path.join(out_dir, f'rank{rank}.pt'))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs at least 2 GPUs')
def test_sharded_solves_equal_the_unsharded_solve(tmp_path):
    import torch.multiprocessing as mp
    from safe_exploration_b200 import workloads
    world = 2
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    ranks = [torch.load(os.path.join(tmp_path, f'rank{r}.pt')) for r in range(world)]
    dev = 'cuda:0'
    for (n_train, N, T_len, E, I) in CASES:
        w = workloads.pendulum_workload(n_train=n_train, num_rollouts=N, time_horizon=T_len, num_elites=E,
                                        num_iterations=I)
        ssm, mpc = workloads.build_solver(w, device=dev, record='costs', seed=5)
        ref = mpc.solve(torch.as_tensor(w.p0, device=dev).unsqueeze(0), seed=5)
        ref = {k: v.cpu() for k, v in ref.items() if torch.is_tensor(v)}
        n_l = N // world
        for r, got in enumerate(ranks):
            g = got[(n_train, N)]
            assert torch.equal(g['elite_ids'], ref['elite_ids']), (n_train, N, r)  # every rank: the same decision
            assert torch.equal(g['best_id'], ref['best_id'])
            assert torch.allclose(g['means'], ref['means'], rtol=1e-12, atol=1e-15)
            assert torch.allclose(g['stds'], ref['stds'], rtol=1e-12, atol=1e-15)
            assert torch.allclose(g['actions'], ref['actions'], rtol=1e-12, atol=1e-15)
            assert torch.equal(g['constraint'], ref['constraint'][:, :, r * n_l:(r + 1) * n_l])
            assert torch.allclose(g['objective'], ref['objective'][:, :, r * n_l:(r + 1) * n_l], rtol=1e-12, atol=0)
    # instance sharding == one batched solve of all instances (the candidate RNG is keyed by GLOBAL instance id)
    c4 = ranks[0]['c4']
    w = workloads.pendulum_workload(n_train=300, num_rollouts=512, time_horizon=4, num_elites=16, num_iterations=3)
    M = c4['p0'].shape[0]
    ssm, mpc = workloads.build_solver(w, device=dev, record='none', num_instances=M)
    one = mpc.solve(c4['p0'].to(dev), seed=50)
    assert torch.equal(one['found'].bool().cpu(), c4['found']) and bool(c4['found'].any())
    sel = c4['found']
    assert torch.equal(one['actions'].cpu()[sel], c4['plans'][sel])
