"""CPU, world_size 2, gloo: the host-side logic of population sharding -- shard arithmetic, the
all-gather buffer layout, identical global elite selection on every rank and regeneration of elite actions from the
counter-based RNG by GLOBAL candidate id.  The per-shard rollouts are evaluated by the oracle here (no GPU); the
device path uses the same layout (csrc/cem_kernels.cuh CostView) and is covered by the 2-GPU run of bench.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle
from safe_exploration_b200 import dist as sedist
from safe_exploration_b200 import workloads

T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))


def test_shard_arithmetic():
    assert sedist.shard_range(16384, 3, 8) == (6144, 2048)
    with pytest.raises(ValueError):
        sedist.shard_range(100, 0, 8)
    parts = [sedist.shard_instances(1024, r, 8) for r in range(8)]
    assert parts[0] == (0, 128) and parts[-1] == (896, 128)
    parts = [sedist.shard_instances(10, r, 4) for r in range(4)]
    assert parts == [(0, 3), (3, 3), (6, 2), (8, 2)] and sum(c for _, c in parts) == 10
    assert sedist.owner_of(5000, 16384, 8) == (2, 904)


def test_pack_unpack_layout():
    W, M, Nl = 4, 3, 5
    obj = torch.arange(M * W * Nl, dtype=torch.float64).reshape(M, W * Nl)
    con = -obj
    gathered = torch.stack([sedist.pack_costs(obj[:, r * Nl:(r + 1) * Nl], con[:, r * Nl:(r + 1) * Nl])
                            for r in range(W)])
    o2, c2 = sedist.unpack_costs(gathered, W, M, Nl)
    assert torch.equal(o2, obj) and torch.equal(c2, con)


def _problem(w):
    gp = oracle.OracleGpCemSSM(w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    return oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                 w.act_b)


def _worker(rank, world, port, out_dir):
    torch.set_default_dtype(torch.double)
    torch.set_num_threads(1)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        w = workloads.pendulum_workload(n_train=60, num_rollouts=64, time_horizon=3, num_elites=8, num_iterations=3)
        prob = _problem(w)
        first, count = sedist.shard_range(w.num_rollouts, rank, world)
        seed, tnu = 99, w.time_horizon * w.n_u
        means, stds = torch.zeros(w.time_horizon, w.n_u), torch.ones(w.time_horizon, w.n_u)
        hist = []
        for it in range(w.num_iterations):
            # this rank draws ONLY its shard, keyed by global candidate id
            eps = T(oracle.philox_normal_noise(seed, it, first, count, w.time_horizon, w.n_u))
            actions = means + stds * eps
            r = oracle.perform_rollouts(prob, T(w.p0), actions)
            send = sedist.pack_costs(r.objective_costs.unsqueeze(0), r.constraint_costs.unsqueeze(0))
            recv = [torch.empty_like(send) for _ in range(world)]
            dist.all_gather(recv, send)  # the one collective of the path
            obj, con = sedist.unpack_costs(torch.stack(recv), world, 1, count)
            ids = oracle.select_elites(obj[0], con[0], w.num_elites)
            # elites may live on another rank: regenerate their action sequences from the RNG instead of gathering
            el = torch.stack([means + stds * T(oracle.philox_normal_noise(seed, it, int(g), 1, w.time_horizon, w.n_u))[0]
                              for g in ids])
            means, stds = el.mean(dim=0), el.std(dim=0)
            hist.append((ids.clone(), means.clone(), stds.clone()))
        torch.save(hist, os.path.join(out_dir, f'rank{rank}.pt'))
    finally:
        dist.destroy_process_group()


def test_population_sharding_world2_matches_single_process(tmp_path):
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    h0 = torch.load(os.path.join(tmp_path, 'rank0.pt'))
    h1 = torch.load(os.path.join(tmp_path, 'rank1.pt'))
    w = workloads.pendulum_workload(n_train=60, num_rollouts=64, time_horizon=3, num_elites=8, num_iterations=3)
    noise = np.stack([oracle.philox_normal_noise(99, i, 0, 64, 3, 1) for i in range(3)])
    ref = oracle.cem_solve(_problem(w), T(w.p0), T(noise), w.num_elites)
    for it in range(3):
        for h in (h0, h1):  # every rank takes the same decision, equal to the unsharded solve
            assert torch.equal(h[it][0], ref['elite_ids'][it])
            assert torch.allclose(h[it][1], ref['means'][it], rtol=1e-12, atol=1e-15)
            assert torch.allclose(h[it][2], ref['stds'][it], rtol=1e-12, atol=1e-15)


# ---- instance sharding (BASELINE config C4): split + gather, world_size 2 over gloo ---------------------------------------
class _FakeSolver:
    """Stands for a fused B200ConstrainedCemMpc in the host-side split / gather logic: its "plan" for an instance is a
    function of the instance's state and GLOBAL index, so a wrong split, a wrong offset or a wrong gather order shows."""

    class _Ssm:
        device, dtype = torch.device('cpu'), torch.float64

    def __init__(self):
        self._ssm, self._time_horizon, self._action_dimen = self._Ssm(), 3, 1
        self.first_instance = 0

    def solve(self, p0, seed=None):
        m = p0.shape[0]
        gid = torch.arange(self.first_instance, self.first_instance + m, dtype=torch.float64)
        plan = (p0.sum(dim=1) + 100.0 * gid + seed).view(m, 1, 1).expand(m, 3, 1).clone()
        return dict(actions=plan, found=(gid.long() % 2 == 0).to(torch.int32))


def _instance_worker(rank, world, port, out_dir):
    torch.set_default_dtype(torch.double)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        p0_all = torch.arange(10, dtype=torch.float64).view(5, 2)
        plans, found = sedist.solve_instances_sharded(_FakeSolver(), p0_all, seed=7)
        torch.save((plans, found), os.path.join(out_dir, f'inst{rank}.pt'))
    finally:
        dist.destroy_process_group()


def test_instance_sharding_world2_split_and_gather(tmp_path):
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    mp.spawn(_instance_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    p0_all = torch.arange(10, dtype=torch.float64).view(5, 2)
    want = _FakeSolver().solve(p0_all, seed=7)  # all five instances on one "GPU"
    for rank in range(2):  # 3 + 2 instances; every rank ends up with all results in global order
        plans, found = torch.load(os.path.join(tmp_path, f'inst{rank}.pt'))
        assert torch.equal(plans, want['actions']) and torch.equal(found, want['found'].bool())
