"""torchrun --nproc-per-node W tools/multigpu_parity.py : population-sharded solve over W GPUs == single-GPU solve.

Every rank solves the same MPC problem with its shard of the population (one ncclAllGather of candidate costs per CEM
iteration, issued by libsecem.so); rank 0 also solves it unsharded.  Elite ids, refitted means/stds and the returned
action sequence must be identical (bit-exact integers, 1e-12 floats: only fp summation order inside one kernel)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from safe_exploration_b200 import dist as sedist, workloads

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = f'cuda:{local}'
dist.init_process_group('nccl', device_id=torch.device(dev))
ok = True
for (n_train, N, T, E, I) in [(100, 512, 5, 16, 4), (300, 4096, 6, 64, 3)]:
    w = workloads.pendulum_workload(n_train=n_train, num_rollouts=N, time_horizon=T, num_elites=E, num_iterations=I)
    ssm, mpc = workloads.build_solver(w, device=dev, record='costs', seed=5)
    comm = sedist.attach_population_sharding(mpc)
    p0 = torch.as_tensor(w.p0, device=dev).unsqueeze(0)
    res = mpc.solve(p0, seed=5)
    torch.cuda.synchronize()
    # gather every rank's view of the decisions
    views = [torch.empty_like(res['elite_ids']) for _ in range(world)]
    dist.all_gather(views, res['elite_ids'].contiguous())
    same = all(torch.equal(v, views[0]) for v in views)
    if rank == 0:
        ssm1, mpc1 = workloads.build_solver(w, device=dev, record='costs', seed=5)
        ref = mpc1.solve(p0, seed=5)
        eq_ids = torch.equal(ref['elite_ids'], res['elite_ids'])
        eq_mean = torch.allclose(ref['means'], res['means'], rtol=1e-12, atol=1e-15)
        eq_act = torch.allclose(ref['actions'], res['actions'], rtol=1e-12, atol=1e-15) and \
            torch.equal(ref['best_id'], res['best_id'])
        n_l = N // world
        eq_cost = torch.equal(ref['constraint'][:, :, :n_l], res['constraint']) and \
            torch.allclose(ref['objective'][:, :, :n_l], res['objective'], rtol=1e-12, atol=0)
        print(f'W={world} n={n_train} N={N}: ranks agree={same} elite_ids={eq_ids} means={eq_mean} '
              f'actions={eq_act} shard costs={eq_cost} best={int(res["best_id"][0])}', flush=True)
        ok = ok and same and eq_ids and eq_mean and eq_act and eq_cost
    mpc._destroy()
    sedist.destroy_nccl_comm(comm)
dist.barrier()
if rank == 0:
    print('MULTIGPU PARITY', 'PASS' if ok else 'FAIL', flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
