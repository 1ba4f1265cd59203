import sys; sys.path.insert(0,'/root/repo')
import torch
from safe_exploration_b200 import workloads
for N in (64, 125, 250, 500, 1000, 2000):
    w = workloads.pendulum_workload(n_train=100, num_rollouts=N, time_horizon=5, num_elites=max(3, N // 25), num_iterations=8)
    ssm, mpc = workloads.build_solver(w, device='cuda:0', record='none', seed=3)
    p0 = torch.as_tensor(w.p0, device='cuda:0').reshape(1, 2)
    for _ in range(3): mpc.solve(p0, sync=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): mpc.solve(p0, sync=False)
    e1.record(); torch.cuda.synchronize()
    print(N, round(e0.elapsed_time(e1) / 50 * 1000 / 8, 1), 'us per CEM iteration', flush=True)
