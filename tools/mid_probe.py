"""Solve time of mid-size models (between the fused small-model rollout and the throughput regime).

    python tools/mid_probe.py [n_train N T] ...
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from safe_exploration_b200 import workloads

DEV = 'cuda:0'
cases = [(200, 500, 5), (400, 500, 5), (400, 2048, 10), (1000, 2048, 10)]
if len(sys.argv) > 3:
    cases = [tuple(int(v) for v in sys.argv[1:4])]
for n, N, T in cases:
    w = workloads.pendulum_workload(n_train=n, num_rollouts=N, time_horizon=T, num_elites=max(3, N // 32))
    ssm, mpc = workloads.build_solver(w, device=DEV, record='none', seed=3)
    p0 = torch.as_tensor(w.p0, device=DEV).reshape(1, w.n_s)
    for _ in range(3):
        mpc.solve(p0, sync=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        res = mpc.solve(p0, sync=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f'n={n} N={N} T={T}: {ms:.3f} ms per solve, {ms * 1e3 / (8 * T):.1f} us per horizon step, '
          f'{ssm.gemm_info()}', flush=True)
