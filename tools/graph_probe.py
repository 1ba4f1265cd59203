"""How much of a small solve is launch overhead?  Times mpc.solve() eagerly and as a replayed CUDA graph (captured
with torch.cuda.graph around the same call; the solve only enqueues work, so it captures as is).

    python tools/graph_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from safe_exploration_b200 import workloads

DEV = 'cuda:0'


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, w in (('C1', workloads.pendulum_workload(n_train=100, num_rollouts=500, time_horizon=5, num_elites=20)),
                ('n=500 N=2048 T=10', workloads.pendulum_workload(n_train=500, num_rollouts=2048, time_horizon=10,
                                                                  num_elites=32))):
    ssm, mpc = workloads.build_solver(w, device=DEV, record='none', seed=3)
    p0 = torch.as_tensor(w.p0, device=DEV).reshape(1, w.n_s)
    eager = timed(lambda: mpc.solve(p0, sync=False), 50)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        mpc.solve(p0, sync=False)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        res = mpc.solve(p0, sync=False)
    replay = timed(graph.replay, 50)
    print(f'{name}: eager {eager:.3f} ms, graph replay {replay:.3f} ms, found={int(res["found"].sum())}', flush=True)
