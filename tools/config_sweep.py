"""Measures the other BASELINE.json configs on one GPU (bench.py carries the headline config C2):

  C1  pendulum, n_train 100, T 5, N 500                      (the reference's own CPU-sized case)
  C3  cart-pole, n_train 5000, T 15, N 65536 on ONE GPU      (BASELINE shards it over 2/4/8; see bench.py --gpus)
  C4  128 independent pendulum MPC instances x N 4096         (one GPU's share of the 1024-instance config)
  C5  reachability microbench: 2^20 ellipsoids, T 1, n_train x n_s sweep

    python tools/config_sweep.py [C1 C3 C4 C5]   -> one JSON line per measurement (CUDA-event timing, 1 warm-up)
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from safe_exploration_b200 import reachability, workloads
from safe_exploration_b200.ssm import B200GpCemSSM

DEV = 'cuda:0'


def timed(fn, reps=2):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def solve_config(name, w, num_instances=1, reps=2):
    ssm, mpc = workloads.build_solver(w, device=DEV, record='none', seed=3, num_instances=num_instances)
    rng = np.random.default_rng(3)
    p0 = torch.as_tensor(w.p0 + (0.05 * rng.standard_normal((num_instances, w.n_s)) if num_instances > 1 else 0.0),
                         device=DEV).reshape(num_instances, w.n_s)
    ms, res = timed(lambda: mpc.solve(p0, sync=False), reps)
    steps = num_instances * w.num_rollouts * w.time_horizon * w.num_iterations
    impl, digits = ssm.gemm_info()
    print(json.dumps({'config': name, 'n_s': w.n_s, 'n_train': w.n_train, 'horizon': w.time_horizon,
                      'population': w.num_rollouts, 'instances': num_instances, 'iterations': w.num_iterations,
                      'solve_ms': ms, 'candidate_steps_per_s': steps / (ms * 1e-3), 'variance_gemm': impl,
                      'digit_planes': digits, 'found': int(res['found'].sum()),
                      'numeric_flags': res['flags'].cpu().tolist()}), flush=True)
    mpc._destroy()
    ssm._destroy()


def c5(n_train, n_s, T, count=1 << 20):
    w = workloads.reach_microbench_workload(n_train=n_train, n_s=n_s, num_ellipsoids=count, time_horizon=T)
    ssm = B200GpCemSSM(None, w.n_s, w.n_u, lengthscale=w.lengthscale, outputscale=w.outputscale, noise=w.noise, device=DEV)
    ssm.update_model(torch.as_tensor(w.x_train, device=DEV), torch.as_tensor(w.y_train, device=DEV), replace_old=True)
    g = torch.Generator(device='cpu').manual_seed(4)
    p = (0.1 * torch.randn(count, n_s, generator=g, dtype=torch.float64)).to(DEV)
    a = torch.randn(count, n_s, n_s, generator=g, dtype=torch.float64)
    q = (0.2 * (a @ a.transpose(1, 2) / n_s + 0.1 * torch.eye(n_s, dtype=torch.float64))).to(DEV)
    u = (torch.rand(count, T, 1, generator=g, dtype=torch.float64) * 2 - 1).to(DEV)
    del a
    T_ = lambda x: torch.as_tensor(x, device=DEV)
    fn = lambda: reachability.multistep_reachability(p, ssm, T_(w.k_fb), u, T_(w.l_mu), T_(w.l_sigma), q_0=q,
                                                     c_safety=w.beta_safety, a=T_(w.a), b=T_(w.b), return_all=False)
    ms, _ = timed(fn, 1)
    impl, digits = ssm.gemm_info()
    print(json.dumps({'config': 'C5', 'n_s': n_s, 'n_train': n_train, 'horizon': T, 'ellipsoids': count,
                      'ms': ms, 'candidate_steps_per_s': count * T / (ms * 1e-3), 'variance_gemm': impl,
                      'digit_planes': digits}), flush=True)
    ssm._destroy()


def main():
    which = sys.argv[1:] or ['C1', 'C3', 'C4', 'C5']
    if 'C1' in which:
        solve_config('C1', workloads.pendulum_workload(n_train=100, num_rollouts=500, time_horizon=5, num_elites=20,
                                                       name='C1'), reps=20)
    if 'C3' in which:
        solve_config('C3 (whole population on one GPU)', workloads.cartpole_workload(), reps=1)
    if 'C4' in which:
        solve_config('C4 (128 of the 1024 instances)', workloads.pendulum_workload(num_rollouts=4096, num_elites=64),
                     num_instances=128, reps=1)
    if 'C5' in which:
        for n_train, n_s in ((256, 2), (1024, 2), (4096, 2), (16384, 2), (1024, 4), (4096, 4), (1024, 6), (4096, 6)):
            c5(n_train, n_s, 1)
        c5(1024, 2, 10)


if __name__ == '__main__':
    t0 = time.time()
    main()
    print(f'# done in {time.time() - t0:.0f} s', flush=True)
