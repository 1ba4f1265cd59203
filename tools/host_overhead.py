"""Where does the host time of a small solve go?  cProfile of B200CemSafeMPC.get_action at C1."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import safe_exploration_b200 as se
from safe_exploration_b200 import workloads

w = workloads.pendulum_workload(n_train=100, num_rollouts=500, time_horizon=5, num_elites=20, name='C1')
ssm, mpc = workloads.build_solver(w, device='cuda:0', record='none', seed=3)


class Env:
    n_s, n_u = w.n_s, w.n_u
    l_mu, l_sigm = w.l_mu, w.l_sigma
    objective_spec = None


class Conf:
    mpc_time_horizon, cem_num_rollouts = w.time_horizon, w.num_rollouts
    cem_num_elites, cem_num_iterations = w.num_elites, w.num_iterations
    use_prior_model, cem_record = True, 'none'


class Gain:
    def get_control_matrix_pytorch(self):
        return torch.as_tensor(w.k_fb)


solver = se.B200CemSafeMPC(ssm, mpc._constraints, Env, Conf, {'lin_model': (w.a, w.b)}, None, None, w.beta_safety,
                           safe_policy=lambda s: np.zeros(w.n_u), lqr=Gain(), mpc=mpc, verbose=False)
state = np.zeros(2)
for _ in range(20):
    solver.get_action(state)
import time
t0 = time.perf_counter()
for _ in range(500):
    solver.get_action(state)
print(f'get_action: {(time.perf_counter() - t0) / 500 * 1e3:.3f} ms per call')
pr = cProfile.Profile()
pr.enable()
for _ in range(500):
    solver.get_action(state)
pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(22)
