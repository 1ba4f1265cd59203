"""Synthetic workloads of the BASELINE.json configs: environment constants restated from the reference's
simulators (which are out of scope themselves) plus seeded synthetic GP training sets of the named shapes.

Host-side setup only (NumPy / SciPy); nothing here is on the hot path.
"""
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import scipy.signal
import scipy.spatial

from .safempc import dlqr


@dataclass
class Workload:
    name: str
    n_s: int
    n_u: int
    x_train: np.ndarray  # [n, n_s + n_u]
    y_train: np.ndarray  # [n, n_s]  (residual of the prior linear model)
    lengthscale: np.ndarray  # [n_s, d]
    outputscale: np.ndarray  # [n_s]
    noise: np.ndarray  # [n_s]
    a: np.ndarray
    b: np.ndarray
    k_fb: np.ndarray
    l_mu: np.ndarray
    l_sigma: np.ndarray
    h_mat: np.ndarray
    h_vec: np.ndarray
    act_a: np.ndarray
    act_b: np.ndarray
    beta_safety: float
    time_horizon: int
    num_rollouts: int
    num_elites: int
    num_iterations: int
    p0: np.ndarray
    meta: dict = field(default_factory=dict)

    @property
    def n_train(self) -> int:
        return self.x_train.shape[0]

    def flop_per_candidate_step(self) -> float:
        """F_step of SURVEY.md 8(d): n_s [n^2 + n (5 d + 8)]."""
        n, d = self.n_train, self.n_s + self.n_u
        return self.n_s * (n * n + n * (5 * d + 8))


def box_polytope(lo, hi):
    lo, hi = np.atleast_1d(lo).astype(float), np.atleast_1d(hi).astype(float)
    n = lo.size
    return np.vstack((np.eye(n), -np.eye(n))), np.concatenate((hi, -lo))


# ---- InvertedPendulum (safe_exploration/environments/environments.py:405-847) ----------------------------------
def pendulum_linear_model(m=0.15, l=0.5, g=9.82, b=0.0, dt=0.05):
    """linearize_discretize at the origin (:315-353) of the dynamics jacobian (:543-563); states (d_theta, theta)."""
    inertia = m * l ** 2
    a_ct = np.array([[b / inertia, g / l], [1.0, 0.0]])
    b_ct = np.array([[1.0 / inertia], [0.0]])
    a_d, b_d, _, _, _ = scipy.signal.cont2discrete((a_ct, b_ct, np.eye(2), np.zeros((2, 1))), dt)
    return a_d, b_d


def pendulum_polytope(simple_constraints=False):
    """_init_safety_constraints (:779-831): convex hull of four corners -> h_mat x <= h_vec (4 faces)."""
    max_rad, max_dtheta, max_dtheta_theta_0 = np.deg2rad(20), 1.2, 0.8
    if simple_constraints:
        corners = np.array([[-max_dtheta_theta_0, max_rad], [max_dtheta_theta_0, max_rad],
                            [max_dtheta_theta_0, -max_rad], [-max_dtheta_theta_0, -max_rad]])
    else:
        corners = np.array([[-max_dtheta, max_rad], [max_dtheta_theta_0, 0.0], [max_dtheta, -max_rad],
                            [-max_dtheta_theta_0, 0.0]])
    eq = scipy.spatial.ConvexHull(corners).equations
    return eq[:, :2].copy(), -eq[:, 2].copy()


# ---- CartPole (environments.py:866-1180) -------------------------------------------------------------------------
def cartpole_norms(l=0.5, g=9.82, u_min=-4.0, u_max=4.0):
    norm_x = np.array([3.0, 5.0, np.deg2rad(30), 2 * np.sqrt(g / l)])
    norm_u = np.array([u_max - u_min])
    return norm_x, norm_u


def cartpole_linear_model(l=0.5, m=0.5, M=0.5, b=0.1, g=9.82, dt=0.1):
    """_jac_dynamics (:1043-1060) + normalised linearize_discretize (:315-353)."""
    a_ct = np.array([[0, 1, 0, 0], [0, 0, .5 * g * m / M, -b * .5 / (M * l)], [0, 0, 0, 1],
                     [0, 0, g * (m + M) / (l * M), -b * (m + M) / (m * M * l ** 2)]], dtype=float)
    b_ct = np.array([0, 1. / M, 0, 1 / (M * l)], dtype=float).reshape(-1, 1)
    norm_x, norm_u = cartpole_norms(l, g)
    a_ct = np.diag(1 / norm_x) @ a_ct @ np.diag(norm_x)
    b_ct = np.diag(1 / norm_x) @ b_ct @ np.diag(norm_u)
    a_d, b_d, _, _, _ = scipy.signal.cont2discrete((a_ct, b_ct, np.eye(4), np.zeros((4, 1))), dt)
    return a_d, b_d


def cartpole_polytope():
    """_init_safety_constraints (:1062-1116) with normalisation (:1148-1176): 9 faces in normalised coordinates."""
    rows = [[0., 0., 7.25, 1.], [0., 0., -7.25, -1.], [0., 0., -1.25, -1.], [0., 0., 1.25, 1.], [0., 1., 0., 0.],
            [0., -1., 0., 0.], [1., 0., 0., 0.], [-1., 0., 0., 0.], [1., 2.0, 0., 0.]]
    h_vec = np.array([1., 1., 1., 1., 1.66, 1.66, 2.6, 4.0, 3.0])
    norm_x, _ = cartpole_norms()
    return np.array(rows) @ np.diag(norm_x), h_vec


def cartpole_obstacle_polytope():
    """The obstacle polytope of _init_safety_constraints (:1118-1139), normalised like the safety polytope: cart
    position in [-10, 3], pole angle within +-90 degrees -- plus the row the reference stacks in between by accident
    (`h_1_mat`, `h_1_vec` still hold the SAFETY face -7.25 theta - dtheta <= 1 at that point of the function)."""
    rows = [[1., 0., 0., 0.], [-1., 0., 0., 0.], [0., 0., -7.25, -1.], [0., 0., 1., 0.], [0., 0., -1., 0.]]
    h_vec = np.array([3.0, 10.0, 1.0, np.deg2rad(90), np.deg2rad(90)])
    norm_x, _ = cartpole_norms()
    return np.array(rows) @ np.diag(norm_x), h_vec


# ---- synthetic GP data ------------------------------------------------------------------------------------------------
def _synthetic_targets(x: np.ndarray, n_s: int, rng: np.random.Generator, scale: float) -> np.ndarray:
    """A fixed smooth residual-dynamics function + observation noise."""
    d = x.shape[1]
    y = np.empty((x.shape[0], n_s))
    for o in range(n_s):
        w1 = np.cos(0.7 * (o + 1) * np.arange(1, d + 1))
        w2 = np.sin(1.3 * (o + 1) * np.arange(1, d + 1))
        y[:, o] = scale * (np.sin(x @ w1 * 1.5) + 0.5 * (x @ w2) * x[:, -1])
    return y + 0.01 * scale * rng.standard_normal(y.shape)


def pendulum_workload(n_train=2000, time_horizon=10, num_rollouts=16384, num_elites=None, num_iterations=8, seed=1,
                      name='C2') -> Workload:
    """C1 / C2 of SURVEY.md 8(d): InvertedPendulum safe CEM-MPC with an exact GP on n_train points."""
    rng = np.random.default_rng(seed)
    a, b = pendulum_linear_model()
    k_lqr, _, _ = dlqr(a, b, np.diag([1., 2.]), 25.0 * np.eye(1))  # episodic_pendulum_cem.py:57-58
    h_mat, h_vec = pendulum_polytope(False)
    lo, hi = np.array([-1.2, -0.4, -1.0]), np.array([1.2, 0.4, 1.0])
    x = rng.uniform(lo, hi, size=(n_train, 3))
    y = _synthetic_targets(x / hi, 2, rng, 0.05)
    act_a, act_b = box_polytope([-1.0], [1.0])
    return Workload(
        name=name, n_s=2, n_u=1, x_train=x, y_train=y,
        lengthscale=np.array([[1.2, 0.4, 1.0], [1.0, 0.5, 1.2]]), outputscale=np.array([0.05, 0.04]) ** 2 * 4,
        noise=np.array([1e-5, 1e-5]), a=a, b=b, k_fb=-k_lqr, l_mu=np.array([0.05, 0.02]),
        l_sigma=np.array([0.05, 0.02]), h_mat=h_mat, h_vec=h_vec, act_a=act_a, act_b=act_b, beta_safety=2.0,
        time_horizon=time_horizon, num_rollouts=num_rollouts,
        num_elites=num_elites if num_elites is not None else max(3, num_rollouts // 64),
        num_iterations=num_iterations, p0=np.zeros(2), meta=dict(env='InvertedPendulum', seed=seed))


def cartpole_workload(n_train=5000, time_horizon=15, num_rollouts=65536, num_elites=None, num_iterations=8, seed=2,
                      name='C3', lipschitz=0.01) -> Workload:
    """C3: CartPole (4 states) safe CEM-MPC, GP on n_train points.

    l_mu = l_sigma = 0.01 instead of the environment's 0.05 (environments.py:936-941): with 0.05 the remainder term
    l_mu * lambda_max(q (I + k_fb^T k_fb)) squares the ellipsoid every step and the shape matrices of the initial
    (std = 1) candidates overflow float64 around step 11 of 15 -- in the reference's own arithmetic too (the oracle
    fails the same way), where `_fix_zeros_nans` then raises.  The reference runs this environment with horizons 2-5."""
    rng = np.random.default_rng(seed)
    a, b = cartpole_linear_model()
    k_lqr, _, _ = dlqr(a, b, np.diag([2., 6., 12., 4.]), 40.0 * np.eye(1))  # defaultconfig_episode.py:69-71
    h_mat, h_vec = cartpole_polytope()
    hi = np.array([0.6, 0.3, 0.25, 0.25, 0.5])
    x = rng.uniform(-hi, hi, size=(n_train, 5))
    y = _synthetic_targets(x / hi, 4, rng, 0.02)
    act_a, act_b = box_polytope([-0.5], [0.5])  # u_min_norm, u_max_norm (environments.py:122-128)
    ls = np.tile(np.array([0.6, 0.35, 0.3, 0.3, 0.6]), (4, 1)) * np.array([[1.0], [1.1], [0.9], [1.2]])
    return Workload(
        name=name, n_s=4, n_u=1, x_train=x, y_train=y, lengthscale=ls, outputscale=np.full(4, 0.02 ** 2 * 4),
        noise=np.full(4, 2e-6), a=a, b=b, k_fb=-k_lqr, l_mu=np.full(4, lipschitz), l_sigma=np.full(4, lipschitz), h_mat=h_mat,
        h_vec=h_vec, act_a=act_a, act_b=act_b, beta_safety=2.0, time_horizon=time_horizon, num_rollouts=num_rollouts,
        num_elites=num_elites if num_elites is not None else max(3, num_rollouts // 64),
        num_iterations=num_iterations, p0=np.zeros(4), meta=dict(env='CartPole', seed=seed))


def reach_microbench_workload(n_train=1024, n_s=2, num_ellipsoids=1 << 20, time_horizon=1, seed=4) -> Workload:
    """C5: multi-step ellipsoid propagation, n_train 256-16k, state dim 2-6, n_u = 1."""
    rng = np.random.default_rng(seed + 31 * n_s + n_train)
    d = n_s + 1
    x = rng.uniform(-1, 1, size=(n_train, d))
    y = _synthetic_targets(x, n_s, rng, 0.05)
    m = rng.standard_normal((n_s, n_s))
    a = 0.9 * m / np.abs(np.linalg.eigvals(m)).max()
    b = 0.1 * rng.standard_normal((n_s, 1))
    k_fb = rng.uniform(-1, 0, size=(1, n_s))
    h_mat, h_vec = box_polytope(-np.ones(n_s), np.ones(n_s))
    act_a, act_b = box_polytope([-1.0], [1.0])
    return Workload(
        name=f'C5-n{n_train}-ns{n_s}', n_s=n_s, n_u=1, x_train=x, y_train=y,
        lengthscale=np.full((n_s, d), 0.8), outputscale=np.full(n_s, 0.01), noise=np.full(n_s, 1e-5), a=a, b=b,
        k_fb=k_fb, l_mu=np.full(n_s, 0.05), l_sigma=np.full(n_s, 0.05), h_mat=h_mat, h_vec=h_vec, act_a=act_a,
        act_b=act_b, beta_safety=2.0, time_horizon=time_horizon, num_rollouts=num_ellipsoids, num_elites=3,
        num_iterations=1, p0=np.zeros(n_s), meta=dict(env='microbench', seed=seed))


def build_solver(w: Workload, dtype=None, device='cuda:0', record='costs', pick_best=False, seed=0,
                 num_instances=1, gemm_impl=0, rank=0, world_size=1, first_instance=0):
    """Workload -> (B200GpCemSSM, B200ConstrainedCemMpc) wired exactly like CemSafeMPC.__init__ does."""
    import torch
    from .cem import ActionConstraint, B200ConstrainedCemMpc, TorchPolytope
    from .reachability import SafeProblem
    from .safempc import B200SafeDynamics, EllipsoidTerminalConstraint
    from .ssm import B200GpCemSSM
    dtype = torch.float64 if dtype is None else dtype
    ssm = B200GpCemSSM(None, w.n_s, w.n_u, lengthscale=w.lengthscale, outputscale=w.outputscale, noise=w.noise,
                       dtype=dtype, device=device, gemm_impl=gemm_impl)
    ssm.update_model(torch.as_tensor(w.x_train, device=device), torch.as_tensor(w.y_train, device=device),
                     replace_old=True)
    problem = SafeProblem(w.n_s, w.n_u, w.l_mu, w.l_sigma, k_fb=w.k_fb, a=w.a, b=w.b, c_safety=w.beta_safety)
    constraints = [ActionConstraint(TorchPolytope((w.act_a, w.act_b))),
                   EllipsoidTerminalConstraint(w.n_s, w.h_mat, w.h_vec)]
    mpc = B200ConstrainedCemMpc(B200SafeDynamics(ssm, problem), constraints, w.n_s + w.n_s ** 2, w.n_u,
                                w.time_horizon, w.num_rollouts, w.num_elites, w.num_iterations, seed=seed,
                                pick_best=pick_best, record=record, num_instances=num_instances, rank=rank,
                                world_size=world_size, first_instance=first_instance)
    return ssm, mpc
