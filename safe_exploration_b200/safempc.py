"""`B200CemSafeMPC`: drop-in for the reference's `CemSafeMPC` (safe_exploration/safempc_cem.py:159-344) -- same
constructor arguments, same `get_action` state machine (FOUND_SOLUTION -> PREVIOUS_SOLUTION -> SAFE_CONTROLLER), same
`update_model` prior subtraction -- with the optimisation delegated to `B200ConstrainedCemMpc`.
"""
from abc import ABC, abstractmethod
from enum import Enum
from functools import lru_cache
from typing import Callable, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import scipy.linalg
import torch
from numpy import ndarray
from torch import Tensor

from . import _lib
from .cem import ActionConstraint, B200ConstrainedCemMpc, Constraint, DynamicsFunc, TorchPolytope, box2torchpoly
from .reachability import SafeProblem, is_ellipsoid_inside_polytope, multistep_reachability, onestep_reachability
from .ssm import B200GpCemSSM, CemSSM, assert_shape

STATE_PENALTY = 10  # per ellipsoid outside its polytope (safempc_cem.py:132)


class MpcResult(Enum):
    """safempc_cem.py:23-27."""
    FOUND_SOLUTION = 1
    PREVIOUS_SOLUTION = 2
    SAFE_CONTROLLER = 3


class SafeMPC(ABC):
    """safe_exploration/safempc.py:8-77 (the subset of abstract members CemSafeMPC implements)."""

    @property
    @abstractmethod
    def state_dimen(self) -> int:
        pass

    @property
    @abstractmethod
    def action_dimen(self) -> int:
        pass

    @abstractmethod
    def get_action(self, state: ndarray):
        pass

    @abstractmethod
    def update_model(self, x: ndarray, y: ndarray, opt_hyp=False, replace_old=True, reinitialize_solver=True) -> None:
        pass

    @abstractmethod
    def collect_metrics(self) -> Dict[str, float]:
        pass


class PQFlattener:
    """[p | vec(q)] <-> (p, q) (safempc_cem.py:30-76); an all-zero q for the whole batch means point states."""

    def __init__(self, state_dimen: int):
        self._state_dimen = state_dimen

    def flatten(self, p: Tensor, q: Optional[Tensor]) -> Tensor:
        """(p [N, n_s], q [N, n_s, n_s] | None) -> [N, n_s + n_s^2]; a missing q is written as zeros (point states)."""
        n, count = self._state_dimen, p.size(0)
        assert_shape(p, (count, n))
        if q is None:
            return torch.cat((p, p.new_zeros((count, n * n))), dim=1)
        assert_shape(q, (count, n, n))
        return torch.cat((p, q.reshape(count, n * n)), dim=1)

    def unflatten(self, flat: Tensor) -> Tuple[Tensor, Optional[Tensor]]:
        """Inverse of `flatten`; q is None when every shape matrix of the batch is exactly zero."""
        n, count = self._state_dimen, flat.size(0)
        assert_shape(flat, (count, self.get_flat_state_dimen()))
        p, q_flat = flat[:, :n], flat[:, n:]
        return p, (q_flat.reshape(count, n, n) if bool(q_flat.ne(0).any()) else None)

    def get_flat_state_dimen(self):
        n = self._state_dimen
        return n * (n + 1)


def _outside_cost(flattener: 'PQFlattener', states: Tensor, polytope_a: Tensor, polytope_b: Tensor):
    """STATE_PENALTY for every ellipsoid of `states` [K, n_s + n_s^2] that is not inside the polytope
    (safempc_cem.py:127-132)."""
    if states.shape[0] == 0:
        return 0
    p, q = flattener.unflatten(states)
    inside = is_ellipsoid_inside_polytope(p, q, polytope_a.to(p), polytope_b.to(p))
    return (~inside.bool()).sum() * STATE_PENALTY


class EllipsoidStateConstraint(Constraint):
    """safempc_cem.py:116-132.  NOTE (quirk Q1): like the reference, only the LAST ellipsoid of the trajectory is
    checked; pass `all_states=True` for the intended all-states semantics."""

    def __init__(self, state_dimen: int, safe_polytope_a: np.ndarray, safe_polytope_b: np.ndarray, device: str = None,
                 all_states: bool = False):
        self._pq_flattener = PQFlattener(state_dimen)
        self._polytope_a = torch.as_tensor(np.asarray(safe_polytope_a, dtype=np.float64), device=device)
        self._polytope_b = torch.as_tensor(np.asarray(safe_polytope_b, dtype=np.float64), device=device).reshape(-1, 1)
        self.all_states = all_states

    def __call__(self, trajectory, actions) -> float:
        checked = trajectory[1:] if self.all_states else trajectory[-1:]
        return _outside_cost(self._pq_flattener, checked, self._polytope_a, self._polytope_b)

    def batch_call(self, trajectories: Tensor, actions: Tensor) -> Tensor:
        """All rollouts at once (interface path of B200ConstrainedCemMpc): trajectories [N, T+1, S] -> costs [N]."""
        return _batch_inside_cost(self._pq_flattener, trajectories[:, 1:] if self.all_states else trajectories[:, -1:],
                                  self._polytope_a, self._polytope_b)


class EllipsoidTerminalConstraint(Constraint):
    """safempc_cem.py:102-113."""

    def __init__(self, state_dimen: int, safe_polytope_a: np.ndarray, safe_polytope_b: np.ndarray, device: str = None):
        self._constraint = EllipsoidStateConstraint(state_dimen, safe_polytope_a, safe_polytope_b, device)
        self._polytope_a = self._constraint._polytope_a
        self._polytope_b = self._constraint._polytope_b

    def __call__(self, trajectory, actions) -> float:
        return self._constraint(trajectory[-1:], actions[-1:])  # the wrapped constraint looks at the last slot

    def batch_call(self, trajectories: Tensor, actions: Tensor) -> Tensor:
        return self._constraint.batch_call(trajectories, actions)


class EllipsoidObstacleConstraint(Constraint):
    """The intermediate-state constraint of the reference's CasADi solver (safempc_simple.py:373-383): every ellipsoid
    BEFORE the terminal one must lie inside the obstacle polytope h_mat_obs x <= h_obs.  The reference's CEM path
    reads that polytope and drops it (safempc_cem.py:137); here it is an opt-in constraint, 10 per violating state
    like the other state constraints."""

    def __init__(self, state_dimen: int, obstacle_polytope_a: np.ndarray, obstacle_polytope_b: np.ndarray,
                 device: str = None):
        self._pq_flattener = PQFlattener(state_dimen)
        self._polytope_a = torch.as_tensor(np.asarray(obstacle_polytope_a, dtype=np.float64), device=device)
        self._polytope_b = torch.as_tensor(np.asarray(obstacle_polytope_b, dtype=np.float64),
                                           device=device).reshape(-1, 1)

    def __call__(self, trajectory, actions) -> float:
        # slot 0 is the (point) initial state, the last slot the terminal ellipsoid
        return _outside_cost(self._pq_flattener, trajectory[1:-1], self._polytope_a, self._polytope_b)

    def batch_call(self, trajectories: Tensor, actions: Tensor) -> Tensor:
        if trajectories.shape[1] <= 2:
            return torch.zeros((trajectories.shape[0],), dtype=trajectories.dtype, device=trajectories.device)
        return _batch_inside_cost(self._pq_flattener, trajectories[:, 1:-1], self._polytope_a, self._polytope_b)


def _batch_inside_cost(flattener: PQFlattener, states: Tensor, polytope_a: Tensor, polytope_b: Tensor) -> Tensor:
    """10 x #{ellipsoids of each rollout not strictly inside the polytope}: states [N, K, S] -> [N]."""
    N, K, S = states.shape
    n_s = flattener._state_dimen
    flat = states.reshape(N * K, S)
    p = flat[:, :n_s].contiguous()
    q = flat[:, n_s:].reshape(N * K, n_s, n_s).contiguous()
    inside = is_ellipsoid_inside_polytope(p, q, polytope_a.to(p), polytope_b.to(p)).reshape(N, K)
    return float(STATE_PENALTY) * (K - inside.sum(dim=1)).to(states.dtype)


def construct_constraints(conf, env) -> List[Constraint]:
    """safempc_cem.py:135-146: [ActionConstraint(normalised action box), Ellipsoid{State|Terminal}Constraint]."""
    h_mat_safe, h_safe, h_mat_obs, h_obs = env.get_safety_constraints(normalize=True)
    boxes = np.array([np.array(xs) for xs in zip(env.u_min_norm, env.u_max_norm)])
    action_constraint = ActionConstraint(box2torchpoly(boxes))
    if conf.use_state_constraint:
        constraint = EllipsoidStateConstraint(env.n_s, h_mat_safe, h_safe)
    else:
        constraint = EllipsoidTerminalConstraint(env.n_s, h_mat_safe, h_safe)
    constraints = [action_constraint, constraint]
    # opt-in (no such key in the reference's configs): pose the obstacle constraint the CasADi solvers pose
    if getattr(conf, 'use_obstacle_constraint', False) and h_mat_obs is not None:
        constraints.append(EllipsoidObstacleConstraint(env.n_s, h_mat_obs, h_obs))
    return constraints


class B200SafeDynamics(DynamicsFunc):
    """The dynamics plug-in of CemSafeMPC (safempc_cem.py:288-302) as an object the fused kernels understand: it holds
    the GP and the constants of one ellipsoid step.  Calling it performs one batched step on the device, so it also
    works as a plain `DynamicsFunc`."""

    def __init__(self, ssm: CemSSM, problem: SafeProblem, objective_fn: Optional[Callable] = None):
        """:param ssm: a B200GpCemSSM (evaluated inside the kernels) or any other CemSSM (its predict_* methods are
                       called per step and the ellipsoid algebra runs on the device: the interface path)
        :param objective_fn: optional `env.objective_cost_function` (safempc_cem.py:304-312): called per step with the
                       new centres [N, n_s]; a `None` return selects the default -sum(variance)"""
        self.ssm = ssm
        self.problem = problem
        self.objective_fn = objective_fn
        self._pq = PQFlattener(ssm.num_states)

    @property
    def fusable(self) -> bool:
        """Can the whole rollout run inside se_cem_solve?  Needs the B200 GP and a built-in objective selector."""
        return isinstance(self.ssm, B200GpCemSSM) and self.objective_fn is None

    def _objective_cost(self, p_next: Tensor, sigma: Tensor) -> Tensor:
        """safempc_cem.py:304-312."""
        pr = self.problem
        if self.objective_fn is not None:
            cost = self.objective_fn(p_next)
            return -torch.sum(sigma, dim=1) if cost is None else cost
        if pr.objective == _lib.SE_OBJ_NEG_VARIANCE:
            return -torch.sum(sigma, dim=1)
        if pr.objective == _lib.SE_OBJ_ABS_STATE1:
            return torch.abs(pr.objective_theta - p_next[:, 1])
        return -p_next[:, -1]

    def __call__(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        pr = self.problem
        ps, qs = self._pq.unflatten(states)
        dev, dt = states.device, states.dtype
        t = lambda x: torch.as_tensor(x, dtype=dt, device=dev)
        p_next, q_next, sigma = onestep_reachability(ps.contiguous(), self.ssm, actions, t(pr.l_mu), t(pr.l_sigma), qs,
                                                     k_fb=t(pr.k_fb), a=t(pr.a), b=t(pr.b), verbose=0,
                                                     c_safety=pr.c_safety)
        return self._pq.flatten(p_next, q_next), self._objective_cost(p_next, sigma)

    def rollout_batch(self, initial_states: Tensor, actions: Tensor):
        """All T steps of N rollouts in ONE device call (`se_reach_multistep`) when the model is the B200 GP and the
        rollouts start from point states -- the case of a callable `env.objective_cost_function`: the objective is then
        evaluated per step on the recorded centres.  -> (trajectories [N, T+1, S], objective_costs [N]), or None when
        the per-step `__call__` has to be used."""
        if not isinstance(self.ssm, B200GpCemSSM):
            return None
        ps, qs = self._pq.unflatten(initial_states)
        if qs is not None:
            return None
        pr = self.problem
        N, T = actions.shape[0], actions.shape[1]
        _, _, p_all, q_all, s_all = multistep_reachability(ps.contiguous(), self.ssm, pr.k_fb, actions, pr.l_mu,
                                                           pr.l_sigma, None, c_safety=pr.c_safety, a=pr.a, b=pr.b)
        n_s = self.ssm.num_states
        traj = torch.empty((N, T + 1, n_s + n_s * n_s), dtype=p_all.dtype, device=p_all.device)
        traj[:, 0] = self._pq.flatten(ps.to(p_all.dtype), None)
        traj[:, 1:, :n_s] = p_all
        traj[:, 1:, n_s:] = q_all.reshape(N, T, n_s * n_s)
        cost = torch.zeros((N,), dtype=p_all.dtype, device=p_all.device)
        for t in range(T):
            cost += self._objective_cost(p_all[:, t], s_all[:, t]).to(cost)
        return traj, cost


def can_fold(constraint) -> bool:
    """Is this one of the constraint classes the fused kernels evaluate themselves?"""
    return isinstance(constraint, (ActionConstraint, EllipsoidTerminalConstraint, EllipsoidStateConstraint,
                                   EllipsoidObstacleConstraint))


_OBJ_CALLABLE = -1  # not a built-in selector: env.objective_cost_function is called on the recorded centres


def classify_objective(env) -> Tuple[int, float, Optional[Callable]]:
    """Which objective does `env` ask for (safempc_cem.py:304-312)?  -> (selector, theta, callable | None).

    An explicit `env.objective_spec = (kind, theta)` wins.  Otherwise `env.objective_cost_function` is probed on a few
    fixed centres: `None` selects the default -sum(variance); a function that reproduces one of the reference's two
    environment objectives exactly -- |theta - p[:, 1]| (InvertedPendulum, environments.py:505-510) or -p[:, -1]
    (LunarLander, lunarlander.py:111-113) -- selects the kernel's built-in form of it (theta is read off the probe,
    and re-read before every solve, so an advancing target is followed); anything else stays a Python callable and
    the solver evaluates it per step on the recorded centres (interface path)."""
    names = {'neg_variance': _lib.SE_OBJ_NEG_VARIANCE, 'abs_state1': _lib.SE_OBJ_ABS_STATE1,
             'neg_last_state': _lib.SE_OBJ_NEG_LAST_STATE}
    spec = getattr(env, 'objective_spec', None)
    if spec is not None:
        kind, theta = spec
        return names[kind], float(theta), None
    fn = getattr(env, 'objective_cost_function', None)
    if fn is None:
        return _lib.SE_OBJ_NEG_VARIANCE, 0.0, None
    n_s = env.n_s
    gen = torch.Generator().manual_seed(20240229)
    probe = torch.randn((6, n_s), generator=gen, dtype=torch.float64)
    probe[0] = 0.0
    try:
        out = fn(probe)
    except Exception:  # e.g. a function that insists on its own device: treat it as opaque
        return _OBJ_CALLABLE, 0.0, fn
    if out is None:
        return _lib.SE_OBJ_NEG_VARIANCE, 0.0, None
    out = torch.as_tensor(out, dtype=torch.float64).reshape(-1)
    if out.shape[0] == probe.shape[0]:
        if torch.equal(out, -probe[:, -1]):
            return _lib.SE_OBJ_NEG_LAST_STATE, 0.0, None
        if n_s >= 2:
            for theta in (float(out[0]), -float(out[0])):  # probe[0, 1] = 0, so out[0] = |theta|
                if torch.equal(out, torch.abs(theta - probe[:, 1])):
                    return _lib.SE_OBJ_ABS_STATE1, theta, None
    return _OBJ_CALLABLE, 0.0, fn


def fold_constraints(problem: SafeProblem, constraints: Sequence[Constraint]) -> SafeProblem:
    """Translate the reference's constraint objects into the device-side problem description."""
    kw = {}
    for c in constraints:
        if isinstance(c, ActionConstraint):
            kw.update(act_a=c._self_region._A.cpu().numpy(), act_b=c._self_region._b.cpu().numpy(),
                      act_penalty=float(c._penalty))
        elif isinstance(c, EllipsoidTerminalConstraint):
            kw.update(h_mat=c._polytope_a.cpu().numpy(), h_vec=c._polytope_b.cpu().numpy().reshape(-1),
                      all_states_constraint=False)
        elif isinstance(c, EllipsoidStateConstraint):
            kw.update(h_mat=c._polytope_a.cpu().numpy(), h_vec=c._polytope_b.cpu().numpy().reshape(-1),
                      all_states_constraint=c.all_states)
        elif isinstance(c, EllipsoidObstacleConstraint):
            kw.update(h_mat_obs=c._polytope_a.cpu().numpy(), h_vec_obs=c._polytope_b.cpu().numpy().reshape(-1))
        else:
            raise TypeError(f'{type(c).__name__} cannot be evaluated inside the fused CEM kernels; supported: '
                            f'ActionConstraint, EllipsoidTerminalConstraint, EllipsoidStateConstraint, '
                            f'EllipsoidObstacleConstraint')
    return problem.with_(**kw)


def dlqr(a, b, q, r):
    """Infinite-horizon discrete LQR gain, u = -K x (safe_exploration/utils.py:23-38)."""
    a, b, q, r = (np.atleast_2d(np.asarray(m, dtype=np.float64)) for m in (a, b, q, r))
    x = scipy.linalg.solve_discrete_are(a, b, q, r)
    k = np.linalg.solve(b.T @ x @ b + r, b.T @ x @ a)
    return k, x, np.linalg.eigvals(a - b @ k)


class LqrFeedbackController:
    """safempc_simple.py:1105-1129: cached k_fb = -K_lqr."""

    def __init__(self, wx_feedback_cost, wu_feedback_cost, n_s: int, n_u: int, linearized_model_a, linearized_model_b,
                 conf=None):
        self._args = (linearized_model_a, linearized_model_b, wx_feedback_cost, wu_feedback_cost)

    @lru_cache()
    def get_control_matrix(self):
        k_lqr, _, _ = dlqr(*self._args)
        return -k_lqr

    def get_control_matrix_pytorch(self):
        return torch.as_tensor(self.get_control_matrix())


class B200CemSafeMPC(SafeMPC):
    """Safe MPC on constrained CEM, evaluated on the B200.  Constructor arguments as CemSafeMPC (safempc_cem.py:162-200);
    `env` needs n_s, n_u, l_mu, l_sigm, objective_cost_function (and optionally `objective_spec`), `conf` needs
    mpc_time_horizon, cem_num_rollouts, cem_num_elites, cem_num_iterations, use_prior_model."""

    def __init__(self, ssm: CemSSM, constraints: Sequence[Constraint], env, conf, opt_env, wx_feedback_cost,
                 wu_feedback_cost, beta_safety: float, safe_policy: Callable[[ndarray], ndarray],
                 lqr: Optional[LqrFeedbackController] = None, mpc: Optional[B200ConstrainedCemMpc] = None,
                 verbose: bool = True) -> None:
        super().__init__()
        self._env = env
        self._conf = conf
        self._constraints = list(constraints)
        self._state_dimen = env.n_s
        self._action_dimen = env.n_u
        self._pq_flattener = PQFlattener(env.n_s)
        self._ssm = ssm
        self._mpc_time_horizon = conf.mpc_time_horizon
        self._beta_safety = beta_safety
        self._safe_policy = safe_policy
        self._use_prior_model = conf.use_prior_model
        self._verbose = verbose
        a_lin, b_lin = opt_env['lin_model']
        self.lin_model = opt_env['lin_model']
        self._linearized_model_a = np.asarray(a_lin, dtype=np.float64).reshape(env.n_s, env.n_s)
        self._linearized_model_b = np.asarray(b_lin, dtype=np.float64).reshape(env.n_s, env.n_u)
        if lqr is None:
            lqr = LqrFeedbackController(wx_feedback_cost, wu_feedback_cost, env.n_s, env.n_u, self._linearized_model_a,
                                        self._linearized_model_b, conf=conf)
        self._lqr = lqr
        # safempc_cem.py:291-296: zeros (not identity) when the prior is disabled (quirk Q9)
        a = self._linearized_model_a if self._use_prior_model else np.zeros_like(self._linearized_model_a)
        b = self._linearized_model_b if self._use_prior_model else np.zeros_like(self._linearized_model_b)
        k_fb = lqr.get_control_matrix_pytorch() if hasattr(lqr, 'get_control_matrix_pytorch') else lqr
        self._base_problem = SafeProblem(env.n_s, env.n_u, env.l_mu, env.l_sigm, k_fb=k_fb, a=a, b=b,
                                         c_safety=beta_safety,
                                         clamp_nonpositive=bool(getattr(conf, 'cem_clamp_nonpositive_variance', False)))
        self._objective_class = None
        self._mpc_injected = mpc is not None
        self._mpc = mpc
        self._dynamics = None
        self._refresh_objective()
        self._last_mpc_actions = np.empty((0, self.action_dimen))
        self._mpc_actions_executed = 0

    def _refresh_objective(self) -> None:
        """safempc_cem.py:304-312 reads `env.objective_cost_function` on every step; here the objective is classified
        (classify_objective) before every solve, so a target that advances during an episode
        (environments.py:488-515) is followed: a new theta updates the device-side problem, a change of class
        (built-in <-> callable) rebuilds the optimiser."""
        kind, theta, fn = classify_objective(self._env)
        cls = (kind, fn is not None)
        if cls == self._objective_class:
            if fn is None and self._mpc is not None and getattr(self._mpc, 'fused', False):
                self._mpc.set_objective(kind, theta)
            elif self._dynamics is not None:
                self._dynamics.objective_fn = fn
            return
        self._objective_class = cls
        problem = self._base_problem.with_(objective=max(kind, 0), objective_theta=theta)
        self._dynamics = B200SafeDynamics(self._ssm, problem, objective_fn=fn)
        if self._mpc_injected:
            return
        conf, env = self._conf, self._env
        if self._mpc is not None and hasattr(self._mpc, '_destroy'):
            self._mpc._destroy()
        self._mpc = B200ConstrainedCemMpc(self._dynamics, constraints=self._constraints,
                                          state_dimen=self._pq_flattener.get_flat_state_dimen(), action_dimen=env.n_u,
                                          time_horizon=self._mpc_time_horizon, num_rollouts=conf.cem_num_rollouts,
                                          num_elites=conf.cem_num_elites, num_iterations=conf.cem_num_iterations,
                                          record=getattr(conf, 'cem_record', 'costs'))

    @property
    def state_dimen(self) -> int:
        return self._state_dimen

    @property
    def action_dimen(self) -> int:
        return self._action_dimen

    @property
    def safety_trajectory_length(self) -> int:
        return self._mpc_time_horizon

    @property
    def performance_trajectory_length(self) -> int:
        return 0

    @property
    def x_train(self) -> ndarray:
        x_train = self._ssm.x_train
        if x_train is None:
            return np.empty((0, self._state_dimen + self._action_dimen))
        return x_train.detach().cpu().numpy()

    def init_solver(self, cost_func=None):
        pass

    def get_action(self, state: ndarray) -> Tuple[ndarray, MpcResult]:
        """safempc_cem.py:231-263."""
        assert_shape(state, (self._state_dimen,))
        self._refresh_objective()
        if hasattr(self._mpc, 'get_actions_point'):
            # point state in, plan out: one small copy each way (the solver reuses its buffers, so the library replays
            # its CUDA graph of the solve)
            mpc_actions = self._mpc.get_actions_point(state)
        else:
            state_batch = torch.tensor(state, dtype=getattr(self._ssm, 'dtype', torch.float64),
                                       device=getattr(self._ssm, 'device', 'cuda:0')).unsqueeze(0)
            mpc_actions, _ = self._mpc.get_actions(self._pq_flattener.flatten(state_batch, None), reuse_outputs=True)
            mpc_actions = mpc_actions.detach().cpu().numpy() if mpc_actions is not None else mpc_actions
        action, result = self._follow_plan(mpc_actions, state)
        return action, result

    def _follow_plan(self, plan: Optional[ndarray], state: ndarray) -> Tuple[ndarray, MpcResult]:
        """The fall-back chain of safempc_cem.py:243-263: a fresh plan is adopted and its first action returned; without
        one the stored plan is followed one more step; when that is used up the safe policy takes over."""
        if plan is not None:
            self._say('Found solution')
            self._last_mpc_actions, self._mpc_actions_executed = plan, 1
            return plan[0], MpcResult.FOUND_SOLUTION
        step = self._mpc_actions_executed
        if step < self._last_mpc_actions.shape[0]:
            self._say('Using existing solution')
            self._mpc_actions_executed = step + 1
            return self._last_mpc_actions[step], MpcResult.PREVIOUS_SOLUTION
        self._say('Using safe controller')
        return self._safe_policy(state), MpcResult.SAFE_CONTROLLER

    def get_action_batch(self, states: ndarray) -> Tuple[ndarray, List[MpcResult]]:
        """M scenarios side by side: what episode_runner.do_rollout (episode_runner.py:206-231) does when it calls
        `solver.get_action(state)` once per scenario and step, as ONE device call (BASELINE config C4: independent MPC
        solves batched).  Every scenario keeps its own stored plan, so the FOUND / PREVIOUS / SAFE_CONTROLLER state
        machine of safempc_cem.py:231-263 runs per scenario exactly as in get_action.

        :param states: [M, n_s]; call it with the same M every step of a batched rollout
        :returns: actions [M, n_u], one MpcResult per scenario
        """
        states = np.asarray(states, dtype=np.float64)
        m = states.shape[0]
        assert_shape(states, (m, self._state_dimen))
        if getattr(self, '_batch_plans', None) is None or len(self._batch_plans) != m:
            self._batch_plans = [np.empty((0, self.action_dimen)) for _ in range(m)]
            self._batch_executed = [0] * m
        self._refresh_objective()
        if getattr(self._mpc, 'fused', True):
            dev_states = torch.as_tensor(states, dtype=self._ssm.dtype, device=self._ssm.device)
            plans, found = self._mpc.get_actions_batch(dev_states, reuse_outputs=True)
            plans, found = plans.detach().cpu().numpy(), found.cpu().numpy()
        else:  # interface path (Python callbacks): one optimisation per scenario, like the reference's loop
            plans = np.zeros((m, self._mpc_time_horizon, self.action_dimen))
            found = np.zeros((m,), dtype=bool)
            for i in range(m):
                plan = self._mpc.get_actions_point(states[i])
                if plan is not None:
                    plans[i], found[i] = plan, True
        actions = np.empty((m, self.action_dimen))
        results = []
        for i in range(m):
            if found[i]:
                actions[i] = plans[i, 0]
                self._batch_plans[i], self._batch_executed[i] = plans[i].copy(), 1
                results.append(MpcResult.FOUND_SOLUTION)
            elif self._batch_executed[i] < self._batch_plans[i].shape[0]:
                actions[i] = self._batch_plans[i][self._batch_executed[i]]
                self._batch_executed[i] += 1
                results.append(MpcResult.PREVIOUS_SOLUTION)
            else:
                actions[i] = self._safe_policy(states[i])
                results.append(MpcResult.SAFE_CONTROLLER)
        return actions, results

    def _say(self, msg: str) -> None:
        if self._verbose:
            print(msg)

    def get_action_verbose(self, state: ndarray):
        raise NotImplementedError

    def update_model(self, x: ndarray, y: ndarray, opt_hyp=False, replace_old=True, reinitialize_solver=True) -> None:
        """safempc_cem.py:314-327: the SSM learns the residual of the prior linear model."""
        x_s = x[:, :self.state_dimen]
        x_u = x[:, self.state_dimen:]
        y_error = y - self.eval_prior(x_s, x_u) if self._use_prior_model else y
        dev = getattr(self._ssm, 'device', 'cuda:0')
        self._ssm.update_model(torch.tensor(x, device=dev), torch.tensor(y_error, device=dev), opt_hyp, replace_old)

    def information_gain(self) -> Union[ndarray, List[None]]:
        return [None] * self.state_dimen

    def ssm_predict(self, z: ndarray) -> Tuple[ndarray, ndarray]:
        mean, sigma = self._ssm.predict_raw(torch.tensor(z, device=getattr(self._ssm, 'device', 'cuda:0')))
        return mean.detach().cpu().numpy(), sigma.detach().cpu().numpy()

    def eval_prior(self, states: ndarray, actions: ndarray):
        return np.dot(states, self._linearized_model_a.T) + np.dot(actions, self._linearized_model_b.T)

    def collect_metrics(self) -> Dict[str, float]:
        return self._ssm.collect_metrics()
