"""GPU tests of the plug-in boundary: arbitrary `DynamicsFunc` / `Constraint` / `CemSSM` objects and an
environment's own `objective_cost_function` behind the B200 classes, the reference's own known answers for the
optimiser, model updates between solves, and the numeric flags of `_fix_zeros_nans`.

The interface path runs the reference-shaped loop (constrained_cem_mpc.py:145-280) on device tensors: callbacks are
Python, everything else (sampling, elite selection + refit, final pick, ellipsoid algebra) is the library's kernels.
"""
from unittest import mock

import numpy as np
import pytest
import torch

import safe_exploration_b200 as se
from oracle import oracle
from safe_exploration_b200 import _lib, workloads
from safe_exploration_b200.reachability import ReachabilityNumericsError

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))
G = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64)).to(device=DEV)


def rel_err(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = np.maximum(np.abs(b), 0 if scale is None else scale)
    den = np.where(den == 0, 1.0, den)
    return float(np.max(np.abs(a - b) / den)) if a.size else 0.0


# ---- the reference's own optimiser tests through the B200 class (constrained_cem_mpc_test.py) ---------------------------
class BasicDynamics(se.DynamicsFunc):
    """constrained_cem_mpc_test.py:81-88: f(s, a) = s + a with a constant objective cost per step."""

    def __init__(self, objective_cost):
        self._objective_cost = objective_cost

    def __call__(self, states, actions):
        assert states.size(0) == actions.size(0), 'Batch dimensions must match.'
        assert states.is_cuda and actions.is_cuda
        return states + actions, torch.full((states.size(0),), self._objective_cost, device=states.device)


class FakeConstraint(se.Constraint):
    """constrained_cem_mpc_test.py:71-76: an arbitrary Constraint object, evaluated per rollout."""

    def __call__(self, trajectory, actions) -> float:
        return 5.0 if trajectory[-1][0].item() == 1 and trajectory[-1][1].item() == 1 else 0.0


def test_toy_dynamics_known_answers_through_the_interface_path():
    """constrained_cem_mpc_test.py:153-227: stds = 0, means ((1,0),(0,1),(0,0)) -> trajectory (0,0),(1,0),(1,1),(1,1);
    objective 15 x 3; constraint cost 5."""
    rf = se.RolloutFunction(BasicDynamics(15.0), [FakeConstraint()], 2, 2, 3, 5)
    means = G([[1, 0], [0, 1], [0, 0]])
    rollouts = rf.perform_rollouts((G([0, 0]), means, torch.zeros((3, 2), device=DEV)))
    assert rollouts.trajectories.shape == (5, 4, 2) and rollouts.actions.shape == (5, 3, 2)
    assert rollouts.trajectories[0].cpu().tolist() == [[0, 0], [1, 0], [1, 1], [1, 1]]
    assert rollouts.actions[0].cpu().tolist() == [[1, 0], [0, 1], [0, 0]]
    assert float(rollouts.objective_costs[0]) == 15 * 3
    assert float(rollouts.constraint_costs[0]) == 5
    assert rollouts.trajectories.is_cuda and rollouts.constraint_costs.is_cuda
    # the same objects through the optimiser class: everything is infeasible (cost 5 only when the walk ends in (1,1);
    # with random candidates it is 0), so get_actions returns the first rollout's actions
    mpc = se.B200ConstrainedCemMpc(BasicDynamics(1.0), [FakeConstraint()], 2, 2, 3, 64, 8, 2, seed=3)
    assert not mpc.fused
    actions, by_time = mpc.get_actions(G([0, 0]))
    assert len(by_time) == 2 and actions is not None and actions.shape == (3, 2)
    assert torch.equal(actions, by_time[-1].actions[0])  # quirk Q3: first feasible rollout of the last iteration
    # the candidates are the library sampler's: iteration 0 is means 0, stds 1
    eps = se.sample_noise(3, 0, 1, 64, 0, 64, 3, 2, device=DEV)[0]
    assert torch.equal(by_time[0].actions, eps)


def _mock_rollouts(*rollouts):
    rf = mock.Mock()
    rf.perform_rollouts.side_effect = list(rollouts)
    return rf


def test_reference_selection_and_refit_known_answers_with_an_injected_rollout_function():
    """constrained_cem_mpc_test.py:230-300: rollouts fail the constraints -> elites by constraint cost (means 2, stds
    std([1, 3], ddof=1)); rollouts pass -> elites by objective among the feasible (means 6, stds std([10, 2], ddof=1))."""
    g = lambda x: torch.tensor(x, dtype=torch.float64, device=DEV)
    r1 = se.Rollouts(g([[[0, 0], [1, 0]]] * 3), g([[[10.0, 0.0]], [[1.0, 0.0]], [[3.0, 0.0]]]), g([8., 9., 10.]),
                     g([3., 2., 1.]))
    r2 = se.Rollouts(g([[[0, 0], [100, 0]]] * 3), g([[[100.0, 0.0]]] * 3), g([20.] * 3), g([20.] * 3))
    rf = _mock_rollouts(r1, r2)
    mpc = se.B200ConstrainedCemMpc(None, [], 2, 2, 1, 3, 2, 2, rollout_function=rf)
    mpc.optimize_trajectories(g([0, 0]))
    _, means, stds = rf.perform_rollouts.call_args_list[1][0][0]
    assert means[0][0].item() == 2 and means[0][1].item() == 0
    assert np.isclose(stds[0][0].item(), np.std([1.0, 3.0], ddof=1)) and stds[0][1] == 0

    r1 = se.Rollouts(g([[[0, 0], [1, 0]]] * 4),
                     g([[[10., 0.], [0., 0.]], [[3., 0.], [0., 0.]], [[1., 0.], [0., 0.]], [[2., 0.], [0., 0.]]]),
                     g([8., 7., 10., 9.]), g([0., 1., 0., 0.]))
    r2 = se.Rollouts(g([[[0, 0], [100, 0]]] * 4), g([[[100., 0.], [0., 0.]]] * 4), g([20.] * 4), g([0., 1., 0., 0.]))
    rf = _mock_rollouts(r1, r2)
    mpc = se.B200ConstrainedCemMpc(None, [], 2, 2, 1, 3, 2, 2, rollout_function=rf)
    mpc.optimize_trajectories(g([0, 0]))
    _, means, stds = rf.perform_rollouts.call_args_list[1][0][0]
    assert means[0][0].item() == 6 and means[0][1].item() == 0
    assert np.isclose(stds[0][0].item(), np.std([10.0, 2.0], ddof=1)) and stds[0][1] == 0


def test_reference_get_actions_known_answers_with_an_injected_rollout_function():
    """constrained_cem_mpc_test.py:302-394: None when nothing is feasible; the only / best feasible rollout's actions."""
    g = lambda x: torch.tensor(x, dtype=torch.float64, device=DEV)
    none = se.Rollouts(g([[[0, 0], [1, 0]]] * 3), g([[[0.0, 0.0]]] * 3), g([0., 0., 0.]), g([3., 2., 1.]))
    mpc = se.B200ConstrainedCemMpc(None, [], 2, 2, 1, 3, 2, 1, rollout_function=_mock_rollouts(none))
    actions, by_time = mpc.get_actions(g([0.0, 0.0]))
    assert actions is None and len(by_time) == 1 and torch.equal(by_time[0].constraint_costs, none.constraint_costs)
    for obj2, con2 in (([1., 20., 10.], [3., 0., 2.]), ([1., 5., 10.], [3., 0., 0.])):
        r1 = se.Rollouts(g([[[0, 0], [1, 0], [1, 0]]] * 3), g([[[0.0, 0.0], [0.0, 0.0]]] * 3), g([0.] * 3), g([3., 2., 1.]))
        r2 = se.Rollouts(g([[[0, 0], [1, 0], [1, 0]]] * 3),
                         g([[[0.0, 0.0], [0.0, 0.0]], [[1.0, 1.5], [3.0, 2.5]], [[0.0, 0.0], [0.0, 0.0]]]), g(obj2), g(con2))
        mpc = se.B200ConstrainedCemMpc(None, [], 2, 2, 1, 3, 2, 2, rollout_function=_mock_rollouts(r1, r2))
        actions, _ = mpc.get_actions(g([0.0, 0.0]))
        assert actions.cpu().tolist() == [[1.0, 1.5], [3.0, 2.5]]


# ---- a CemSSM that is not the B200 GP ------------------------------------------------------------------------------------
class ForeignSSM(se.CemSSM):
    """A CemSSM the kernels know nothing about (stands for the MC-dropout models of ssm_cem/dropout_ssm_cem.py).  It
    answers from a B200 GP internally so that the interface path can be compared with the fused path number for number."""

    def __init__(self, gp):
        super().__init__(gp.num_states, gp.num_actions)
        self._gp = gp
        self.calls = 0

    def predict_with_jacobians(self, states, actions):
        self.calls += 1
        return self._gp.predict_with_jacobians(states, actions)

    def predict_without_jacobians(self, states, actions):
        self.calls += 1
        return self._gp.predict_without_jacobians(states, actions)

    def predict_raw(self, z):
        return self._gp.predict_raw(z)

    def _update_model(self, x_train, y_train):
        pass

    def _train_model(self, x_train, y_train):
        pass

    def collect_metrics(self):
        return {}

    @property
    def parametric(self):
        return False


class UnknownTerminalConstraint(se.Constraint):
    """An arbitrary Constraint (per-rollout Python call): the terminal ellipsoid rule restated with torch ops."""

    def __init__(self, h_mat, h_vec):
        self._h_mat, self._h_vec = G(h_mat), G(h_vec).reshape(-1)

    def __call__(self, trajectory, actions) -> float:
        last = trajectory[-1]
        p, q = last[:2], last[2:].reshape(2, 2)
        d = self._h_mat @ p + torch.sqrt(torch.einsum('mi,ij,mj->m', self._h_mat, q, self._h_mat)) - self._h_vec
        return 10.0 * float((d >= 0).any())


def test_foreign_ssm_and_unknown_constraint_equal_the_fused_path():
    """Interface path (per-step Python predict + se_reach_step, per-rollout Python constraint) == fused se_cem_solve on
    the same candidates: identical elites and constraint costs, trajectories to rounding."""
    w = workloads.pendulum_workload(n_train=150, num_rollouts=96, time_horizon=4, num_elites=8, num_iterations=3)
    ssm, fused = workloads.build_solver(w, record='full', seed=9)
    res = fused.solve(G(w.p0).unsqueeze(0), seed=9)
    foreign = ForeignSSM(ssm)
    prob = se.SafeProblem(2, 1, w.l_mu, w.l_sigma, k_fb=w.k_fb, a=w.a, b=w.b, c_safety=w.beta_safety)
    cons = [se.ActionConstraint(se.TorchPolytope((w.act_a, w.act_b))), UnknownTerminalConstraint(w.h_mat, w.h_vec)]
    mpc = se.B200ConstrainedCemMpc(se.B200SafeDynamics(foreign, prob), cons, 6, 1, w.time_horizon, w.num_rollouts,
                                   w.num_elites, w.num_iterations, seed=9)
    assert not mpc.fused
    state = se.PQFlattener(2).flatten(G(w.p0).unsqueeze(0), None)[0]
    actions, by_time = mpc.get_actions(state)
    assert foreign.calls == w.time_horizon * w.num_iterations
    for i, r in enumerate(by_time):
        assert torch.equal(r.constraint_costs, res['constraint'][i, 0])
        assert rel_err(r.actions.cpu(), res['rollout_actions'][i, 0].cpu(), 1e-3) < 1e-12  # mean + std * eps: FMA or not
        assert rel_err(r.objective_costs.cpu(), res['objective'][i, 0].cpu()) < 1e-10
        assert rel_err(r.trajectories.cpu(), res['trajectories'][i, 0].cpu(), 1e-9) < 1e-10
        assert torch.equal(mpc.last_info['elite_ids'][i], res['elite_ids'][i, 0])
    assert mpc.last_info['best_id'] == int(res['best_id'][0])
    assert (actions is None) == (int(res['found'][0]) == 0)
    if actions is not None:
        assert rel_err(actions.cpu(), res['actions'][0].cpu(), 1e-3) < 1e-12


# ---- the environment's own objective ------------------------------------------------------------------------------------
class _Env:
    n_s, n_u = 2, 1

    def __init__(self, w, fn):
        self.l_mu, self.l_sigm = w.l_mu, w.l_sigma
        self._fn = fn

    def objective_cost_function(self, ps):
        return self._fn(ps)


class _Conf:
    use_prior_model, cem_record = True, 'full'

    def __init__(self, w):
        self.mpc_time_horizon, self.cem_num_rollouts = w.time_horizon, w.num_rollouts
        self.cem_num_elites, self.cem_num_iterations = w.num_elites, w.num_iterations


def _solver(w, ssm, fn, constraints):
    return se.B200CemSafeMPC(ssm, constraints, _Env(w, fn), _Conf(w), {'lin_model': (w.a, w.b)}, None, None,
                             w.beta_safety, safe_policy=lambda s: np.zeros(1), lqr=torch.as_tensor(w.k_fb),
                             verbose=False)


def test_environment_objective_functions_are_recognised_or_called():
    """safempc_cem.py:304-312.  The reference's two environment objectives select the kernel's built-in forms (theta
    read off the function, and followed when it advances, environments.py:488-515); any other function is called per
    step on the recorded centres (interface path) -- checked against the oracle with the same function."""
    w = workloads.pendulum_workload(n_train=300, num_rollouts=256, time_horizon=4, num_elites=8, num_iterations=3)
    ssm, fused = workloads.build_solver(w, record='full', seed=13)
    cons = fused._constraints
    theta = [0.12]
    pendulum = lambda ps: torch.abs(torch.full_like(ps[:, 1], theta[0]) - ps[:, 1])  # environments.py:505-510
    assert se.classify_objective(_Env(w, pendulum))[:2] == (_lib.SE_OBJ_ABS_STATE1, 0.12)
    assert se.classify_objective(_Env(w, lambda ps: -ps[:, -1]))[:2] == (_lib.SE_OBJ_NEG_LAST_STATE, 0.0)
    assert se.classify_objective(_Env(w, lambda ps: None))[:2] == (_lib.SE_OBJ_NEG_VARIANCE, 0.0)
    solver = _solver(w, ssm, pendulum, cons)
    assert solver._mpc.fused and solver._mpc._problem.objective_theta == 0.12
    theta[0] = -0.05  # the environment reached its target and advanced to the next one
    solver.get_action(np.zeros(2))
    assert solver._mpc.fused and solver._mpc._problem.objective_theta == -0.05

    # a function the kernels do not know: smooth, non-separable in the state
    custom = lambda ps: (ps[:, 0] - 0.03) ** 2 + 0.5 * torch.abs(ps[:, 1] + 0.02)
    solver = _solver(w, ssm, custom, cons)
    assert not solver._mpc.fused
    state = se.PQFlattener(2).flatten(G(w.p0).unsqueeze(0), None)[0]
    solver._mpc.seed = 13
    actions, by_time = solver._mpc.get_actions(state)

    class Problem(oracle.SafeCemProblem):
        def objective_cost(self, p_next, sigma):
            return custom(p_next)

    gp = oracle.OracleGpCemSSM(2, 1, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    prob = Problem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a, w.act_b)
    noise = np.stack([oracle.philox_normal_noise(13, i, 0, w.num_rollouts, w.time_horizon, 1)
                      for i in range(w.num_iterations)])
    ref = oracle.cem_solve(prob, T(w.p0), T(noise), w.num_elites)
    for i, r in enumerate(ref['rollouts']):
        assert np.array_equal(by_time[i].constraint_costs.cpu().numpy(), r.constraint_costs.numpy())
        assert rel_err(by_time[i].objective_costs.cpu(), r.objective_costs) < 1e-8
        assert rel_err(by_time[i].trajectories.cpu(), r.trajectories, 1e-6) < 1e-7
        assert np.array_equal(solver._mpc.last_info['elite_ids'][i].cpu().numpy(), ref['elite_ids'][i].numpy())
    assert solver._mpc.last_info['best_id'] == ref['best_id']
    assert max((r.constraint_costs == 0).sum().item() for r in ref['rollouts']) >= w.num_elites


# ---- get_action -> update_model -> get_action on ONE solver ----------------------------------------------------------------
def test_model_updates_between_solves_never_reuse_stale_state():
    """The episode loop (episode_runner.py:206-231, 122-127): a solver that keeps solving while its model is replaced
    (new data, new hyper-parameters -- the GP handle is destroyed and re-created, usually at the same address) must
    give what a freshly built solver gives: no stale CEM handle, no stale CUDA graph."""
    w = workloads.pendulum_workload(n_train=300, num_rollouts=256, time_horizon=4, num_elites=8, num_iterations=3)
    ssm, mpc = workloads.build_solver(w, record='none', seed=0)
    p0 = G(w.p0).unsqueeze(0)

    def fresh(updates, ls, os_, nz):
        """A new model + solver that saw the same update_model calls since the last full refresh."""
        s2 = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
        for xs, ys, replace in updates:
            s2.update_model(G(xs), G(ys), replace_old=replace)
        prob = se.SafeProblem(2, 1, w.l_mu, w.l_sigma, k_fb=w.k_fb, a=w.a, b=w.b, c_safety=w.beta_safety)
        return se.B200ConstrainedCemMpc(se.B200SafeDynamics(s2, prob), mpc._constraints, 6, 1, w.time_horizon,
                                        w.num_rollouts, w.num_elites, w.num_iterations, record='none'), s2

    rng = np.random.default_rng(3)
    x, y = w.x_train, w.y_train
    updates = [(x, y, True)]
    for step in range(4):
        for _ in range(3):  # first sighting runs eagerly, the second is captured, the third replays the graph
            got = mpc.solve(p0, seed=100 + step, reuse_outputs=True)
            got = {k: v.clone() for k, v in got.items() if k in ('actions', 'means', 'stds', 'elite_ids', 'found')}
        ref_mpc, keep = fresh(updates, ssm.lengthscale, ssm.outputscale, ssm.noise)
        want = ref_mpc.solve(p0, seed=100 + step)
        for k in got:
            assert torch.equal(got[k], want[k]), (step, k)
        if step % 2 == 0:  # same size, new hyper-parameters and targets (what opt_hyp=True amounts to)
            ssm.lengthscale = ssm.lengthscale * 1.07
            ssm.outputscale = ssm.outputscale * 0.9
            y = y + 0.002 * rng.standard_normal(y.shape)
            ssm.update_model(G(x), G(y), replace_old=True)
            updates = [(x, y, True)]
        else:  # appended samples (n changes, n_pad does not): the incremental factor update
            x_add = rng.uniform(-0.3, 0.3, (5, 3))
            y_add = 0.01 * rng.standard_normal((5, 2))
            ssm.update_model(G(x_add), G(y_add), replace_old=False)
            updates.append((x_add, y_add, False))
            x, y = np.vstack((x, x_add)), np.vstack((y, y_add))
    # a stale GP handle is refused loudly by the library itself
    import ctypes as C
    stale = C.c_void_p(None)
    cs, cfg = mpc._problem.c_struct(), _lib.SeCemConfig(1, 256, 4, 8, 3, 0, 0, 256)
    old = ssm.handle
    _lib.check(_lib.load().se_cem_create(C.byref(stale), old, C.byref(cs), C.byref(cfg)))
    ssm.update_model(G(x), G(y), replace_old=True)
    out = _lib.SeCemOutputs()
    rc = _lib.load().se_cem_solve(stale, _lib.ptr(p0), None, C.c_uint64(1), C.byref(out), _lib.stream_ptr(DEV))
    assert rc != 0 and b'destroyed or replaced' in _lib.load().se_last_error()
    _lib.load().se_cem_destroy(stale)


# ---- _fix_zeros_nans on the device (gp_reachability_pytorch.py:234-243) ----------------------------------------------------
def _zero_variance_gp(monkeypatch):
    """A model whose predictive variance is EXACTLY zero everywhere: outputscale = noise = 0 at the boundary, factors
    taken from an identity covariance (k* = 0, so s + noise - ||L^-1 k*||^2 = 0 + 0 - 0)."""
    # (float64 DMMA kernels: the digit planes of the tensor-core path are scaled by outputscale, which must be > 0 there)
    gp = se.B200GpCemSSM(None, 2, 1, np.ones((2, 3)), np.zeros(2), np.zeros(2), device=DEV, gemm_impl=_lib.SE_GEMM_MMA)
    monkeypatch.setattr(gp, 'kernel_matrix',
                        lambda x, *a, **k: torch.eye(x.shape[0], dtype=torch.float64, device=DEV).repeat(2, 1, 1))
    rng = np.random.default_rng(0)
    gp.update_model(G(rng.uniform(-1, 1, (300, 3))), G(0.01 * rng.standard_normal((300, 2))), replace_old=True)
    return gp


def test_exact_zero_and_nan_variances_through_reach_step_follow_the_batch_rule():
    """`se_reach_step` on caller-given posteriors: any NaN raises (the reference's ValueError); an exact zero in the batch
    sets EVERY entry <= 0 to 1e-5 (negatives included) and warns; without an exact zero a negative variance survives,
    turns into NaN at the next sqrt and raises there -- all three as oracle.fix_zeros_nans / onestep_reachability do."""

    class Fixed(se.CemSSM):
        parametric = False

        def __init__(self, mean, var):
            super().__init__(2, 1)
            self.m, self.v = mean, var

        def predict_without_jacobians(self, s, a):
            return self.m, self.v

        def predict_with_jacobians(self, s, a):
            return self.m, self.v, torch.zeros((s.shape[0], 2, 3), dtype=s.dtype, device=s.device)

        predict_raw = _update_model = _train_model = collect_metrics = None

    p, u = G(np.zeros((4, 2))), G(np.zeros((4, 1)))
    l = G([0.05, 0.02])
    mean = G(0.01 * np.arange(8).reshape(4, 2))
    var = G([[1e-4, 2e-4], [0.0, 3e-4], [-2e-5, 1e-4], [5e-4, 1e-4]])  # one exact zero, one negative
    ref_ssm = Fixed(mean.cpu(), var.cpu())
    for q in (None, G(np.tile(1e-3 * np.eye(2), (4, 1, 1)))):
        p1, q1, sig = se.onestep_reachability(p, Fixed(mean, var), u, l, l, q, k_fb=G([[-0.1, -0.2]]), c_safety=2.0)
        rp, rq, rs = oracle.onestep_reachability(p.cpu(), ref_ssm, u.cpu(), l.cpu(), l.cpu(), None if q is None else q.cpu(),
                                                 T([[-0.1, -0.2]]), 2.0)
        assert sig.cpu().tolist() == rs.tolist() and sig[1, 0] == 1e-5 and sig[2, 0] == 1e-5
        assert rel_err(q1.cpu(), rq, 1e-12) < 1e-12 and rel_err(p1.cpu(), rp, 1e-3) < 1e-12
    neg_only = G([[1e-4, 2e-4], [1e-4, 3e-4], [-2e-5, 1e-4], [5e-4, 1e-4]])
    with pytest.raises(ValueError):
        se.onestep_reachability(p, Fixed(mean, neg_only), u, l, l, None, c_safety=2.0)
    with pytest.raises(ValueError):
        oracle.onestep_reachability(p.cpu(), Fixed(mean.cpu(), neg_only.cpu()), u.cpu(), l.cpu(), l.cpu(), None, None, 2.0)
    with_nan = var.clone()
    with_nan[3, 1] = float('nan')
    with pytest.raises(ReachabilityNumericsError, match='nan'):
        se.onestep_reachability(p, Fixed(mean, with_nan), u, l, l, None, c_safety=2.0)


def test_numeric_flags_through_the_gp_kernels_and_the_cem_solve(monkeypatch, capsys, tmp_path):
    """Exact-zero variances produced BY the kernels (se_reach_onestep, se_cem_solve): counted, replaced by 1e-5, the
    solve carries on with a warning.  A non-finite state: NaN variances -> ValueError, from both entry points."""
    monkeypatch.chdir(tmp_path)
    gp = _zero_variance_gp(monkeypatch)
    p, u = G(0.1 * np.ones((8, 2))), G(np.zeros((8, 1)))
    l = G([0.05, 0.02])
    p1, q1, sig = se.onestep_reachability(p, gp, u, l, l, None, c_safety=2.0)
    assert (sig == 1e-5).all() and 'found 0' in capsys.readouterr().out
    assert torch.allclose(q1, torch.diag_embed(torch.full((8, 2), 2 * 4.0 * 1e-5, device=DEV, dtype=torch.float64)))
    w = workloads.pendulum_workload(n_train=300, num_rollouts=128, time_horizon=3, num_elites=8, num_iterations=2)
    prob = se.SafeProblem(2, 1, w.l_mu, w.l_sigma, k_fb=w.k_fb, a=w.a, b=w.b, c_safety=2.0)
    cons = [se.ActionConstraint(se.TorchPolytope((w.act_a, w.act_b))), se.EllipsoidTerminalConstraint(2, w.h_mat, w.h_vec)]
    mpc = se.B200ConstrainedCemMpc(se.B200SafeDynamics(gp, prob), cons, 6, 1, 3, 128, 8, 2, record='full')
    res = mpc.solve(G(np.zeros((1, 2))), seed=1)
    assert int(res['flags'][0]) == 0 and int(res['flags'][1]) == 2 * 3 * 128 * 2  # iterations x steps x candidates x n_s
    assert 'found 0' in capsys.readouterr().out
    assert torch.isfinite(res['trajectories']).all()

    # non-finite states: a NaN enters the GP -> NaN variances -> the reference's ValueError, from both entry points and on
    # both variance engines (an infinite state turns into NaN within a step in the reference; here it is reported at once)
    for impl in (_lib.SE_GEMM_AUTO, _lib.SE_GEMM_MMA):
        ssm, mpc = workloads.build_solver(w, record='costs', seed=0, gemm_impl=impl)
        for bad in (float('nan'), float('inf')):
            with pytest.raises(ReachabilityNumericsError):
                mpc.solve(G([[bad, 0.0]]), seed=2)
        with pytest.raises(ReachabilityNumericsError):
            se.onestep_reachability(G([[0.0, float('nan')], [0.0, 0.0]]), ssm, G([[0.0], [0.1]]), l, l, None,
                                    c_safety=2.0)
        mpc.solve(G(np.zeros((1, 2))), seed=3)  # and the solver is usable afterwards
    # the reference dumps the model and the offending inputs before it raises (gp_reachability_pytorch.py:256-266)
    dump = torch.load(tmp_path / 'negative_variance_state.pt')
    assert set(dump) == {'gp_model', 'gp_likelihood', 'state', 'action', 'x_train', 'y_train'}
    assert dump['x_train'].shape == (300, 3) and torch.isnan(dump['state']).any()


def test_clamp_nonpositive_switch_of_the_cem_solve(monkeypatch):
    """se_problem.clamp_nonpositive: the solve-wide form of the reference's batch rule (a variance <= 0 becomes 1e-5 and is
    counted) -- here on the zero-variance model, where strict and clamped mode must agree."""
    gp = _zero_variance_gp(monkeypatch)
    w = workloads.pendulum_workload(n_train=300, num_rollouts=128, time_horizon=3, num_elites=8, num_iterations=2)
    cons = [se.ActionConstraint(se.TorchPolytope((w.act_a, w.act_b))), se.EllipsoidTerminalConstraint(2, w.h_mat, w.h_vec)]
    out = []
    for clamp in (False, True):
        prob = se.SafeProblem(2, 1, w.l_mu, w.l_sigma, k_fb=w.k_fb, a=w.a, b=w.b, c_safety=2.0, clamp_nonpositive=clamp)
        mpc = se.B200ConstrainedCemMpc(se.B200SafeDynamics(gp, prob), cons, 6, 1, 3, 128, 8, 2, record='full')
        out.append(mpc.solve(G(np.zeros((1, 2))), seed=1))
    assert torch.equal(out[0]['trajectories'], out[1]['trajectories'])
    assert out[0]['flags'].tolist() == out[1]['flags'].tolist()


# ---- batched runner (episode_runner.py:165-324) -------------------------------------------------------
class _LinearEnv:
    """A deterministic stand-in for the reference's Environment: x' = a x + b u, observation = x' - (a x + b u) + 0."""
    n_s, n_u = 2, 1

    def __init__(self, w, x0, fail_at=None):
        self._a, self._b, self._x0, self._fail_at = w.a, w.b, np.asarray(x0, dtype=np.float64), fail_at
        self.l_mu, self.l_sigm = w.l_mu, w.l_sigma
        self.steps = 0

    def objective_cost_function(self, ps):
        return None

    def reset(self, mean=None, std=None):
        self.state, self.steps = self._x0.copy(), 0
        return self.state

    def step(self, action):
        action = np.asarray(action, dtype=np.float64).reshape(1)
        self.state = self._a @ self.state + self._b @ action
        self.steps += 1
        done = self._fail_at is not None and self.steps >= self._fail_at
        return action, self.state.copy(), self.state.copy(), done, 0

    def collect_metrics(self):
        return {'steps': self.steps}


class _Metrics:
    def __init__(self):
        self.scalars, self.non_scalars = {}, []

    def log_scalar(self, name, value, step):
        self.scalars[name] = value

    def log_non_scalars(self, d, step):
        self.non_scalars.append(d)


def test_batched_runner_equals_single_scenario_runs():
    """8 scenarios x 5 steps through do_rollout_batch (one batched solve per step) == 8 do_rollout runs, action for
    action: the solver of scenario i is keyed to global instance i, so both draw the same candidates."""
    from safe_exploration_b200.runner import do_rollout, do_rollout_batch
    w = workloads.pendulum_workload(n_train=300, num_rollouts=256, time_horizon=4, num_elites=8, num_iterations=3)
    ssm, proto = workloads.build_solver(w, record='none')
    rng = np.random.default_rng(12)
    x0 = 0.03 * rng.standard_normal((8, 2))
    x0[5] = [3.0, 3.0]  # far outside the safe set: never feasible -> safe controller
    mk_envs = lambda: [_LinearEnv(w, x0[i], fail_at=3 if i == 2 else None) for i in range(8)]

    def solver_for(env):
        return se.B200CemSafeMPC(ssm, proto._constraints, env, _Conf(w), {'lin_model': (w.a, w.b)}, None, None,
                                 w.beta_safety, safe_policy=lambda s: -0.1 * s[:1], lqr=torch.as_tensor(w.k_fb),
                                 verbose=False)

    envs = mk_envs()
    batch_solver = solver_for(envs[0])
    batch_solver._mpc.seed = 700
    sinks = [_Metrics() for _ in envs]
    batched = do_rollout_batch(envs, 5, batch_solver, metrics=sinks, cost=lambda s: float(np.abs(s).sum()))
    singles = []
    for i, env in enumerate(mk_envs()):
        solver = solver_for(env)
        solver._mpc.seed, solver._mpc.first_instance = 700, i
        m = _Metrics()
        singles.append((do_rollout(env, 5, scenario_id=i, metrics=m, solver=solver,
                                   cost=lambda s: float(np.abs(s).sum())), m))
    kinds = set()
    for i in range(8):
        (xx, yy, cc, codes, failed), (sx, sy, sc, scodes, sfailed), m1 = batched[i], singles[i][0], singles[i][1]
        assert np.array_equal(xx, sx) and np.array_equal(yy, sy) and cc == sc, i
        assert np.array_equal(codes, scodes) and failed == sfailed
        assert xx.shape[0] == (4 if i == 2 else 6)  # scenario 2 ends after three steps, the others run all five
        for name in ('episode_length', 'mpc_found_solution_count', 'mpc_previous_solution_count',
                     'safe_controller_fallback_count', 'env_result'):
            assert sinks[i].scalars[name] == m1.scalars[name], (i, name)
        assert sinks[i].scalars['mean_time_in_solver'] > 0 and sinks[i].non_scalars[0] == {'steps': xx.shape[0] - 1}
        kinds.update(codes[1:, 0].tolist())
    assert kinds == {0.0, 1.0}  # both MPC plans and safe-controller steps occurred
