"""GPU parity tests of the model-refresh kernels: covariance assembly and the gradient of
the exact marginal log likelihood, against the float64 oracle (autograd through the Cholesky factorisation).

Tolerances: covariance entries 1e-13 relative to the prior scale (one branch-free exp, <= 2 ulp); marginal likelihood
1e-10 relative; gradient 1e-8 relative to the largest entry of the output's gradient (K^-1 carries cond(K) * eps).
"""
import numpy as np
import pytest
import torch

import safe_exploration_b200 as se
from oracle import oracle
from safe_exploration_b200 import workloads

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))
G = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64)).to(DEV)


def problem(n, n_s, n_u, seed, noise=1e-2):
    g = torch.Generator().manual_seed(seed)
    d = n_s + n_u
    x = torch.rand((n, d), generator=g, dtype=torch.float64) * 2 - 1
    y = torch.stack([torch.sin((o + 2.0) * x[:, o % d]) * 0.3 + 0.1 * x[:, (o + 1) % d] for o in range(n_s)], dim=1)
    y = y + 0.05 * torch.randn((n, n_s), generator=g, dtype=torch.float64)
    ls = 0.4 + torch.rand((n_s, d), generator=g, dtype=torch.float64)
    os_ = 0.05 + torch.rand((n_s,), generator=g, dtype=torch.float64)
    nz = noise * (1 + torch.rand((n_s,), generator=g, dtype=torch.float64))
    return x, y, ls, os_, nz


@pytest.mark.parametrize('n,n_s,n_u', [(1, 1, 1), (63, 2, 1), (64, 1, 2), (65, 2, 1), (200, 4, 1), (333, 6, 3), (130, 3, 2)])
def test_kernel_matrix_matches_oracle(n, n_s, n_u):
    x, y, ls, os_, nz = problem(n, n_s, n_u, 10 + n)
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV)
    k = gp.kernel_matrix(x.to(DEV)).cpu()
    ref = oracle.rbf_ard(x, x, ls, os_) + torch.diag_embed(nz.view(-1, 1).expand(-1, n))
    scale = (os_ + nz).view(-1, 1, 1)
    assert float(((k - ref).abs() / scale).max()) < 1e-13
    assert torch.equal(k, k.transpose(1, 2)), 'the assembled covariance must be exactly symmetric'


def test_kernel_matrix_rejects_bad_input():
    x, y, ls, os_, nz = problem(10, 2, 1, 3)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    with pytest.raises(se.SecemError):
        gp.kernel_matrix(x)  # host tensor: no CPU fallback
    bad = ls.clone()
    bad[0, 0] = 0.0
    with pytest.raises(se.SecemError):
        gp.kernel_matrix(x.to(DEV), bad, os_, nz)


@pytest.mark.parametrize('n,n_s,n_u', [(1, 1, 1), (5, 2, 1), (64, 2, 1), (65, 1, 2), (129, 2, 1), (300, 4, 1), (257, 6, 3)])
def test_mll_gradient_matches_oracle_autograd(n, n_s, n_u):
    x, y, ls, os_, nz = problem(n, n_s, n_u, 100 + n)
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV)
    mll, grad, alpha = gp.mll_and_grad(x.to(DEV), y.to(DEV))
    ref_mll, ref_grad = oracle.exact_mll_and_grad(x, y, ls, os_, nz)
    assert float(((mll.cpu() - ref_mll).abs() / ref_mll.abs()).max()) < 1e-10
    scale = ref_grad.abs().max(dim=1, keepdim=True).values
    assert float(((grad.cpu() - ref_grad).abs() / scale).max()) < 1e-8


def test_mll_gradient_is_deterministic():
    x, y, ls, os_, nz = problem(500, 2, 1, 7)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    a = gp.mll_and_grad(x.to(DEV), y.to(DEV))[1]
    b = gp.mll_and_grad(x.to(DEV), y.to(DEV))[1]
    assert torch.equal(a, b)


def test_training_follows_the_oracle_adam_trajectory():
    """GpCemSSM._train_model (gp_ssm_cem.py:103-129): same losses and hyper-parameters as the oracle's autograd + Adam
    loop after 25 iterations (1e-7: Adam divides by sqrt(v), which amplifies gradient rounding in the first steps)."""
    class Conf:
        exact_gp_training_iterations = 25
        exact_gp_kernel = 'rbf'
    x, y, ls, os_, nz = problem(150, 2, 1, 42)
    gp = se.B200GpCemSSM(Conf(), 2, 1, ls, os_, nz, device=DEV)
    gp.update_model(x.to(DEV), y.to(DEV), opt_hyp=True, replace_old=True)
    rls, ros, rnz, rlosses = oracle.train_hyperparameters(x, y, ls, os_, nz, 25)
    losses = gp.collect_metrics()['losses']
    assert len(losses) == 25 and losses[-1] < losses[0]
    assert np.allclose(losses, rlosses, rtol=1e-7, atol=0)
    assert np.allclose(gp.lengthscale.numpy(), rls.numpy(), rtol=1e-7)
    assert np.allclose(gp.outputscale.numpy(), ros.numpy(), rtol=1e-7)
    assert np.allclose(gp.noise.numpy(), rnz.numpy(), rtol=1e-7)
    # and the refreshed model predicts with the trained hyper-parameters
    ogp = oracle.OracleGpCemSSM(2, 1, rls, ros, rnz)
    ogp.update_model(x, y, replace_old=True)
    z = x[:40] * 0.9
    m, v = gp.predict_raw(z.to(DEV))
    om, ov = ogp.predict_raw(z)
    assert np.allclose(m.cpu().numpy(), om.numpy(), rtol=1e-6, atol=1e-9)
    assert np.allclose(v.cpu().numpy(), ov.numpy(), rtol=1e-6, atol=1e-12)


def test_pendulum_workload_gradient_at_full_size():
    """Size-independent property at BASELINE's C2 model size (n = 2000): the analytic gradient agrees with a central
    finite difference of the marginal likelihood along a random direction in log-hyper-parameter space."""
    w = workloads.pendulum_workload(num_rollouts=256)
    x, y = G(w.x_train), G(w.y_train)
    ls = torch.as_tensor(w.lengthscale, dtype=torch.float64).reshape(2, 3)
    os_ = torch.as_tensor(w.outputscale, dtype=torch.float64).reshape(2)
    nz = torch.as_tensor(w.noise, dtype=torch.float64).reshape(2) * 100  # keep cond(K) ~ 1e5 for the difference quotient
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    mll, grad, _ = gp.mll_and_grad(x, y, ls, os_, nz)
    theta = torch.cat((ls, os_.unsqueeze(1), nz.unsqueeze(1)), dim=1)
    g = torch.Generator().manual_seed(5)
    direction = torch.randn(theta.shape, generator=g, dtype=torch.float64)
    eps = 1e-5
    def f(t):
        return gp.mll_and_grad(x, y, t[:, :3], t[:, 3], t[:, 4])[0].cpu()
    fd = (f(theta * torch.exp(eps * direction)) - f(theta * torch.exp(-eps * direction))) / (2 * eps)
    an = (grad.cpu() * theta * direction).sum(1)
    assert float(((fd - an).abs() / an.abs()).max()) < 1e-5


def test_create_solver_end_to_end_matches_oracle(golden_cem):
    """The config-selected path (utils_config.py:58-135, cem_ssm = 'exact_gp_b200'): update_model with the prior
    subtracted (safempc_cem.py:314-327), one get_action, and the same solve restated by the oracle on the candidates
    the device drew (Philox stream of seed 0): identical constraint decisions, elites and returned plan."""
    from test_host_cabi import _CemConf, _PendulumEnv
    g = golden_cem

    class Conf(_CemConf):
        mpc_time_horizon, cem_num_rollouts, cem_num_elites, cem_num_iterations = 4, 128, 12, 3
        cem_record = 'full'
    solver, safe_policy = se.create_solver(Conf(), _PendulumEnv(), lengthscale=g['lengthscale'],
                                           outputscale=g['outputscale'], noise=g['noise_var'], device=DEV)
    a, b = solver.lin_model
    x, y_res = g['x_train'], g['y_train']
    solver.update_model(x, y_res + x[:, :2] @ np.asarray(a).T + x[:, 2:] @ np.asarray(b).T, replace_old=True)
    action, result = solver.get_action(g['p0'])
    info = solver._mpc.last_info

    noise = np.stack([oracle.philox_normal_noise(0, i, 0, 128, 4, 1) for i in range(3)])
    ogp = oracle.OracleGpCemSSM(2, 1, g['lengthscale'], g['outputscale'], g['noise_var'])
    ogp.update_model(T(x), T(y_res), replace_old=True)
    env = _PendulumEnv()
    h, v = env.get_safety_constraints()[:2]
    k, _, _ = se.dlqr(a, b, Conf.lqr_wx_cost, Conf.lqr_wu_cost)
    prob = oracle.SafeCemProblem(ogp, env.l_mu, env.l_sigm, -k, a, b, 2.0, h, v[:, 0], g['act_a'], g['act_b'])
    ref = oracle.cem_solve(prob, T(g['p0']), T(noise), 12)
    assert np.array_equal(info['elite_ids'][:, 0].cpu().numpy(), ref['elite_ids'].numpy())
    for i, r in enumerate(ref['rollouts']):
        assert np.array_equal(info['constraint'][i, 0].cpu().numpy(), r.constraint_costs.numpy())
        assert np.allclose(info['objective'][i, 0].cpu().numpy(), r.objective_costs.numpy(), rtol=1e-8, atol=0)
    if ref['best_id'] >= 0:
        assert result == se.MpcResult.FOUND_SOLUTION
        assert np.allclose(action, ref['rollouts'][-1].actions[ref['best_id'], 0].numpy(), rtol=1e-12)
    else:
        assert result == se.MpcResult.SAFE_CONTROLLER and np.allclose(action, safe_policy(g['p0']))
