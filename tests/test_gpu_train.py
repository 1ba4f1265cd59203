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


# ---- on-device factor refresh (se_gp_factor / se_gp_kinv): SURVEY.md 8f rank 1 -----------------------------------------
def _factor_through_cabi(k, first_block=0, w_old=None, wt_old=None):
    """k [n_s, n, n] (device, float64) -> (W_pad, W^T_pad, diag, info) through the C ABI, identity on the padding."""
    from safe_exploration_b200 import _lib
    lib = _lib.load()
    n_s, n = k.shape[0], k.shape[1]
    n_pad = -(-n // 128) * 128
    k_pad = torch.zeros((n_s, n_pad, n_pad), dtype=torch.float64, device=DEV)
    k_pad[:, :n, :n] = k
    k_pad.diagonal(dim1=1, dim2=2)[:, n:] = 1.0
    w = torch.zeros_like(k_pad)
    wt = torch.zeros_like(k_pad)
    if w_old is not None:
        keep = first_block * 128
        w[:, :keep, :keep] = w_old[:, :keep, :keep]
        wt[:, :keep, :keep] = wt_old[:, :keep, :keep]
    diag = torch.ones((n_s, n_pad), dtype=torch.float64, device=DEV)
    ws = torch.empty(lib.se_gp_factor_workspace(n_s, n_pad) // 8, dtype=torch.float64, device=DEV)
    info = torch.zeros(n_s, dtype=torch.int32, device=DEV)
    _lib.check(lib.se_gp_factor(n_s, n_pad, first_block, _lib.ptr(k_pad), _lib.ptr(w), _lib.ptr(wt), _lib.ptr(diag),
                                _lib.ptr(ws), _lib.ptr(info), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    return w, wt, diag, info


@pytest.mark.parametrize('n,n_s,n_u', [(1, 1, 1), (100, 2, 1), (128, 1, 1), (129, 2, 1), (300, 4, 1), (700, 2, 1), (1100, 3, 2)])
def test_native_factor_matches_a_float64_cholesky(n, n_s, n_u):
    """W = L^-1 from the blocked on-device factorisation against torch's float64 Cholesky + triangular solve ON THE CPU
    (LAPACK): W L = I to 1e-9, W to 1e-9 of its largest entry per output (K has condition number ~ n s / noise = 1e5),
    diag(L) to 1e-12, W^T the exact transpose, zeros above the diagonal, the identity on the padding."""
    x, y, ls, os_, nz = problem(n, n_s, n_u, 500 + n)
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV)
    k = gp.kernel_matrix(x.to(DEV))
    w, wt, diag, info = _factor_through_cabi(k)
    assert int(info.abs().max()) == 0
    chol = torch.linalg.cholesky(k.cpu())
    ref = torch.linalg.solve_triangular(chol, torch.eye(n, dtype=torch.float64).expand(n_s, -1, -1), upper=False)
    got = w[:, :n, :n].cpu()
    assert float((got - ref).abs().amax(dim=(1, 2)).div(ref.abs().amax(dim=(1, 2))).max()) < 1e-9
    eye_err = torch.matmul(got, chol) - torch.eye(n, dtype=torch.float64)
    assert float(eye_err.abs().max()) < 1e-9
    assert float((diag[:, :n].cpu() / chol.diagonal(dim1=1, dim2=2) - 1).abs().max()) < 1e-12
    assert torch.equal(wt, w.transpose(1, 2))
    assert float(torch.triu(w, diagonal=1).abs().max()) == 0.0
    n_pad = w.shape[1]
    if n_pad > n:
        pad = w[:, n:, :]
        assert torch.equal(pad[:, :, n:], torch.eye(n_pad - n, dtype=torch.float64, device=DEV).expand(n_s, -1, -1))
        assert float(pad[:, :, :n].abs().max()) == 0.0


def test_native_factor_extends_an_old_factor_block_row_by_block_row():
    """update_model(replace_old=False) (ssm_cem.py:72-99): passing the first block row that changed reproduces the
    factor of the grown training set -- bit for bit, since every block row is computed from the rows above it by the same
    kernels in the same order."""
    x, y, ls, os_, nz = problem(700, 2, 1, 77)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    k_all = gp.kernel_matrix(x.to(DEV))
    w_full, wt_full, diag_full, _ = _factor_through_cabi(k_all)
    for n_old in (128, 300, 650):
        w_old, wt_old, _, _ = _factor_through_cabi(k_all[:, :n_old, :n_old].contiguous())
        fb = n_old // 128
        big_w = torch.zeros_like(w_full)
        big_wt = torch.zeros_like(w_full)
        big_w[:, :w_old.shape[1], :w_old.shape[1]] = w_old
        big_wt[:, :w_old.shape[1], :w_old.shape[1]] = wt_old
        w_inc, wt_inc, diag_inc, info = _factor_through_cabi(k_all, first_block=fb, w_old=big_w, wt_old=big_wt)
        assert int(info.abs().max()) == 0
        assert torch.equal(w_inc, w_full) and torch.equal(wt_inc, wt_full)
        assert torch.equal(diag_inc[:, fb * 128:], diag_full[:, fb * 128:])


def test_native_inverse_covariance_matches_float64_and_feeds_the_gradient():
    from safe_exploration_b200 import _lib
    lib = _lib.load()
    n, n_s = 450, 3
    x, y, ls, os_, nz = problem(n, n_s, 1, 901)
    gp = se.B200GpCemSSM(None, n_s, 1, ls, os_, nz, device=DEV)
    k = gp.kernel_matrix(x.to(DEV))
    w, wt, _, _ = _factor_through_cabi(k)
    n_pad = w.shape[1]
    kinv = torch.full((n_s, n_pad, n_pad), float('nan'), dtype=torch.float64, device=DEV)
    _lib.check(lib.se_gp_kinv(n_s, n_pad, _lib.ptr(wt), _lib.ptr(kinv), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    ref = torch.linalg.inv(k.cpu())
    got = torch.tril(kinv[:, :n, :n]).cpu()
    assert torch.isfinite(got).all()
    assert float((got - torch.tril(ref)).abs().amax(dim=(1, 2)).div(ref.abs().amax(dim=(1, 2))).max()) < 1e-9
    # the same model through both factor engines: marginal likelihood and gradient agree
    a = se.B200GpCemSSM(None, n_s, 1, ls, os_, nz, device=DEV, factor_impl='native').mll_and_grad(x.to(DEV), y.to(DEV))
    b = se.B200GpCemSSM(None, n_s, 1, ls, os_, nz, device=DEV, factor_impl='cusolver').mll_and_grad(x.to(DEV), y.to(DEV))
    assert float(((a[0] - b[0]).abs() / b[0].abs()).max()) < 1e-10
    scale = b[1].abs().amax(dim=1, keepdim=True)
    assert float(((a[1] - b[1]).abs() / scale).max()) < 1e-8


def test_native_factor_reports_a_matrix_that_is_not_positive_definite():
    k = torch.eye(200, dtype=torch.float64, device=DEV).unsqueeze(0).repeat(2, 1, 1)
    k[1, 150, 150] = -1.0
    _, _, _, info = _factor_through_cabi(k)
    assert info.cpu().tolist() == [0, 151]
    gp = se.B200GpCemSSM(None, 2, 1, device=DEV)
    k_pad = torch.eye(256, dtype=torch.float64, device=DEV).unsqueeze(0).repeat(2, 1, 1)
    k_pad[1, 150, 150] = -1.0
    with pytest.raises(torch.linalg.LinAlgError, match='output 1'):  # the host wrapper raises like torch.linalg.cholesky
        gp._run_factor(k_pad, torch.zeros_like(k_pad), torch.zeros_like(k_pad), 0)


def test_update_model_engines_agree_and_append_reuses_block_rows():
    """The whole `update_model` through both factor engines: predictions agree to 1e-9 / 1e-8 (variance), and an append
    reports the incremental path."""
    x, y, ls, os_, nz = problem(900, 2, 1, 31)
    z = torch.rand((257, 3), dtype=torch.float64, device=DEV) * 2 - 1
    out = {}
    for impl in ('native', 'cusolver'):
        gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, factor_impl=impl)
        gp.update_model(x[:500].to(DEV), y[:500].to(DEV), replace_old=True)
        assert gp.last_update == 'full'
        gp.update_model(x[500:].to(DEV), y[500:].to(DEV))
        assert gp.last_update == 'incremental'
        out[impl] = gp.predict_with_jacobians(z[:, :2].contiguous(), z[:, 2:].contiguous())
    for a, b, tol in zip(out['native'], out['cusolver'], (1e-9, 1e-8, 1e-9)):
        assert float((a - b).abs().max() / b.abs().max()) < tol


def test_native_factor_properties_at_the_full_c3_model_size():
    """BASELINE C3's model (cart-pole, n = 5000, n_s = 4: 40 block rows, the long DMMA products and the side stream at
    their full size) checked through size-independent properties instead of a CPU factorisation: W K W^T = I and
    W^T W K = I to 1e-8 (cond K ~ 1e6 here), W^T the exact transpose, 2 sum log diag L = log det K to 1e-10."""
    w = workloads.cartpole_workload(num_rollouts=256)
    x = torch.as_tensor(w.x_train, device=DEV)
    n = x.shape[0]
    gp = se.B200GpCemSSM(None, w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise, device=DEV)
    k = gp.kernel_matrix(x)
    w_pad, wt_pad, diag, info = _factor_through_cabi(k)
    assert int(info.abs().max()) == 0
    wm = w_pad[:, :n, :n]
    assert torch.equal(wt_pad, w_pad.transpose(1, 2))
    eye = torch.eye(n, dtype=torch.float64, device=DEV)
    for o in range(w.n_s):
        assert float((wm[o] @ k[o] @ wm[o].t() - eye).abs().max()) < 1e-8
        assert float((wm[o].t() @ (wm[o] @ k[o]) - eye).abs().max()) < 1e-8
    sign, logdet = torch.linalg.slogdet(k)
    assert bool((sign > 0).all())
    assert float(((2 * diag[:, :n].log().sum(-1) - logdet) / logdet.abs()).abs().max()) < 1e-10
