"""CPU: the oracle (oracle/oracle.py) against the golden vectors produced by the VERBATIM reference
(tests/golden/make_golden.py) and against the reference's own known-answer tests."""
import numpy as np
import pytest
import torch

from oracle import oracle

T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))


def test_remainder_overapproximation_matches_reference(golden_algebra):
    g = golden_algebra  # test_utils.py:35-59, including the non-symmetric q
    for i in range(2):
        u_mu, u_sigma = oracle.compute_remainder_overapproximations(T(g['rem_q'][i:i + 1]), T(g['rem_k_fb'][i]),
                                                                    T(g['rem_l_mu']), T(g['rem_l_sigma']))
        assert np.allclose(u_mu.numpy()[0], g['rem_u_mu'][i], rtol=1e-12)
        assert np.allclose(u_sigma.numpy()[0], g['rem_u_sigma'][i], rtol=1e-12)
    # value quoted in SURVEY.md 8c for case 1
    assert np.allclose(g['rem_u_mu'][0], [4.70869342, 23.54346712])
    assert np.allclose(g['rem_u_sigma'][0], [0.43399048, 0.65098572])


def test_sum_and_rectangle_match_reference(golden_algebra):
    g = golden_algebra
    p, q = oracle.sum_two_ellipsoids(T(g['sum_p_a']), T(g['sum_q_a']), T(g['sum_p_b']), T(g['sum_q_b']))
    assert np.allclose(p.numpy(), g['sum_p'], rtol=1e-14) and np.allclose(q.numpy(), g['sum_q'], rtol=1e-14)
    assert np.allclose(oracle.ellipsoid_from_rectangle(T(g['rect_ub'])).numpy(), g['rect_q'], rtol=1e-15)
    with pytest.raises(AssertionError):  # test_utils_ellipsoid.py:106-125: u_b <= 0 is rejected
        oracle.ellipsoid_from_rectangle(T([[0.1, -0.2]]))


def test_polytope_checks_match_reference(golden_algebra):
    g = golden_algebra
    d = oracle.lin_ellipsoid_safety_distance(T(g['sd_p']), T(g['sd_q']), T(g['sd_h_mat']), T(g['sd_h_vec']))
    assert np.allclose(d.numpy(), g['sd_d'], rtol=1e-14)
    inside = oracle.is_ellipsoid_inside_polytope(T(g['sd_p']), T(g['sd_q']), T(g['sd_h_mat']), T(g['sd_h_vec']))
    assert np.array_equal(inside.numpy(), g['sd_inside'])
    # test_gp_reachability_pytorch.py:183-219: (5,5) inside; (0,0) touching -> unsafe; (20,20) outside
    assert inside.numpy()[3] and not inside.numpy()[0] and not inside.numpy()[4]


@pytest.mark.parametrize('tag', ['untrained', 'trained'])
@pytest.mark.parametrize('lin', ['lin', 'nolin'])
@pytest.mark.parametrize('branch', ['ell', 'pt'])
def test_onestep_matches_reference(golden_onestep, tag, lin, branch):
    g = golden_onestep
    gp = oracle.OracleGpCemSSM(2, 1, g[f'{tag}_lengthscale'], g[f'{tag}_outputscale'], g[f'{tag}_noise'])
    gp.update_model(T(g['x_train']), T(g['y_train']), replace_old=True)
    a, b = (T(g['a']), T(g['b'])) if lin == 'lin' else (None, None)
    q = T(g['q']) if branch == 'ell' else None
    p1, q1, sig = oracle.onestep_reachability(T(g['p']), gp, T(g['u']), T(g['l_mu']), T(g['l_sigma']), q, T(g['k_fb']),
                                              float(g['c_safety']), a, b)
    key = f'{tag}_{lin}_{branch}'
    assert np.allclose(p1.numpy(), g[f'{key}_p1'], rtol=1e-12, atol=0)
    assert np.allclose(q1.numpy(), g[f'{key}_q1'], rtol=1e-11, atol=0)
    assert np.allclose(sig.numpy(), g[f'{key}_sigma'], rtol=1e-12, atol=0)
    # third return value == predictive variance (test_gp_reachability_pytorch.py:139-159)
    _, var = gp.predict_without_jacobians(T(g['p']), T(g['u']))
    assert np.allclose(sig.numpy(), var.numpy(), rtol=1e-13)


def test_gp_jacobian_matches_autograd(golden_onestep):
    g = golden_onestep
    gp = oracle.OracleGpCemSSM(2, 1, g['trained_lengthscale'], g['trained_outputscale'], g['trained_noise'])
    gp.update_model(T(g['x_train']), T(g['y_train']), replace_old=True)
    z = torch.cat((T(g['p'][:4]), T(g['u'][:4])), dim=1)
    jac_auto = torch.autograd.functional.jacobian(lambda zz: gp.predict_raw(zz)[0].t(), z)  # [N, n_s, N, d]
    jac_auto = torch.stack([jac_auto[i, :, i, :] for i in range(4)])
    _, _, jac = gp.predict_with_jacobians(z[:, :2], z[:, 2:])
    assert torch.allclose(jac, jac_auto, rtol=1e-9, atol=1e-12)


def test_multistep_matches_reference_numpy(golden_multistep):
    g = golden_multistep
    gp = oracle.OracleGpCemSSM(2, 1, g['lengthscale'], g['outputscale'], g['noise'])
    gp.update_model(T(g['x_train']), T(g['y_train']), replace_old=True)
    _, _, p_all, q_all, _ = oracle.multistep_reachability(T(g['p0']), gp, T(g['k_fb']), T(g['k_ff']), T(g['l_mu']),
                                                          T(g['l_sigma']), None, float(g['c_safety']), T(g['a']),
                                                          T(g['b']))
    assert np.allclose(p_all.numpy(), g['p_all'], rtol=1e-11, atol=0)
    assert np.allclose(q_all.numpy(), g['q_all'], rtol=1e-10, atol=0)


def _cem_problem(g):
    gp = oracle.OracleGpCemSSM(2, 1, g['lengthscale'], g['outputscale'], g['noise_var'])
    gp.update_model(T(g['x_train']), T(g['y_train']), replace_old=True)
    return oracle.SafeCemProblem(gp, g['l_mu'], g['l_sigma'], g['k_fb'], g['a'], g['b'], float(g['beta']), g['h_mat'],
                                 g['h_vec'], g['act_a'], g['act_b'])


@pytest.mark.parametrize('loops', [False, True])
def test_cem_solve_matches_reference(golden_cem, loops):
    g = golden_cem
    res = oracle.cem_solve(_cem_problem(g), T(g['p0']), T(g['eps']), int(g['num_elites']), loops=loops)
    assert np.allclose(res['means'].numpy(), g['means'], rtol=1e-10, atol=1e-13)
    assert np.allclose(res['stds'].numpy(), g['stds'], rtol=1e-10, atol=1e-13)
    for i, r in enumerate(res['rollouts']):
        assert np.array_equal(r.constraint_costs.numpy(), g['constraint'][i])
        assert np.allclose(r.objective_costs.numpy(), g['objective'][i], rtol=1e-10, atol=0)
        assert np.allclose(r.trajectories.numpy(), g['trajectories'][i], rtol=1e-9, atol=1e-14)
    assert np.array_equal(res['elite_ids'].numpy(), g['elite_ids'])
    assert bool(g['found']) and np.allclose(res['actions'].numpy(), g['actions'], rtol=1e-10, atol=1e-13)


def test_select_elites_known_answers():
    """constrained_cem_mpc_test.py:230-301: infeasible -> lowest constraint cost (mean 2, std of [1,3]);
    feasible -> lowest objective among feasible (mean 6, std of [10,2])."""
    actions = T([[[1.]], [[5.]], [[3.]], [[7.]]])
    ids = oracle.select_elites(T([1., 1., 1., 1.]), T([1., 5., 2., 9.]), 2)
    mean, std = oracle.refit(actions, ids)
    assert mean.item() == 2.0 and np.isclose(std.item(), np.std([1, 3], ddof=1))
    actions = T([[[10.]], [[4.]], [[2.]], [[8.]]])
    ids = oracle.select_elites(T([1., 0.5, 2., 9.]), T([0., 3., 0., 0.]), 2)
    mean, std = oracle.refit(actions, ids)
    assert mean.item() == 6.0 and np.isclose(std.item(), np.std([10, 2], ddof=1))


def test_pick_quirk_q3_first_feasible():
    cc, obj = T([3., 0., 0., 0.]), T([1., 10., 5., 2.])
    assert oracle.pick_first_feasible(cc) == 1  # what the reference's expression returns
    assert oracle.pick_best_feasible(obj, cc) == 3
    assert oracle.pick_first_feasible(T([1., 2.])) == -1


def test_action_constraint_known_answer():
    """test_safempc_cem.py:59-71: normalised CartPole box [-0.5, 0.5]; 0.2 inside, -5 and 6 outside -> cost 6."""
    a_mat, b_vec = T([[1.], [-1.]]), T([0.5, 0.5])
    assert oracle.action_constraint_costs(T([[[0.2], [-5.], [6.]]]), a_mat, b_vec).item() == 6.0
    assert oracle.action_constraint_costs(T([[[0.5], [-0.5]]]), a_mat, b_vec).item() == 0.0  # boundary is inside


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    f = lambda c, k: oracle.philox4x32_10(np.array(c, dtype=np.uint32), np.array(k, dtype=np.uint32)).tolist()
    assert f([0] * 4, [0] * 2) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert f([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert f([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_oracle_mll_gradient_against_finite_differences():
    """The oracle's marginal-likelihood gradient (autograd through the Cholesky factorisation, the route the
    reference's loss.backward() takes, gp_ssm_cem.py:121) agrees with central differences; its Adam loop lowers the
    loss.  (The RBF-ARD arithmetic itself is gpytorch's: parity unpinned, see oracle.py.)"""
    g = torch.Generator().manual_seed(0)
    x = torch.rand((40, 3), generator=g, dtype=torch.float64) * 2 - 1
    y = torch.stack((torch.sin(2 * x[:, 0]), 0.3 * x[:, 1] * x[:, 2]), dim=1) + 0.05 * torch.randn((40, 2), generator=g, dtype=torch.float64)
    ls = torch.full((2, 3), 0.7, dtype=torch.float64)
    os_ = torch.tensor([0.5, 0.2], dtype=torch.float64)
    nz = torch.tensor([0.01, 0.02], dtype=torch.float64)
    mll, grad = oracle.exact_mll_and_grad(x, y, ls, os_, nz)
    theta = torch.cat((ls, os_.unsqueeze(1), nz.unsqueeze(1)), dim=1)
    for o in range(2):
        for c in range(5):
            e = torch.zeros_like(theta)
            e[o, c] = 1e-6
            hi, lo = theta + e, theta - e
            fd = (oracle.exact_mll(x, y, hi[:, :3], hi[:, 3], hi[:, 4])[o] -
                  oracle.exact_mll(x, y, lo[:, :3], lo[:, 3], lo[:, 4])[o]) / 2e-6
            assert abs(fd - grad[o, c]) <= 1e-6 * max(1.0, abs(grad[o, c]))
    _, _, _, losses = oracle.train_hyperparameters(x, y, ls, os_, nz, 30)
    assert losses[-1] < losses[0]
