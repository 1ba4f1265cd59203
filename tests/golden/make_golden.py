"""Generate the golden vectors of tests/golden/ by EXECUTING THE VERBATIM REFERENCE (oracle/ref_shim.py).

Run in the build container only (`python tests/golden/make_golden.py`): /root/reference does not exist on the GPU box.
The one restated component is the GP (`oracle.OracleGpCemSSM`; gpytorch 0.3.2 is unavailable, see oracle/oracle.py);
everything else -- onestep_reachability (torch and NumPy twins), multistep_reachability, the remainder bound, the
ellipsoid sums, the polytope checks, ConstrainedCemMpc and CemSafeMPC -- is the reference's own code.  Every array the
tests need (inputs included) is stored, so the fixtures are self-contained.
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle, ref_shim  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_default_dtype(torch.double)
T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))


class NumpySsm:
    """What test_gp_reachability_pytorch.py:21-71 (CemSSMNumpyWrapper) does: a CemSSM seen through the NumPy
    StateSpaceModel call convention ssm(states [1,n_s], actions [1,n_u]) -> (mu [n_s,1], sigma [n_s], jac [n_s,d])."""

    def __init__(self, ssm):
        self._ssm = ssm
        self.num_states, self.num_actions = ssm.num_states, ssm.num_actions

    def __call__(self, states, actions):
        p, s, j = self._ssm.predict_with_jacobians(T(states), T(actions))
        return p.t().numpy(), s.squeeze(0).numpy(), j.squeeze(0).numpy()


def ellipsoid_algebra(ns):
    out = {}
    # test_utils.py:35-59 (note q_2 is non-symmetric)
    q = torch.tensor([[[2.0, 1.0], [1.0, 3.0]], [[10.0, 9.0], [0.5, 2.9]]])
    k_fb = torch.tensor([[[1.0, 5.0], [5.0, 9.0]], [[1.1, 5.2], [4.5, 6.5]]])
    l_mu, l_sigma = torch.tensor([0.01, 0.05]), torch.tensor([0.02, 0.03])
    u_pt, s_pt = ns.utils.compute_remainder_overapproximations_pytorch(q, k_fb, l_mu.repeat(2, 1), l_sigma.repeat(2, 1))
    u_np = np.stack([ns.utils.compute_remainder_overapproximations(q[i].numpy(), k_fb[i].numpy(), l_mu.numpy(),
                                                                   l_sigma.numpy())[0] for i in range(2)])
    s_np = np.stack([ns.utils.compute_remainder_overapproximations(q[i].numpy(), k_fb[i].numpy(), l_mu.numpy(),
                                                                   l_sigma.numpy())[1] for i in range(2)])
    assert np.allclose(u_pt.numpy(), u_np) and np.allclose(s_pt.numpy(), s_np)
    out.update(rem_q=q.numpy(), rem_k_fb=k_fb.numpy(), rem_l_mu=l_mu.numpy(), rem_l_sigma=l_sigma.numpy(),
               rem_u_mu=u_pt.numpy(), rem_u_sigma=s_pt.numpy())
    # test_utils_ellipsoid.py:19-50
    p_a = torch.tensor([[0.0, 1.0], [10.1, 11.1], [100.1, 101.1]])
    q_a = torch.tensor([[[2.0, 1.0], [1.0, 2.0]], [[12.1, 11.1], [11.1, 12.1]], [[102.1, 101.1], [101.1, 102.1]]])
    p_b = torch.tensor([[2.0, 2.0], [12.1, 12.1], [102.1, 102.1]])
    q_b = torch.tensor([[[4.0, 1.0], [1.0, 2.0]], [[14.1, 11.1], [11.1, 12.1]], [[104.1, 101.1], [101.1, 102.1]]])
    p_s, q_s = ns.utils_ellipsoid.sum_two_ellipsoids_pytorch(p_a, q_a, p_b, q_b)
    for i in range(3):
        pn, qn = ns.utils_ellipsoid.sum_two_ellipsoids(p_a[i].numpy(), q_a[i].numpy(), p_b[i].numpy(), q_b[i].numpy())
        assert np.allclose(pn, p_s[i].numpy()) and np.allclose(qn, q_s[i].numpy())
    out.update(sum_p_a=p_a.numpy(), sum_q_a=q_a.numpy(), sum_p_b=p_b.numpy(), sum_q_b=q_b.numpy(), sum_p=p_s.numpy(),
               sum_q=q_s.numpy())
    # test_utils_ellipsoid.py:106-125
    ub = torch.tensor([[0.1, 0.3, 0.5], [1.0, 1.0, 1.0], [2.0, 0.2, 0.7]])
    out.update(rect_ub=ub.numpy(), rect_q=ns.utils_ellipsoid.ellipsoid_from_rectangle_pytorch(ub).numpy())
    # test_gp_reachability_pytorch.py:162-219
    p = torch.tensor([[0., 0.], [1., 3.], [5., 10.], [5., 5.], [20., 20.]])
    qq = torch.cat((.2 * torch.tensor([[[.6, .21], [.21, .55]], [[.5, .2], [.2, .65]], [[.7, .21], [.21, .59]]]),
                    torch.tensor([[[2., 1.], [1., 2.]], [[2., 1.], [1., 2.]]])))
    a_box = T(np.vstack((np.eye(2), -np.eye(2))))
    b_box = T([[10.], [10.], [0.], [0.]])
    d = ns.reach_pt.lin_ellipsoid_safety_distance(p, qq, a_box, b_box)
    for i in range(5):
        dn = ns.reach_np.lin_ellipsoid_safety_distance(p[i].unsqueeze(1).numpy(), qq[i].numpy(), a_box.numpy(),
                                                       b_box.numpy())
        assert np.allclose(dn.squeeze(1), d[i].numpy())
    inside = ns.reach_pt.is_ellipsoid_inside_polytope(p, qq, a_box, b_box)
    # known answers of test_gp_reachability_pytorch.py:183-219: (5,5) inside, (0,0) and (20,20) not
    assert bool(inside[3]) and not bool(inside[0]) and not bool(inside[4])
    out.update(sd_p=p.numpy(), sd_q=qq.numpy(), sd_h_mat=a_box.numpy(), sd_h_vec=b_box.numpy(), sd_d=d.numpy(),
               sd_inside=inside.numpy())
    np.savez(os.path.join(OUT, 'ellipsoid_algebra.npz'), **out)


def onestep(ns, data):
    """test_gp_reachability_pytorch.py:75-136 with the GP restated; untrained (ln 2) and trained-like hyper-parameters,
    both branches, torch batch vs NumPy single."""
    x, y = data['X'], data['y']
    rng = np.random.default_rng(125)
    n_s, n_u, c_safety = 2, 1, 2.0
    out = dict(x_train=x, y_train=y)
    hypers = {'untrained': (np.full((2, 3), np.log(2.0)), np.full(2, np.log(2.0)), np.full(2, np.log(2.0))),
              'trained': (np.array([[0.9, 0.5, 1.1], [1.2, 0.4, 0.8]]), np.array([0.05, 0.02]), np.array([1e-4, 3e-4]))}
    a, b, k_fb = rng.random((n_s, n_s)), rng.random((n_s, n_u)), rng.random((n_u, n_s))
    l_mu = l_sigma = np.full(n_s, 0.001)
    n = 34
    p = .1 * rng.random((n, n_s))
    u = rng.random((n, n_u))
    m = rng.standard_normal((n, n_s, n_s))
    q = .05 * (m @ m.transpose(0, 2, 1) / n_s + .2 * np.eye(n_s))
    q[:2] = .2 * np.array([[[.6, .21], [.21, .55]], [[.5, .2], [.2, .65]]])  # the reference test's ellipsoids
    out.update(a=a, b=b, k_fb=k_fb, l_mu=l_mu, l_sigma=l_sigma, c_safety=c_safety, p=p, q=q, u=u)
    for tag, (ls, os_, nz) in hypers.items():
        gp = oracle.OracleGpCemSSM(n_s, n_u, ls, os_, nz)
        gp.update_model(T(x), T(y), replace_old=True)
        mean, var, jac = gp.predict_with_jacobians(T(p), T(u))
        out.update({f'{tag}_lengthscale': ls, f'{tag}_outputscale': os_, f'{tag}_noise': nz, f'{tag}_mean': mean.numpy(),
                    f'{tag}_var': var.numpy(), f'{tag}_jac': jac.numpy()})
        for lin in (True, False):
            aa, bb = (T(a), T(b)) if lin else (None, None)
            for branch, qq in (('ell', T(q)), ('pt', None)):
                p1, q1, sig = ns.reach_pt.onestep_reachability(T(p), gp, T(u), T(l_mu), T(l_sigma), qq, T(k_fb),
                                                               c_safety, verbose=0, a=aa, b=bb)
                for i in range(2):  # NumPy twin, one ellipsoid at a time (:121-136)
                    pn, qn = ns.reach_np.onestep_reachability(
                        p[i][:, None], NumpySsm(gp), u[i][None, :], l_mu, l_sigma, None if qq is None else q[i], k_fb,
                        c_safety, verbose=0, a=a if lin else None, b=b if lin else None)
                    assert np.allclose(pn.squeeze(), p1[i].numpy()) and np.allclose(qn, q1[i].numpy())
                # the restated oracle must agree with the verbatim reference
                po, qo, so = oracle.onestep_reachability(T(p), gp, T(u), T(l_mu), T(l_sigma), qq, T(k_fb), c_safety,
                                                         aa, bb)
                assert torch.allclose(po, p1, rtol=1e-12, atol=0) and torch.allclose(qo, q1, rtol=1e-11, atol=0)
                key = f'{tag}_{"lin" if lin else "nolin"}_{branch}'
                out.update({f'{key}_p1': p1.numpy(), f'{key}_q1': q1.numpy(), f'{key}_sigma': sig.numpy()})
    np.savez(os.path.join(OUT, 'onestep_invpend.npz'), **out)


def multistep(ns, data):
    """NumPy multistep_reachability (gp_reachability.py:168-221), T = 3, per-step feedback gains, 3 start states."""
    x, y = data['X'], data['y']
    rng = np.random.default_rng(7)
    n_s, n_u, t_len = 2, 1, 3
    ls, os_, nz = np.array([[0.9, 0.5, 1.1], [1.2, 0.4, 0.8]]), np.array([0.05, 0.02]), np.array([1e-4, 3e-4])
    gp = oracle.OracleGpCemSSM(n_s, n_u, ls, os_, nz)
    gp.update_model(T(x), T(y), replace_old=True)
    a = np.array([[1.0, 0.1], [0.05, 0.9]])
    b = np.array([[0.2], [0.05]])
    k_fb = -0.3 * rng.random((t_len - 1, n_u, n_s))
    k_ff = 0.4 * rng.random((3, t_len, n_u)) - 0.2
    p0 = 0.1 * rng.random((3, n_s))
    l_mu, l_sigma, c_safety = np.array([0.05, 0.02]), np.array([0.05, 0.02]), 2.0
    p_all, q_all = [], []
    for i in range(3):
        _, _, pa, qa = ns.reach_np.multistep_reachability(p0[i][:, None], NumpySsm(gp), k_fb, k_ff[i], l_mu, l_sigma,
                                                          None, c_safety, 0, a, b, None)
        p_all.append(pa), q_all.append(qa)
    p_all, q_all = np.stack(p_all), np.stack(q_all)
    _, _, po, qo, _ = oracle.multistep_reachability(T(p0), gp, T(k_fb), T(k_ff), T(l_mu), T(l_sigma), None, c_safety,
                                                    T(a), T(b))
    assert np.allclose(po.numpy(), p_all, rtol=1e-11, atol=0) and np.allclose(qo.numpy(), q_all, rtol=1e-10, atol=0)
    np.savez(os.path.join(OUT, 'multistep_invpend.npz'), x_train=x, y_train=y, lengthscale=ls, outputscale=os_,
             noise=nz, a=a, b=b, k_fb=k_fb, k_ff=k_ff, p0=p0, l_mu=l_mu, l_sigma=l_sigma, c_safety=c_safety,
             p_all=p_all, q_all=q_all)


def cem(ns, data):
    """The verbatim CemSafeMPC + ConstrainedCemMpc on the pendulum with RNG-fixed candidates."""
    import importlib
    safempc_cem = importlib.import_module('safe_exploration.safempc_cem')
    envs = importlib.import_module('safe_exploration.environments.environments')
    x, y = data['X'], data['y']
    rng = np.random.default_rng(11)
    n_s, n_u = 2, 1
    num_iter, n_roll, t_len, n_elite = 4, 96, 4, 12
    ls, os_, nz = np.array([[0.9, 0.5, 1.1], [1.2, 0.4, 0.8]]), np.array([0.0025, 0.0016]), np.array([1e-5, 2e-5])
    gp = oracle.OracleGpCemSSM(n_s, n_u, ls, os_, nz)
    env = envs.InvertedPendulum(verbosity=0, simple_constraints=False)
    a_lin, b_lin = env.linearize_discretize()
    gp.update_model(T(x), T(0.1 * y), replace_old=True)
    conf = types.SimpleNamespace(device='cpu', plot_cem_optimisation=False, plot_cem_terminal_states=False,
                                 mpc_time_horizon=t_len, use_prior_model=True, cem_num_rollouts=n_roll,
                                 cem_num_elites=n_elite, cem_num_iterations=num_iter, use_state_constraint=True)
    wx, wu = np.diag([1., 2.]), 25 * np.eye(1)
    constraints = safempc_cem.construct_constraints(conf, env)
    solver = safempc_cem.CemSafeMPC(gp, constraints, env, conf, {'lin_model': (a_lin, b_lin)}, wx, wu, beta_safety=2.0,
                                    safe_policy=lambda s: np.zeros(1))
    noise = rng.standard_normal((num_iter, n_roll, t_len, n_u))
    noise[0, :8] *= 0.05  # make sure a few candidates stay inside the action box in the first iteration
    calls = {'i': 0}
    elites = []

    def fixed_sample(self, sample_shape=torch.Size()):
        i = calls['i']
        calls['i'] += 1
        return self.loc + self.scale * T(noise[i])

    orig_sample = torch.distributions.Normal.sample
    orig_select = ns.cem.ConstrainedCemMpc._select_elites

    def spy_select(self, rollouts):
        res = orig_select(self, rollouts)
        elites.append((res.actions.mean(dim=0).numpy().copy(), res.actions.std(dim=0).numpy().copy(),
                       res.objective_costs.numpy().copy(), res.constraint_costs.numpy().copy()))
        return res

    # The reference sorts with torch's default (unstable) sort and constraint costs are small integers, so its own
    # pick among ties is unspecified (verified: x.sort() != x.sort(stable=True) on ties with torch 2.11).  The golden
    # run pins the order the way the drop-in defines it -- ascending (cost, index) -- by making that sort stable.
    orig_sort = torch.Tensor.sort

    def stable_sort(self, *a, **k):
        k.setdefault('stable', True)
        return orig_sort(self, *a, **k)

    torch.Tensor.sort = stable_sort
    torch.distributions.Normal.sample = fixed_sample
    ns.cem.ConstrainedCemMpc._select_elites = spy_select
    try:
        p0 = np.array([0.02, -0.01])
        state = solver._pq_flattener.flatten(T(p0).unsqueeze(0), None)
        actions, rollouts = solver._mpc.get_actions(state)
    finally:
        torch.Tensor.sort = orig_sort
        torch.distributions.Normal.sample = orig_sample
        ns.cem.ConstrainedCemMpc._select_elites = orig_select
    k_fb = solver._lqr.get_control_matrix()
    h_mat, h_safe, _, _ = env.get_safety_constraints(normalize=True)
    act_box = constraints[0]._self_region
    out = dict(x_train=x, y_train=0.1 * y, lengthscale=ls, outputscale=os_, noise_var=nz, a=a_lin, b=b_lin, k_fb=k_fb,
               l_mu=env.l_mu, l_sigma=env.l_sigm, beta=2.0, h_mat=h_mat, h_vec=h_safe.reshape(-1),
               act_a=act_box._A.numpy(), act_b=act_box._b.numpy(), p0=p0, eps=noise, num_elites=n_elite,
               found=actions is not None, actions=np.zeros((t_len, n_u)) if actions is None else actions.numpy(),
               trajectories=np.stack([r.trajectories.numpy() for r in rollouts]),
               rollout_actions=np.stack([r.actions.numpy() for r in rollouts]),
               objective=np.stack([r.objective_costs.numpy() for r in rollouts]),
               constraint=np.stack([r.constraint_costs.numpy() for r in rollouts]),
               means=np.stack([e[0] for e in elites]), stds=np.stack([e[1] for e in elites]))
    # the restated oracle must reproduce the verbatim run
    prob = oracle.SafeCemProblem(gp, env.l_mu, env.l_sigm, k_fb, a_lin, b_lin, 2.0, h_mat, h_safe, act_box._A.numpy(),
                                 act_box._b.numpy())
    res = oracle.cem_solve(prob, T(p0), T(noise), n_elite)
    assert np.allclose(res['means'].numpy(), out['means'], rtol=1e-10, atol=1e-13), 'oracle means != reference'
    assert np.allclose(res['stds'].numpy(), out['stds'], rtol=1e-10, atol=1e-13)
    for i in range(num_iter):
        assert np.array_equal(res['rollouts'][i].constraint_costs.numpy(), out['constraint'][i])
        assert np.allclose(res['rollouts'][i].objective_costs.numpy(), out['objective'][i], rtol=1e-10, atol=0)
        assert np.allclose(res['rollouts'][i].trajectories.numpy(), out['trajectories'][i], rtol=1e-9, atol=1e-14)
    assert (res['actions'] is None) == (actions is None)
    if actions is not None:
        assert np.allclose(res['actions'].numpy(), actions.numpy(), rtol=1e-10, atol=1e-13)
    out['elite_ids'] = res['elite_ids'].numpy()
    out['best_id'] = res['best_id']
    print('cem: found =', actions is not None, 'feasible per iteration =',
          [(c == 0).sum() for c in out['constraint']], 'best', res['best_id'])
    np.savez(os.path.join(OUT, 'cem_pendulum.npz'), **out)


def main():
    assert ref_shim.reference_available(), 'run this where /root/reference exists'
    ns = ref_shim.load()
    data = np.load(os.path.join(ref_shim.REFERENCE_ROOT, 'safe_exploration', 'test', 'invpend_data.npz'))
    data = {'X': np.asarray(data['X'], dtype=np.float64), 'y': np.asarray(data['y'], dtype=np.float64)}
    ellipsoid_algebra(ns)
    onestep(ns, data)
    multistep(ns, data)
    cem(ns, data)
    print('golden vectors written to', OUT)


if __name__ == '__main__':
    main()
