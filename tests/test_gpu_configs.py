"""GPU parity on the shapes of BASELINE.json's own configs that the pendulum-sized tests do not reach.

  * C3 shape: a whole CartPole safe CEM-MPC solve (n_s = 4, 9-face polytope, Jacobi eigenvalue path, tcgen05 variance
    GEMM inside the CEM loop) against `oracle.cem_solve` -- constrained_cem_mpc.py:222-280 + safempc_cem.py:288-312 at
    n_s = 4 (environments.py:1062-1176 polytope).
  * C5 shapes: `se_reach_multistep` at n_s in {4, 6}, n_train in {1024, 4096}, T = 10 against
    `oracle.multistep_reachability` (gp_reachability.py:168-221) on 64 sampled ellipsoids.
  * the largest models: 64 candidates at n = 5000 / n_s = 4 (the C3 model) and a posterior check at n = 16384; the int32
    level sums of the digit GEMM at their worst case (every digit -128, K = 16384).
  * switches nothing else exercises on the device: use_prior_model = False (quirk Q9 zeros), the two environment
    objectives (safempc_cem.py:304-312).

Tolerances: float64 1e-9 relative on means / Jacobians / centres, 1e-8 on the variance and on shape matrices for the
automatic digit choice; integer work (elites, constraint costs, picks, level sums) bit-exact.
"""
import numpy as np
import pytest
import torch

import safe_exploration_b200 as se
from oracle import oracle
from safe_exploration_b200 import _lib, workloads

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))
G = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64)).to(device=DEV)


def rel_err(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = np.maximum(np.abs(b), 0 if scale is None else scale)
    den = np.where(den == 0, 1.0, den)
    return float(np.max(np.abs(a - b) / den)) if a.size else 0.0


def oracle_problem(w, **kw):
    gp = oracle.OracleGpCemSSM(w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    return oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                 w.act_b, **kw)


def philox_noise(seed, w):
    return np.stack([oracle.philox_normal_noise(seed, i, 0, w.num_rollouts, w.time_horizon, w.n_u)
                     for i in range(w.num_iterations)])


def assert_solve_equals_oracle(res, ref, traj_tol=1e-7):
    assert np.array_equal(res['elite_ids'][:, 0].cpu().numpy(), ref['elite_ids'].numpy())  # integer work: bit-exact
    for i, r in enumerate(ref['rollouts']):
        assert np.array_equal(res['constraint'][i, 0].cpu().numpy(), r.constraint_costs.numpy())
        assert rel_err(res['objective'][i, 0].cpu(), r.objective_costs) < 1e-8
        assert rel_err(res['trajectories'][i, 0].cpu(), r.trajectories, 1e-6) < traj_tol
    assert rel_err(res['means'][:, 0].cpu(), ref['means'], 1e-3) < 1e-8
    assert rel_err(res['stds'][:, 0].cpu(), ref['stds'], 1e-3) < 1e-8
    assert int(res['best_id'][0]) == ref['best_id']


# ---- C3 shape: whole solve ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('n_train,t_len,iters,elites', [(400, 15, 3, 32), (600, 8, 6, 4)])
def test_cartpole_cem_solve_matches_oracle_on_the_tensor_core_path(n_train, t_len, iters, elites):
    """(400, T = 15): BASELINE C3's horizon -- no candidate is feasible there, so elites are chosen by constraint cost
    (ties everywhere: costs are 10 + 3 k); (600, T = 8, E = 4): feasible rollouts appear from the third iteration on and
    the last iterations select by objective -- both branches of _select_elites (constrained_cem_mpc.py:241-257)."""
    w = workloads.cartpole_workload(n_train=n_train, num_rollouts=1024, time_horizon=t_len, num_elites=elites,
                                    num_iterations=iters, name='C3-small')
    ssm, mpc = workloads.build_solver(w, record='full', seed=5)
    res = mpc.solve(G(w.p0).unsqueeze(0), seed=5)
    assert ssm.gemm_info()[0] == 'tcgen05-i8'
    ref = oracle.cem_solve(oracle_problem(w), T(w.p0), T(philox_noise(5, w)), w.num_elites)
    # elementwise on [p | vec q] with a floor of 1e-6: the 1e-9-class variance error of the 6-plane GEMM compounds over
    # the horizon through the squared radius (measured 1.2e-7 at T = 8), so 1e-6 here; decisions stay bit-exact
    assert_solve_equals_oracle(res, ref, traj_tol=1e-6)
    feasible = [(r.constraint_costs == 0).sum().item() for r in ref['rollouts']]
    if t_len == 8:
        assert feasible[-1] >= elites and feasible[0] < elites  # both selection modes occurred
        assert ref['best_id'] >= 0


# ---- C5 shapes: multi-step propagation -----------------------------------------------------------------------------------
@pytest.mark.parametrize('n_s,n_train', [(4, 1024), (6, 1024), (4, 4096), (6, 4096)])
def test_multistep_c5_shapes_match_oracle(n_s, n_train):
    w = workloads.reach_microbench_workload(n_train=n_train, n_s=n_s, num_ellipsoids=64, time_horizon=10)
    rng = np.random.default_rng(40 + n_s)
    N, t_len = 64, 10
    p0 = 0.1 * rng.standard_normal((N, n_s))
    a_ = rng.standard_normal((N, n_s, n_s))
    # SURVEY 8(d)'s C5 shapes with a smaller initial ellipsoid (2e-3 instead of 0.2): with l_mu = 0.05 the remainder term
    # squares the radius every step, and from 0.2 the shapes reach 1e155 by step 10, where nothing is being compared
    q0 = 2e-3 * (a_ @ a_.transpose(0, 2, 1) / n_s + 0.1 * np.eye(n_s))
    u = rng.uniform(-1, 1, (N, t_len, 1))
    ssm = se.B200GpCemSSM(None, n_s, 1, w.lengthscale, w.outputscale, w.noise, device=DEV)
    ssm.update_model(G(w.x_train), G(w.y_train), replace_old=True)
    pT, qT, p_all, q_all, s_all = se.multistep_reachability(G(p0), ssm, w.k_fb, G(u), w.l_mu, w.l_sigma, G(q0),
                                                            c_safety=w.beta_safety, a=w.a, b=w.b)
    assert ssm.gemm_info()[0] == 'tcgen05-i8'
    gp = oracle.OracleGpCemSSM(n_s, 1, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    rp, rq, rp_all, rq_all, rs_all = oracle.multistep_reachability(T(p0), gp, T(w.k_fb), T(u), T(w.l_mu), T(w.l_sigma),
                                                                   T(q0), w.beta_safety, T(w.a), T(w.b), T(w.k_fb))
    assert rel_err(p_all.cpu(), rp_all, float(rp_all.abs().max())) < 1e-9
    assert rel_err(s_all.cpu(), rs_all) < 1e-8
    for t in range(t_len):  # shape matrices grow over the horizon: relative to each step's own scale; the 1e-9-class
        # variance error of every step feeds the next radius, so the bound is 1e-8 early and 1e-7 at the end
        tol = 1e-8 if t < 4 else 1e-7
        assert rel_err(q_all[:, t].cpu(), rq_all[:, t], float(rq_all[:, t].abs().max()) * 1e-3) < tol, t
    assert rel_err(qT.cpu(), rq, float(rq.abs().max()) * 1e-3) < 1e-7 and rel_err(pT.cpu(), rp, 1e-2) < 1e-8


# ---- the largest models ------------------------------------------------------------------------------------------------
def test_spot_check_at_the_full_c3_model_size():
    """n = 5000, n_s = 4 (BASELINE C3's model): 64 candidates through one ellipsoid step against the oracle."""
    w = workloads.cartpole_workload(num_rollouts=64)
    assert w.n_train == 5000
    rng = np.random.default_rng(8)
    N = 64
    p = 0.05 * rng.standard_normal((N, 4))
    a_ = rng.standard_normal((N, 4, 4))
    q = 1e-3 * (a_ @ a_.transpose(0, 2, 1) / 4 + 0.1 * np.eye(4))
    u = rng.uniform(-0.5, 0.5, (N, 1))
    ssm = se.B200GpCemSSM(None, 4, 1, w.lengthscale, w.outputscale, w.noise, device=DEV)
    ssm.update_model(G(w.x_train), G(w.y_train), replace_old=True)
    p1, q1, sig = se.onestep_reachability(G(p), ssm, G(u), G(w.l_mu), G(w.l_sigma), G(q), k_fb=G(w.k_fb),
                                          c_safety=w.beta_safety, a=G(w.a), b=G(w.b))
    impl, digits = ssm.gemm_info()
    assert impl == 'tcgen05-i8'
    gp = oracle.OracleGpCemSSM(4, 1, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    rp, rq, rs = oracle.onestep_reachability(T(p), gp, T(u), T(w.l_mu), T(w.l_sigma), T(q), T(w.k_fb), w.beta_safety,
                                             T(w.a), T(w.b))
    assert rel_err(p1.cpu(), rp, float(rp.abs().max())) < 1e-9
    assert rel_err(sig.cpu(), rs) < 1e-8, digits
    assert rel_err(q1.cpu(), rq, float(rq.abs().max()) * 1e-3) < 1e-8


def _oracle_posterior_by_solves(x, y, ls, os_, nz, z):
    """The oracle's posterior (oracle.OracleGpCemSSM._posterior) without materialising L^-1: at n = 16384 the n
    right-hand sides of its `solve_triangular(chol, eye)` are the whole cost, the 128 query columns are all it needs."""
    k = oracle.rbf_ard(x, x, ls, os_)
    k.diagonal(dim1=1, dim2=2).add_(nz.view(-1, 1))
    chol = torch.linalg.cholesky(k)
    del k
    alpha = torch.cholesky_solve(y.t().unsqueeze(-1), chol).squeeze(-1)
    kstar = oracle.rbf_ard(z, x, ls, os_)  # [n_s, N, n]
    mean = torch.einsum('onk,ok->on', kstar, alpha)
    v = torch.linalg.solve_triangular(chol, kstar.transpose(1, 2), upper=False)  # L^-1 k*  [n_s, n, N]
    var = os_.view(-1, 1) - (v * v).sum(1) + nz.view(-1, 1)
    return mean.t(), var.t()


def test_posterior_at_n_16384_and_the_digit_plane_limits():
    """The largest training set the tcgen05 path takes.  6 planes (automatic at outputscale / noise = 10^3) and 7 planes
    against the oracle; 8 planes are refused there because the top level sum could leave int32 (8 x 2^14 x 16384 = 2^31)."""
    n, n_s, n_u = 16384, 1, 1
    rng = np.random.default_rng(77)
    x = rng.uniform(-1, 1, (n, 2))
    y = 0.1 * np.sin(3 * x[:, :1]) + 0.05 * x[:, :1] * x[:, 1:]
    ls, os_, nz = np.array([[0.08, 0.1]]), np.array([0.05]), np.array([5e-5])  # short length scales: K stays well conditioned
    z = rng.uniform(-1, 1, (128, 2))
    z[:64] = x[rng.integers(0, n, 64)] + 1e-4
    want_m, want_v = _oracle_posterior_by_solves(T(x), T(y), T(ls), T(os_), T(nz), T(z))
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV, incremental=False)
    gp.update_model(G(x), G(y), replace_old=True)
    for digits, tol in ((0, 1e-8), (7, 1e-8)):
        gp.set_tc_digits(digits)
        mean, var = gp.predict_without_jacobians(G(z[:, :1]), G(z[:, 1:]))
        impl, used = gp.gemm_info()
        assert impl == 'tcgen05-i8' and used == (digits or 6)
        assert rel_err(mean.cpu(), want_m, float(want_m.abs().max())) < 1e-9
        # the float64 factorisations of both sides carry cond(K) eps themselves at this size: 1e-7 on the variance
        assert rel_err(var.cpu(), want_v) < 1e-7, (digits, rel_err(var.cpu(), want_v))
    gp.set_tc_digits(8)
    with pytest.raises(_lib.SecemError, match='int32'):
        gp.predict_without_jacobians(G(z[:, :1]), G(z[:, 1:]))


@pytest.mark.parametrize('S,nk', [(7, 512), (8, 256), (6, 512)])
def test_int32_level_sums_at_the_worst_case(S, nk):
    """Every digit of both operands at -128 over the longest contraction each plane count is allowed (K = 32 nk):
    |D_l| = (l + 1) 2^14 K must come back exactly -- (S = 7, K = 16384): 7 x 2^28 < 2^31; (S = 8, K = 8192): 2^30."""
    lib = _lib.load()
    a = torch.full((nk * S * 128 * 32,), -128, dtype=torch.int8, device=DEV)
    b = torch.full((nk * S * 64 * 32,), -128, dtype=torch.int8, device=DEV)
    out = torch.zeros((S, 128, 64), dtype=torch.int32, device=DEV)
    _lib.check(lib.se_debug_tc_probe(_lib.ptr(a), _lib.ptr(b), nk, S, 128, 256, _lib.ptr(out), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    got = out.cpu().numpy().astype(np.int64)
    for l in range(S):
        assert (got[l] == (l + 1) * 16384 * 32 * nk).all(), (l, got[l, 0, 0])
    assert got.max() < 2 ** 31


# ---- switches --------------------------------------------------------------------------------------------------------
def test_cem_without_prior_model_uses_zeros_not_identity():
    """use_prior_model = False (safempc_cem.py:291-296, quirk Q9): a = b = 0, so p_1 = mu alone."""
    w = workloads.pendulum_workload(n_train=300, num_rollouts=512, time_horizon=4, num_elites=16, num_iterations=3)
    w.a, w.b = np.zeros_like(w.a), np.zeros_like(w.b)
    ssm, mpc = workloads.build_solver(w, record='full', seed=21)
    res = mpc.solve(G(w.p0 + 0.01).unsqueeze(0), seed=21)
    ref = oracle.cem_solve(oracle_problem(w, use_prior_model=False), T(w.p0 + 0.01), T(philox_noise(21, w)),
                           w.num_elites)
    assert_solve_equals_oracle(res, ref)
    # and through the solver class: conf.use_prior_model = False must produce the same device-side problem
    class Env:
        n_s, n_u, l_mu, l_sigm = 2, 1, w.l_mu, w.l_sigma
        objective_cost_function = staticmethod(lambda ps: None)

    class Conf:
        mpc_time_horizon, cem_num_rollouts, cem_num_elites, cem_num_iterations = 4, 512, 16, 3
        use_prior_model, cem_record = False, 'full'

    a_true, b_true = workloads.pendulum_linear_model()
    solver = se.B200CemSafeMPC(ssm, mpc._constraints, Env, Conf, {'lin_model': (a_true, b_true)}, None, None,
                               w.beta_safety, safe_policy=lambda s: np.zeros(1), lqr=torch.as_tensor(w.k_fb),
                               verbose=False)
    assert not solver._dynamics.problem.a.any() and not solver._dynamics.problem.b.any()
    res2 = solver._mpc.solve(G(w.p0 + 0.01).unsqueeze(0), seed=21)
    assert torch.equal(res2['elite_ids'], res['elite_ids']) and torch.equal(res2['trajectories'], res['trajectories'])


@pytest.mark.parametrize('kind,theta', [('abs_state1', 0.12), ('neg_last_state', 0.0)])
def test_environment_objective_selectors_match_oracle(kind, theta):
    """safempc_cem.py:304-312 with the environment's own objective: |theta* - p[:, 1]| (InvertedPendulum,
    environments.py:505-510) and -p[:, -1] (LunarLander, lunarlander.py:111-113), summed over the horizon."""
    w = workloads.pendulum_workload(n_train=300, num_rollouts=512, time_horizon=5, num_elites=8, num_iterations=4)
    ssm, mpc = workloads.build_solver(w, record='full', seed=31)
    mpc.set_objective({'abs_state1': _lib.SE_OBJ_ABS_STATE1, 'neg_last_state': _lib.SE_OBJ_NEG_LAST_STATE}[kind], theta)
    res = mpc.solve(G(w.p0).unsqueeze(0), seed=31)
    ref = oracle.cem_solve(oracle_problem(w, objective=kind, objective_theta=theta), T(w.p0), T(philox_noise(31, w)),
                           w.num_elites)
    assert_solve_equals_oracle(res, ref)
    feasible = [(r.constraint_costs == 0).sum().item() for r in ref['rollouts']]
    assert max(feasible) >= w.num_elites  # the objective decided at least one selection


@pytest.mark.parametrize('n_s', [3, 4, 6])
def test_remainder_bound_with_nonsymmetric_q_in_higher_dimensions(n_s):
    """test_utils.py:35-59 generalised: the reference takes r^2 = max Re lambda(q B) from a general `torch.eig`
    (utils.py:175-185, 651-665), so a NON-symmetric shape matrix handed to onestep_reachability must go through a general
    eigenvalue routine for n_s >= 3 as well (csrc/eig_small.cuh), not through the symmetrised Jacobi form."""
    rng = np.random.default_rng(60 + n_s)
    n, N = 140, 96
    x = rng.uniform(-1, 1, (n, n_s + 1))
    y = 0.1 * np.sin(x @ rng.standard_normal((n_s + 1, n_s)))
    ls, os_, nz = np.full((n_s, n_s + 1), 0.9), np.full(n_s, 0.05), np.full(n_s, 5e-5)
    a_ = rng.standard_normal((N, n_s, n_s))
    q = 1e-2 * (a_ @ a_.transpose(0, 2, 1) / n_s + 0.1 * np.eye(n_s))
    q = q + 0.05 * np.abs(q).max(axis=(1, 2), keepdims=True) * rng.standard_normal((N, n_s, n_s))  # clearly non-symmetric
    k_fb = rng.uniform(-1, 0, (1, n_s))
    bmat = np.eye(n_s) + k_fb.T @ k_fb
    real = np.array([np.abs(np.linalg.eigvals(qi @ bmat).imag).max() == 0 for qi in q])
    assert real.sum() >= N // 4  # set-up check only: enough cases with a real spectrum remain (47 of 96 at n_s = 6)
    q = q[real]
    N = q.shape[0]
    p, u = 0.1 * rng.standard_normal((N, n_s)), rng.uniform(-1, 1, (N, 1))
    l = np.full(n_s, 0.05)
    ssm = se.B200GpCemSSM(None, n_s, 1, ls, os_, nz, device=DEV)
    ssm.update_model(G(x), G(y), replace_old=True)
    p1, q1, sig = se.onestep_reachability(G(p), ssm, G(u), G(l), G(l), G(q), k_fb=G(k_fb), c_safety=2.0)
    gp = oracle.OracleGpCemSSM(n_s, 1, ls, os_, nz)
    gp.update_model(T(x), T(y), replace_old=True)
    rp, rq, rs = oracle.onestep_reachability(T(p), gp, T(u), T(l), T(l), T(q), T(k_fb), 2.0)
    assert rel_err(p1.cpu(), rp, 1e-3) < 1e-9 and rel_err(sig.cpu(), rs) < 1e-8
    assert rel_err(q1.cpu(), rq, float(rq.abs().max()) * 1e-3) < 1e-8
    # and it matters: the symmetrised matrix has a different spectrum
    sym = np.array([np.linalg.eigvalsh(0.5 * (qi + qi.T) @ bmat @ np.eye(n_s)).max() if False else
                    np.linalg.eigvals(0.5 * (qi + qi.T) @ bmat).real.max() for qi in q])
    gen = np.array([np.linalg.eigvals(qi @ bmat).real.max() for qi in q])
    assert np.abs(sym / gen - 1).max() > 1e-4
