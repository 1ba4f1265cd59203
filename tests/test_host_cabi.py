"""CPU: the C-ABI library loads and exports every symbol include/secem.h declares (no compute calls without a GPU),
and the host-side mirror of the reference's interfaces behaves like the reference's own unit tests say."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import safe_exploration_b200 as se
from safe_exploration_b200 import _lib, workloads
from safe_exploration_b200.safempc import fold_constraints

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'secem.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(se_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f'{name} declared in include/secem.h but not exported by libsecem.so'
        assert name in _lib.SIGNATURES, f'{name} has no ctypes signature'
    assert lib.se_version() == 100


def test_struct_layouts_match_header(tmp_path):
    """ctypes mirrors against the C compiler's view of include/secem.h: sizes and the offset of every field."""
    import subprocess
    mirrors = {'se_problem': _lib.SeProblem, 'se_cem_config': _lib.SeCemConfig, 'se_cem_outputs': _lib.SeCemOutputs}
    lines = []
    for cname, cls in mirrors.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for field, _ in cls._fields_:
            lines.append(f'printf("{cname}.{field} %zu\\n", offsetof({cname}, {field}));')
    src = tmp_path / 'layout.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "secem.h"\nint main(void) {\n' +
                   '\n'.join(lines) + '\nreturn 0; }\n')
    exe = tmp_path / 'layout'
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(['gcc', '-I', os.path.join(root, 'include'), str(src), '-o', str(exe)], check=True)
    got = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True,
                                                       text=True).stdout.splitlines())
    for cname, cls in mirrors.items():
        assert int(got[cname]) == ctypes.sizeof(cls), cname
        for field, _ in cls._fields_:
            assert int(got[f'{cname}.{field}']) == getattr(cls, field).offset, f'{cname}.{field}'
    assert ctypes.sizeof(_lib.SeCemConfig) == 36
    assert ctypes.sizeof(_lib.SeProblem) == 168  # 160 before clamp_nonpositive, 144 before the obstacle polytope


def test_argument_errors_are_reported_without_a_gpu():
    lib = _lib.load()
    out = ctypes.c_void_p(None)
    rc = lib.se_gp_create(ctypes.byref(out), 9, 1, 10, _lib.SE_F64, None, None, None, None, None, None, None)
    assert rc != 0 and b'n_s' in lib.se_last_error()
    with pytest.raises(_lib.SecemError):
        _lib.check(rc)


def test_no_cpu_fallback():
    ssm = se.B200GpCemSSM(None, 2, 1)
    with pytest.raises(_lib.SecemError):
        ssm.predict_without_jacobians(torch.zeros(1, 2), torch.zeros(1, 1))  # no model / not on a CUDA device
    with pytest.raises(_lib.SecemError):
        se.is_ellipsoid_inside_polytope(torch.zeros(1, 2), torch.eye(2).unsqueeze(0), torch.eye(2), torch.ones(2, 1))


def test_pq_flattener_round_trip():
    """test_safempc_cem.py:23-42."""
    f = se.PQFlattener(3)
    p, q = torch.rand(4, 3), torch.rand(4, 3, 3)
    flat = f.flatten(p, q)
    assert flat.shape == (4, 12) and f.get_flat_state_dimen() == 12
    p2, q2 = f.unflatten(flat)
    assert torch.equal(p, p2) and torch.equal(q, q2)
    p3, q3 = f.unflatten(f.flatten(p, None))
    assert torch.equal(p, p3) and q3 is None


def test_torch_polytope_and_action_constraint():
    """constrained_cem_mpc_test.py:60-136 + test_safempc_cem.py:59-71."""
    poly = se.box2torchpoly([[0., 10.], [0., 10.]])
    assert torch.tensor([5., 5.]) in poly and torch.tensor([10., 0.]) in poly and torch.tensor([11., 5.]) not in poly
    assert poly.contains_points(torch.tensor([[1., 1.], [20., 1.], [5., -1.], [10., 10.]])) == 2
    assert np.allclose(poly.chebXc.numpy(), [5., 5.])
    c = se.ActionConstraint(se.box2torchpoly([[-0.5, 0.5]]))
    assert c(None, torch.tensor([[0.2], [-5.], [6.]])) == 6


def test_construct_constraints_and_folding():
    class Env:
        n_s, n_u = 4, 1
        u_min_norm, u_max_norm = np.array([-0.5]), np.array([0.5])

        def get_safety_constraints(self, normalize=True):
            h, v = workloads.cartpole_polytope()
            return h, v[:, None], None, None

    class Conf:
        use_state_constraint = True

    cs = se.construct_constraints(Conf(), Env())
    assert isinstance(cs[0], se.ActionConstraint) and isinstance(cs[1], se.EllipsoidStateConstraint)
    Conf.use_state_constraint = False
    cs = se.construct_constraints(Conf(), Env())
    assert isinstance(cs[1], se.EllipsoidTerminalConstraint)
    prob = fold_constraints(se.SafeProblem(4, 1, np.ones(4), np.ones(4)), cs)
    st = prob.c_struct()
    assert st.m_safe == 9 and st.m_act == 2 and st.act_penalty == 3.0 and st.state_penalty == 10.0
    assert st.all_states_constraint == 0  # quirk Q1: "state" constraint is terminal-only by default
    with pytest.raises(TypeError):
        fold_constraints(prob, [object()])


def test_workload_constants_match_reference_environments():
    """Values printed by the verbatim reference environments in the build container (make_golden.py)."""
    a, b = workloads.pendulum_linear_model()
    assert np.allclose(a, [[1.02465061, 0.99005578], [0.05041017, 1.02465061]], atol=1e-8)
    assert np.allclose(b, [[1.34427126], [0.03346995]], atol=1e-8)
    h, v = workloads.pendulum_polytope(True)
    assert sorted(np.round(v, 8)) == sorted(np.round([0.34906585, 0.8, 0.8, 0.34906585], 8))
    ac, bc = workloads.cartpole_linear_model()
    assert np.allclose(ac[0], [1.0, 1.66666667e-01, 4.19847224e-03, -5.25453321e-04], atol=1e-8)
    assert np.allclose(bc[:, 0], [0.02653238, 0.31886176, 0.29942205, 0.35589813], atol=1e-8)
    hc, vc = workloads.cartpole_polytope()
    assert hc.shape == (9, 4) and vc.shape == (9,)


def test_golden_cem_constants_equal_workload_restatement(golden_cem):
    g = golden_cem  # a, b, k_fb, polytope come from the verbatim reference environment
    w = workloads.pendulum_workload(n_train=8, num_rollouts=8)
    assert np.allclose(w.a, g['a']) and np.allclose(w.b, g['b']) and np.allclose(w.k_fb, g['k_fb'])
    order = lambda h, v: sorted(map(tuple, np.round(np.hstack((h, v[:, None])), 9).tolist()))
    assert order(w.h_mat, w.h_vec) == order(g['h_mat'], g['h_vec'])
    assert np.allclose(w.act_a, g['act_a']) and np.allclose(w.act_b, g['act_b'])


def test_dlqr_matches_riccati():
    a, b = workloads.pendulum_linear_model()
    k, x, eig = se.dlqr(a, b, np.diag([1., 2.]), 25 * np.eye(1))
    assert np.all(np.abs(eig) < 1)
    resid = a.T @ x @ a - x - a.T @ x @ b @ np.linalg.solve(b.T @ x @ b + 25, b.T @ x @ a) + np.diag([1., 2.])
    assert np.abs(resid).max() < 1e-8


def test_cem_constructor_contract():
    ssm = se.B200GpCemSSM(None, 2, 1)
    dyn = se.B200SafeDynamics(ssm, se.SafeProblem(2, 1, [.1, .1], [.1, .1], k_fb=np.zeros((1, 2)), a=np.eye(2),
                                                  b=np.zeros((2, 1))))
    cons = [se.ActionConstraint(se.box2torchpoly([[-1., 1.]])),
            se.EllipsoidTerminalConstraint(2, *workloads.pendulum_polytope())]
    mpc = se.B200ConstrainedCemMpc(dyn, cons, 6, 1, 5, 64, 8, 3)
    assert mpc._num_iterations == 3 and mpc._rollout_function._constraints[1] is cons[1]  # safempc_cem.py:270, 280
    with pytest.raises(ValueError):
        se.B200ConstrainedCemMpc(dyn, cons, 5, 1, 5, 64, 8, 3)
    # an arbitrary DynamicsFunc selects the interface path (reference-shaped loop on device tensors) ...
    generic = se.B200ConstrainedCemMpc(lambda s, a: (s, a), cons, 6, 1, 5, 64, 8, 3)
    assert not generic.fused and mpc.fused
    assert generic._rollout_function._constraints[1] is cons[1]
    with pytest.raises(_lib.SecemError):  # ... which is not a CPU fallback: host tensors are refused
        generic.get_actions(torch.zeros(6))
    with pytest.raises(TypeError):  # batching / sharding exist on the fused path only
        se.B200ConstrainedCemMpc(lambda s, a: (s, a), cons, 6, 1, 5, 64, 8, 3, num_instances=4)
    with pytest.raises(_lib.SecemError):
        mpc.get_actions(torch.zeros(1, 6))  # no training data -> loud failure, no fallback


def test_update_model_shape_errors():
    ssm = se.B200GpCemSSM(None, 2, 1)
    with pytest.raises(ValueError):  # ssm_cem.py:83-84
        ssm.update_model(torch.zeros(5, 4), torch.zeros(5, 2))
    with pytest.raises(ValueError):
        ssm.update_model(torch.zeros(5, 3), torch.zeros(5, 3))


def test_exp_polynomial_in_the_kernel_source_is_the_derived_one():
    """csrc/common.cuh exp_nb: its Horner constants are exactly what tools/exp_poly.py derives (no hand-edited digits)."""
    import importlib.util
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location('exp_poly', os.path.join(root, 'tools', 'exp_poly.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    coef = mod.coefficients()
    src = open(os.path.join(root, 'safe_exploration_b200', 'csrc', 'common.cuh')).read()
    body = src[src.index('double exp_nb(double x)'):src.index('// The same contract as exp_nb')]
    got = [float(body[body.index('double p = ') + 11:].split(';')[0])]
    got += [float(v) for v in re.findall(r'p = fma\(p, r, ([0-9.e+-]+)\);', body)]
    assert got == coef[::-1]
    # exp_tab: the 64 table entries are 2^(j/64) correctly rounded, the reduction constants are 64 / ln 2 and ln 2 / 64
    from decimal import Decimal, getcontext
    getcontext().prec = 60
    ln2 = Decimal(2).ln()
    tab_src = src[src.index('kExpTab[64] = {') + 15:]
    tab = [float(v) for v in tab_src[:tab_src.index('}')].replace('\n', ' ').split(',')]
    assert tab == [float((ln2 * j / 64).exp()) for j in range(64)]
    body = src[src.index('double exp_tab(double x'):src.index('finite_f64')]
    hi = float(ln2 / 64)
    assert repr(float(Decimal(64) / ln2)) in body and repr(hi) in body and repr(float(ln2 / 64 - Decimal(hi))) in body


class _PendulumEnv:
    """The attributes create_solver reads from an Environment (environments.py), with the pendulum constants that
    workloads.py restates from the reference."""
    n_s, n_u, dt = 2, 1, 0.05
    l_mu, l_sigm = np.array([0.05, 0.02]), np.array([0.05, 0.02])
    u_min_norm, u_max_norm = np.array([-1.0]), np.array([1.0])
    objective_spec = None

    def get_safety_constraints(self, normalize=True):
        h, v = workloads.pendulum_polytope(True)
        return h, v[:, None], None, None

    def linearize_discretize(self):
        return workloads.pendulum_linear_model()

    @staticmethod
    def objective_cost_function(ps):
        return None


class _CemConf:
    solver_type, cem_ssm = 'safempc_cem', 'exact_gp_b200'
    lin_prior, use_prior_model, use_state_constraint = True, True, True
    lqr_wx_cost, lqr_wu_cost = np.diag([1., 2.]), 25 * np.eye(1)
    cem_beta_safety = 2.0
    mpc_time_horizon, cem_num_rollouts, cem_num_elites, cem_num_iterations = 3, 64, 8, 2
    exact_gp_training_iterations, exact_gp_kernel = 10, 'rbf'
    junk_state_dimen = junk_action_dimen = 0
    verbose = 0


def test_create_solver_wires_the_b200_classes():
    """utils_config.py:58-135: the safempc_cem branch builds SSM, constraints and solver, and returns the LQR policy."""
    solver, safe_policy = se.create_solver(_CemConf(), _PendulumEnv())
    assert isinstance(solver, se.B200CemSafeMPC) and isinstance(solver._ssm, se.B200GpCemSSM)
    assert solver.state_dimen == 2 and solver.action_dimen == 1 and solver.safety_trajectory_length == 3
    a, b = workloads.pendulum_linear_model()
    k, _, _ = se.dlqr(a, b, np.diag([1., 2.]), 25 * np.eye(1))
    assert np.allclose(safe_policy(np.array([0.1, -0.2])), -k @ np.array([0.1, -0.2]))
    assert np.allclose(solver.lin_model[0], a) and np.allclose(solver.lin_model[1], b)
    cons = solver._mpc._rollout_function._constraints if hasattr(solver._mpc, '_rollout_function') else None
    assert cons is None or isinstance(cons[1], se.EllipsoidStateConstraint)

    class NoPrior(_CemConf):
        lin_prior = False
    with pytest.raises(ValueError, match='lin_prior'):
        se.create_solver(NoPrior(), _PendulumEnv())

    class Dropout(_CemConf):
        cem_ssm = 'mc_dropout'
    with pytest.raises(ValueError, match='Unknown value for cem_ssm'):
        se.create_solver(Dropout(), _PendulumEnv())

    class Casadi(_CemConf):
        solver_type = 'safempc'
    with pytest.raises(ValueError, match='Unknown solver type'):
        se.create_solver(Casadi(), _PendulumEnv())


def test_runner_host_logic_with_a_scripted_solver():
    """do_rollout / do_rollout_batch (episode_runner.py:165-324): stacks start with the reference's zero row, exit codes
    follow the MpcResult, a finished scenario keeps its slot, metrics carry the reference's names."""
    from safe_exploration_b200.runner import do_rollout, do_rollout_batch
    from safe_exploration_b200.safempc import MpcResult

    class Env:
        n_s, n_u = 2, 1

        def __init__(self, x0, fail_at=None):
            self.x0, self.fail_at, self.t = np.asarray(x0, float), fail_at, 0

        def reset(self, mean=None, std=None):
            self.x, self.t = self.x0.copy(), 0
            return self.x

        def step(self, action):
            self.x = self.x + float(np.asarray(action).reshape(-1)[0])
            self.t += 1
            return np.asarray(action).reshape(1), self.x.copy(), 2 * self.x, self.fail_at == self.t, 7

    class Solver:
        def __init__(self):
            self.batch_shapes = []

        def get_action(self, state):
            return np.array([0.5]), MpcResult.FOUND_SOLUTION if state[0] < 1 else MpcResult.SAFE_CONTROLLER

        def get_action_batch(self, states):
            self.batch_shapes.append(states.shape)
            out = [self.get_action(s) for s in states]
            return np.stack([a for a, _ in out]), [r for _, r in out]

        def collect_metrics(self):
            return {'losses': []}

    class Sink:
        def __init__(self):
            self.s, self.n = {}, []

        def log_scalar(self, k, v, step):
            self.s[k] = v

        def log_non_scalars(self, d, step):
            self.n.append(d)

    sink = Sink()
    xx, yy, cc, codes, failed = do_rollout(Env([0., 0.]), 4, metrics=sink, solver=Solver(), cost=lambda s: s[0])
    assert xx.shape == (5, 3) and not xx[0].any() and np.array_equal(xx[1:, 0], [0., .5, 1., 1.5])
    assert np.array_equal(yy[1:, 0], [1., 2., 3., 4.]) and cc == [.5, 1., 1.5, 2.] and not failed
    assert codes[:, 0].tolist() == [0, 1, 1, 0, 0]
    assert sink.s['episode_length'] == 4 and sink.s['mpc_found_solution_count'] == 2
    assert sink.s['safe_controller_fallback_count'] == 2 and sink.s['env_result'] == 7 and sink.n == [{'losses': []}]
    solver = Solver()
    res = do_rollout_batch([Env([0., 0.]), Env([5., 5.], fail_at=2), Env([0., 0.])], 4, solver)
    assert solver.batch_shapes == [(3, 2)] * 4  # the finished scenario keeps its slot
    assert np.array_equal(res[0][0], xx) and np.array_equal(res[2][0], xx)
    assert res[1][0].shape == (3, 3) and res[1][4] and res[1][3][:, 0].tolist() == [0, 0, 0]


def test_general_eigenvalue_routine_matches_lapack():
    """csrc/eig_small.cuh (host build of the function the kernels call for a non-symmetric shape matrix): the largest real
    eigenvalue of q B as the reference takes it from a general `torch.eig` (utils.py:175-185, 651-665); NaN for a complex
    pair, where the reference asserts.  Includes the reference's own non-symmetric case (test_utils.py:35-59)."""
    lib = _lib.load()

    def eig_max(m):
        m = np.ascontiguousarray(m, dtype=np.float64)
        out = ctypes.c_double(0)
        assert lib.se_debug_eig_max(m.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), m.shape[0], ctypes.byref(out)) == 0
        return out.value

    q = np.array([[.8, .2], [.1, .9]])  # hmm... the 2 x 2 of test_utils.py:35-59 (non-symmetric on purpose)
    k_fb = np.array([[.5, .2]])
    assert np.isclose(eig_max(q @ (np.eye(2) + k_fb.T @ k_fb)), np.linalg.eigvals(q @ (np.eye(2) + k_fb.T @ k_fb)).real.max(),
                      rtol=1e-13)
    rng = np.random.default_rng(5)
    checked = 0
    for trial in range(3000):
        n = int(rng.integers(1, 7))
        a = rng.standard_normal((n, n))
        q = a @ a.T / n + 0.1 * np.eye(n)
        q = q + 0.02 * np.abs(q).max() * rng.standard_normal((n, n))  # clearly non-symmetric
        k = rng.uniform(-1, 0, (1, n))
        m = q @ (np.eye(n) + k.T @ k) if trial % 3 else np.eye(n) * 2.0 + 1e-7 * rng.standard_normal((n, n))
        ev = np.linalg.eigvals(m)
        got = eig_max(m)
        if np.abs(ev.imag).max() > 1e-9 * np.abs(ev).max():
            continue  # complex pair: NaN or the real part, like an assertion in the reference
        checked += 1
        assert abs(got - ev.real.max()) <= 1e-9 * np.abs(ev).max(), (trial, n, got, ev)
    assert checked > 2000
    assert np.isnan(eig_max(np.array([[0., 1., 0.], [-1., 0., 0.], [0., 0., -5.]])))  # rotation block: +-i


def test_committed_traffic_capture_names_the_kernel_the_bench_line_reports():
    """bench.py takes `roofline.traffic` from profiles/ncu_traffic.json and refuses a capture of another kernel: the
    committed capture must be of the kernel the default C2 run launches (the 2-CTA cluster flavour of the variance GEMM),
    on the default workload, with more DRAM bytes than the algorithmic 417 MB and less than twice that."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location('bench_module', os.path.join(ROOT, 'bench.py'))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    name = bench._gemm_kernel_name(6, 16384)
    assert name == 'oz::gp_trisumsq_ozaki_cluster_kernel<6,1,2>'
    assert bench._gemm_kernel_name(6, 128) == 'oz::gp_trisumsq_ozaki_kernel<6,2>'  # one candidate tile: no pairs
    traffic, stamp = bench.profiled_traffic(name, 'C2/N16384')
    assert stamp is not None and stamp.get('kernel') == name, stamp
    assert 417e6 < traffic < 2 * 417e6
    with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
        assert all(os.path.exists(os.path.join(ROOT, e['source'])) for e in json.load(f))
