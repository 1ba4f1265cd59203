"""GPU parity tests of the tcgen05 (kind::i8) variance GEMM -- SE_GEMM_TC, what SE_GEMM_AUTO resolves to for float64
models of 256 .. 16384 training points.

The contraction || L^-1 k* ||^2 is evaluated from balanced radix-256 digit planes with exact int32 accumulation, so
the only error is the truncation after S planes.  Tolerances written here (relative to the variance itself, which on
these workloads is ~10^3 below the prior scale): S = 7, 8 -> 1e-9 (the float64 bar of test_gpu_parity.py); S = 6 and the
automatic choice -> 1e-8; S = 5 -> 1e-5.  The tile primitive itself is integer work and must be bit-exact.
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
TOL = {0: 1e-8, 5: 1e-5, 6: 1e-8, 7: 1e-9, 8: 1e-9}


def rel_err(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = np.maximum(np.abs(b), 0 if scale is None else scale)
    den = np.where(den == 0, 1.0, den)
    return float(np.max(np.abs(a - b) / den)) if a.size else 0.0


def image(planes):
    """[nk][S][rows][32] int8 digit planes -> bytes in shared-memory operand order [row/8][k/16][row%8][k%16]."""
    nk, S, R, K = planes.shape
    return np.ascontiguousarray(planes.reshape(nk, S, R // 8, 8, 2, 16).transpose(0, 1, 2, 4, 3, 5)).reshape(-1)


@pytest.mark.parametrize('S', [3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize('nk', [1, 2, 5])
def test_tile_primitive_is_bit_exact(S, nk):
    """One 128 x 64 tile: level sums D_l = sum_{t+u=l} A_t B_u^T over nk K steps, full signed-byte range."""
    lib = _lib.load()
    rng = np.random.default_rng(100 * S + nk)
    a = rng.integers(-128, 128, size=(nk, S, 128, 32), dtype=np.int8)
    b = rng.integers(-128, 128, size=(nk, S, 64, 32), dtype=np.int8)
    want = np.zeros((S, 128, 64), dtype=np.int64)
    for t in range(S):
        for u in range(S - t):
            for k in range(nk):
                want[t + u] += a[k, t].astype(np.int64) @ b[k, u].astype(np.int64).T
    da, db = torch.from_numpy(image(a)).to(DEV), torch.from_numpy(image(b)).to(DEV)
    out = torch.full((S, 128, 64), -1, dtype=torch.int32, device=DEV)
    _lib.check(lib.se_debug_tc_probe(_lib.ptr(da), _lib.ptr(db), nk, S, 128, 256, _lib.ptr(out), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().astype(np.int64), want)


def _random_model(n_s, n_u, n, seed, ratio=1e3):
    rng = np.random.default_rng(seed)
    d = n_s + n_u
    x = rng.uniform(-1, 1, (n, d))
    y = 0.1 * np.sin(x @ rng.standard_normal((d, n_s)))
    ls = rng.uniform(0.6, 1.4, (n_s, d))
    os_ = rng.uniform(0.02, 0.08, n_s)
    nz = os_ / ratio * rng.uniform(0.8, 1.2, n_s)
    return x, y, ls, os_, nz


@pytest.mark.parametrize('digits', [0, 5, 6, 7, 8])
@pytest.mark.parametrize('n_s,n_u,n,N', [(2, 1, 300, 1000), (4, 1, 700, 513), (6, 2, 257, 129), (1, 3, 1500, 77),
                                         (3, 2, 129, 128)])
def test_tc_predict_matches_oracle(digits, n_s, n_u, n, N):
    """Ragged n and N, every state dimension; mean and Jacobian are float64 FMA work and keep the 1e-9 bar."""
    x, y, ls, os_, nz = _random_model(n_s, n_u, n, n + N)
    z = np.random.default_rng(1).uniform(-1.1, 1.1, (N, n_s + n_u))
    z[: N // 2] = x[np.random.default_rng(2).integers(0, n, N // 2)] + 1e-3  # next to training points: var ~ noise
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC, tc_digits=digits)
    gp.update_model(G(x), G(y), replace_old=True)
    ref = oracle.OracleGpCemSSM(n_s, n_u, ls, os_, nz)
    ref.update_model(T(x), T(y), replace_old=True)
    mean, var, jac = gp.predict_with_jacobians(G(z[:, :n_s]), G(z[:, n_s:]))
    m0, v0, j0 = ref.predict_with_jacobians(T(z[:, :n_s]), T(z[:, n_s:]))
    impl, used = gp.gemm_info()
    assert impl == 'tcgen05-i8' and (used == digits if digits else used in (6, 7, 8))
    assert rel_err(mean.cpu(), m0, np.abs(m0.numpy()).max()) < 1e-9
    assert rel_err(jac.cpu(), j0, np.abs(j0.numpy()).max()) < 1e-9
    assert rel_err(var.cpu(), v0) < TOL[digits], (digits, used)
    m1, v1 = gp.predict_without_jacobians(G(z[:, :n_s]), G(z[:, n_s:]))
    assert torch.equal(v1, var)
    # the two kernel instantiations contract their FMAs differently: agreement at rounding level only
    assert torch.allclose(m1, mean, rtol=1e-10, atol=1e-12 * float(mean.abs().max()))


def test_auto_digits_follow_the_prior_to_noise_ratio():
    """outputscale / noise = 10^3 -> 6 planes, 10^5 -> 7, 10^8 -> 8; the 1e-8 target holds for each."""
    for ratio, want in ((1e3, 6), (1e5, 7), (1e8, 8)):
        x, y, ls, os_, nz = _random_model(2, 1, 400, 5, ratio)
        nz = os_ / ratio
        gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
        gp.update_model(G(x), G(y), replace_old=True)
        z = np.random.default_rng(3).uniform(-1, 1, (256, 3))
        _, var = gp.predict_without_jacobians(G(z[:, :2]), G(z[:, 2:]))
        assert gp.gemm_info() == ('tcgen05-i8', want)
        gp.set_gemm_impl(_lib.SE_GEMM_MMA)
        _, v64 = gp.predict_without_jacobians(G(z[:, :2]), G(z[:, 2:]))
        assert gp.gemm_info()[0] == 'dmma'
        # against the float64 kernel on the same factors; at 10^8 the float64 kernel's own rounding, amplified by the
        # cancellation, is ~1e-7, so only the looser bound can be asked there
        assert rel_err(var.cpu(), v64.cpu()) < (1e-8 if ratio < 1e6 else 1e-5)


def test_auto_impl_selection():
    """n <= 128 stays on the exact DMMA kernel, larger float64 models use the tensor cores, float32 models FFMA."""
    x, y, ls, os_, nz = _random_model(2, 1, 100, 1)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    gp.update_model(G(x), G(y), replace_old=True)
    assert gp.gemm_info()[0] == 'dmma'
    x, y, ls, os_, nz = _random_model(2, 1, 200, 1)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    gp.update_model(G(x), G(y), replace_old=True)
    assert gp.gemm_info()[0] == 'tcgen05-i8'
    gp32 = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, dtype=torch.float32)
    gp32.update_model(G(x), G(y), replace_old=True)
    assert gp32.gemm_info()[0] == 'simt'
    with pytest.raises(_lib.SecemError):
        gp32.set_gemm_impl(_lib.SE_GEMM_TC)


def test_tc_chunked_workspace_and_reruns_are_bit_identical():
    """Integer accumulation and a fixed partial-sum order: chunking the batch or repeating the call changes nothing."""
    x, y, ls, os_, nz = _random_model(2, 1, 500, 9)
    z = G(np.random.default_rng(4).uniform(-1, 1, (1000, 3)))
    mk = lambda ws: se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC, workspace_bytes=ws)
    big, small = mk(None), mk(2 * 256 * 512 * 8)  # 256 rows per chunk
    for gp in (big, small):
        gp.update_model(G(x), G(y), replace_old=True)
    a = big.predict_with_jacobians(z[:, :2].contiguous(), z[:, 2:].contiguous())
    b = small.predict_with_jacobians(z[:, :2].contiguous(), z[:, 2:].contiguous())
    c = big.predict_with_jacobians(z[:, :2].contiguous(), z[:, 2:].contiguous())
    for s, t, r in zip(a, b, c):
        assert torch.equal(s, t) and torch.equal(s, r)


_CLUSTER_SCRIPT = """
import sys, numpy as np, torch
sys.path.insert(0, {root!r})
sys.path.insert(0, {tests!r})
import safe_exploration_b200 as se
from safe_exploration_b200 import _lib
from test_gpu_tc import _random_model, G, DEV
x, y, ls, os_, nz = _random_model({n_s}, 1, {n}, 9)
z = G(np.random.default_rng(4).uniform(-1, 1, ({N}, {n_s} + 1)))
gp = se.B200GpCemSSM(None, {n_s}, 1, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC)
gp.update_model(G(x), G(y), replace_old=True)
mean, var, jac = gp.predict_with_jacobians(z[:, :{n_s}].contiguous(), z[:, {n_s}:].contiguous())
np.save({out!r}, var.cpu().numpy())
"""


@pytest.mark.parametrize('n_s,n,N', [(2, 500, 1024), (4, 333, 2048)])
def test_cluster_gemm_equals_the_one_cta_gemm_bit_for_bit(n_s, n, N, tmp_path):
    """The default variance GEMM runs 2-CTA clusters that share the digit planes of L^-1 through a multicast bulk copy
    (SE_OZ_CLUSTER=12); the one-CTA kernel (0) and the pairs that share the candidates' planes (21: the shallower row
    tile is padded with the exact zeros above the diagonal) form the same integer level sums and reduce them in the same
    order, so the variances must agree to the last bit.  The engine is chosen per process: the other two run in
    subprocesses."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    got = {}
    for shape in ('12', '0', '21'):
        out = str(tmp_path / f'var_{shape}.npy')
        script = _CLUSTER_SCRIPT.format(root=root, tests=os.path.join(root, 'tests'), n_s=n_s, n=n, N=N, out=out)
        env = dict(os.environ, SE_OZ_CLUSTER=shape)
        proc = subprocess.run([sys.executable, '-c', script], env=env, capture_output=True, text=True, timeout=300)
        assert proc.returncode == 0, proc.stderr[-2000:]
        got[shape] = np.load(out)
    assert np.isfinite(got['12']).all() and (got['12'] > 0).all()
    assert np.array_equal(got['12'], got['0']) and np.array_equal(got['12'], got['21'])


def test_tc_model_update_rebuilds_the_digit_image():
    """update_model appends data (ssm_cem.py:72-99): the digit image of the new factor must be used afterwards."""
    x, y, ls, os_, nz = _random_model(2, 1, 600, 11)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC)
    gp.update_model(G(x[:300]), G(y[:300]), replace_old=True)
    z = np.random.default_rng(5).uniform(-1, 1, (200, 3))
    gp.predict_without_jacobians(G(z[:, :2]), G(z[:, 2:]))
    gp.update_model(G(x[300:]), G(y[300:]))
    _, var = gp.predict_without_jacobians(G(z[:, :2]), G(z[:, 2:]))
    ref = oracle.OracleGpCemSSM(2, 1, ls, os_, nz)
    ref.update_model(T(x), T(y), replace_old=True)
    _, v0 = ref.predict_without_jacobians(T(z[:, :2]), T(z[:, 2:]))
    assert rel_err(var.cpu(), v0) < 1e-8


def test_incremental_model_update_equals_full_refactorisation():
    """appending samples (update_model(replace_old=False), ssm_cem.py:72-99) only computes the new rows
    of L^-1; predictions must equal those of a model factorised from scratch, and the path taken is reported."""
    x, y, ls, os_, nz = _random_model(2, 1, 700, 21)
    z = np.random.default_rng(6).uniform(-1, 1, (300, 3))
    inc = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV)
    inc.update_model(G(x[:400]), G(y[:400]), replace_old=True)
    assert inc.last_update == 'full'
    for lo, hi in ((400, 401), (401, 530), (530, 700)):
        inc.update_model(G(x[lo:hi]), G(y[lo:hi]))
        assert inc.last_update == 'incremental' and inc.x_train.shape[0] == hi
    full = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, incremental=False)
    full.update_model(G(x), G(y), replace_old=True)
    assert full.last_update == 'full'
    a = inc.predict_with_jacobians(G(z[:, :2]), G(z[:, 2:]))
    b = full.predict_with_jacobians(G(z[:, :2]), G(z[:, 2:]))
    for s_, t_ in zip(a, b):
        assert rel_err(s_.cpu(), t_.cpu(), 1e-3 * float(t_.abs().max())) < 1e-8
    # changed hyper-parameters or replaced data force a full factorisation
    inc.noise = inc.noise * 2
    inc.update_model(G(x[:5]), G(y[:5]))
    assert inc.last_update == 'full'
    inc.update_model(G(x[:300]), G(y[:300]), replace_old=True)
    assert inc.last_update == 'full'


@pytest.mark.parametrize('pipeline', ['0', '1'])
def test_tc_cem_solve_matches_oracle(pipeline, monkeypatch):
    """Full safe CEM-MPC solve on the tensor-core path (pendulum, n = 400): same elites, same constraint decisions,
    costs to 1e-8 -- with and without the two-stream population pipeline."""
    monkeypatch.setenv('SE_PIPELINE', pipeline)
    w = workloads.pendulum_workload(n_train=400, num_rollouts=1024, time_horizon=6, num_elites=32, num_iterations=3,
                                    name='tc')
    ssm, mpc = workloads.build_solver(w, record='full', seed=11)
    assert ssm.gemm_info()[0] == 'tcgen05-i8'
    res = mpc.solve(G(w.p0).unsqueeze(0), seed=11)
    noise = np.stack([oracle.philox_normal_noise(11, i, 0, w.num_rollouts, w.time_horizon, 1)
                      for i in range(w.num_iterations)])
    gp = oracle.OracleGpCemSSM(2, 1, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    prob = oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                 w.act_b)
    ref = oracle.cem_solve(prob, T(w.p0), T(noise), w.num_elites)
    assert np.array_equal(res['elite_ids'][:, 0].cpu().numpy(), ref['elite_ids'].numpy())
    for i, r in enumerate(ref['rollouts']):
        assert np.array_equal(res['constraint'][i, 0].cpu().numpy(), r.constraint_costs.numpy())
        assert rel_err(res['objective'][i, 0].cpu().numpy(), r.objective_costs.numpy()) < 1e-8
        assert rel_err(res['trajectories'][i, 0].cpu().numpy(), r.trajectories.numpy(), 1e-6) < 1e-7
    assert int(res['best_id'][0]) == ref['best_id']


@pytest.mark.parametrize('hook', ['se_debug_exp', 'se_debug_exp_table'])
def test_branch_free_exp_is_within_two_ulp_of_libm(hook):
    """csrc/common.cuh exp_tab (the table-based exp of the kernel-row kernel) and exp_nb (the polynomial one of the
    model-refresh gradient kernel): <= 2 ulp against numpy's exp over the range the kernel rows use, exact powers at the
    reduction boundaries, and the clamp below -700 stays tiny and positive."""
    lib = _lib.load()
    debug_exp = getattr(lib, hook)
    rng = np.random.default_rng(7)
    x = np.concatenate([rng.uniform(-60.0, 2.0, 400000), rng.uniform(-700.0, 700.0, 100000),
                        np.array([0.0, -0.0, 1e-300, -1e-300, np.log(2.0) / 2, -np.log(2.0) / 2, -699.999, 699.999]),
                        np.arange(-1000, 1001) * np.log(2.0) / 2, np.arange(-40000, 40001, 7) * np.log(2.0) / 128])
    x = x[(x > -700.0) & (x < 700.0)]
    dx = torch.from_numpy(x).to(DEV)
    out = torch.empty_like(dx)
    _lib.check(debug_exp(_lib.ptr(dx), x.size, _lib.ptr(out), _lib.stream_ptr(DEV)))
    got, want = out.cpu().numpy(), np.exp(x)
    ulp = np.abs(got - want) / np.spacing(want)
    assert ulp.max() <= 2.0, (ulp.max(), x[ulp.argmax()])
    tiny = torch.tensor([-700.0, -750.0, -1e5, -1e300, -np.inf], dtype=torch.float64, device=DEV)
    out = torch.empty_like(tiny)
    _lib.check(debug_exp(_lib.ptr(tiny), tiny.numel(), _lib.ptr(out), _lib.stream_ptr(DEV)))
    o = out.cpu().numpy()
    assert np.all(o > 0) and np.all(o < 1e-300)


def test_non_finite_inputs_poison_the_mean():
    """A NaN / Inf state must surface as NaN in the posterior mean (the ellipsoid step's NaN check looks there,
    gp_reachability_pytorch.py:234-243); the other rows of the batch are unaffected."""
    x, y, ls, os_, nz = _random_model(2, 1, 300, 21)
    gp = se.B200GpCemSSM(None, 2, 1, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC)
    gp.update_model(G(x), G(y), replace_old=True)
    z = np.random.default_rng(5).uniform(-1, 1, (64, 3))
    clean = gp.predict_with_jacobians(G(z[:, :2]), G(z[:, 2:]))
    z2 = z.copy()
    z2[3, 0], z2[17, 2], z2[40, 1] = np.nan, np.inf, -np.inf
    mean, var, jac = gp.predict_with_jacobians(G(z2[:, :2]), G(z2[:, 2:]))
    bad = np.zeros(64, dtype=bool)
    bad[[3, 17, 40]] = True
    assert torch.isnan(mean[bad]).all() and torch.isnan(jac[bad]).all()
    assert torch.equal(mean[~bad], clean[0][~bad]) and torch.equal(var[~bad], clean[1][~bad])
    assert torch.equal(jac[~bad], clean[2][~bad])


@pytest.mark.parametrize('digits,tol', [(4, 1e-6), (3, 1e-4)])
@pytest.mark.parametrize('n_s,n_u,n,N', [(2, 1, 300, 1000), (4, 1, 700, 513), (6, 2, 257, 129)])
def test_float32_tolerance_planes(digits, tol, n_s, n_u, n, N):
    """4 / 3 digit planes: BASELINE's float32 tolerance (1e-4 relative, the variance taken relative to the prior scale
    outputscale + noise like in the float32 tests of test_gpu_parity.py) on the tensor cores; 4 planes hold 1e-6.
    Mean and Jacobian are float64 work and keep 1e-9."""
    x, y, ls, os_, nz = _random_model(n_s, n_u, n, n + N)
    z = np.random.default_rng(1).uniform(-1.1, 1.1, (N, n_s + n_u))
    z[: N // 2] = x[np.random.default_rng(2).integers(0, n, N // 2)] + 1e-3
    gp = se.B200GpCemSSM(None, n_s, n_u, ls, os_, nz, device=DEV, gemm_impl=_lib.SE_GEMM_TC, tc_digits=digits)
    gp.update_model(G(x), G(y), replace_old=True)
    ref = oracle.OracleGpCemSSM(n_s, n_u, ls, os_, nz)
    ref.update_model(T(x), T(y), replace_old=True)
    mean, var, jac = gp.predict_with_jacobians(G(z[:, :n_s]), G(z[:, n_s:]))
    m0, v0, j0 = ref.predict_with_jacobians(T(z[:, :n_s]), T(z[:, n_s:]))
    assert gp.gemm_info() == ('tcgen05-i8', digits)
    prior = (os_ + nz)[None, :]
    assert rel_err(var.cpu(), v0, prior) < tol
    assert rel_err(mean.cpu(), m0, np.abs(m0.numpy()).max()) < 1e-9
    assert rel_err(jac.cpu(), j0, np.abs(j0.numpy()).max()) < 1e-9


def test_float32_tolerance_planes_same_decisions_outside_margin():
    """A whole solve with 4 digit planes on RNG-fixed candidates: identical safe/unsafe decisions for every candidate
    whose terminal safety distance is outside a 1e-4 margin band, ellipsoid centres / shapes within 1e-4 relative."""
    w = workloads.pendulum_workload(n_train=400, num_rollouts=1024, time_horizon=6, num_elites=32, num_iterations=1,
                                    name='tc4')
    ssm, mpc = workloads.build_solver(w, record='full', seed=11)
    ssm.set_tc_digits(4)
    res = mpc.solve(G(w.p0).unsqueeze(0), seed=11)
    assert ssm.gemm_info() == ('tcgen05-i8', 4)
    noise = np.stack([oracle.philox_normal_noise(11, 0, 0, w.num_rollouts, w.time_horizon, 1)])
    gp = oracle.OracleGpCemSSM(2, 1, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    prob = oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                 w.act_b)
    ref = oracle.cem_solve(prob, T(w.p0), T(noise), w.num_elites)
    r = ref['rollouts'][0]
    last = r.trajectories[:, -1]
    d = oracle.lin_ellipsoid_safety_distance(last[:, :2], last[:, 2:].reshape(-1, 2, 2), T(w.h_mat),
                                             T(w.h_vec).reshape(-1, 1))
    clear = (d.abs().min(dim=1)[0] > 1e-4).numpy()
    assert clear.sum() >= 0.9 * len(clear)
    assert np.array_equal(res['constraint'][0, 0].cpu().numpy()[clear], r.constraint_costs.numpy()[clear])
    traj = r.trajectories.numpy()
    assert rel_err(res['trajectories'][0, 0].cpu().numpy(), traj, np.abs(traj).max(axis=(0, 1))[None, None, :]) < 1e-4
