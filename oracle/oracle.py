"""TEST INFRASTRUCTURE ONLY -- float64 CPU restatement of the reference's safe CEM-MPC hot path.

This file is the *oracle*: the checker the CUDA path is compared against.  It is never imported by the
product (`safe_exploration_b200/`); only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline
legs may use it.  Every function cites the reference file:line it restates (paths relative to the
reference root, commit 32f0582a; submodule constrained-cem-mpc @ 4e207630).

Pinning status
--------------
* Reachability / ellipsoid algebra / polytope check / CEM selection, refit and return value: PINNED.  The
  functions below are checked against (a) the reference's own known-answer tests and (b) outputs of the
  verbatim reference modules executed in the build container (`oracle/ref_shim.py`), committed as golden
  vectors under `tests/golden/` (generator: `tests/golden/make_golden.py`).
* RBF-ARD exact-GP posterior (mean / variance / mean-Jacobian): PARITY UNPINNED at the gpytorch boundary.
  The arithmetic lives in gpytorch==0.3.2 (setup.py:28), which is not vendored under /root/reference and is
  not installable here, and no reference test fixes a numeric RBF posterior value.  `OracleGpCemSSM` is the
  textbook exact-GP algebra (Cholesky, fp64) with *materialised* hyper-parameters, following the call sites
  safe_exploration/ssm_cem/gp_ssm_cem.py:45-57 (ScaleKernel(RBFKernel-ARD), batch of n_s independent GPs),
  :79-94 (GaussianLikelihood applied => noise added to the predictive variance; outputs transposed to
  [N, n_s]) and safe_exploration/ssm_pytorch/utilities.py:54-85 (Jacobian of the *mean* w.r.t. z).

dtype: float64 throughout, like the reference (experiments/run.py:98).
"""
from __future__ import annotations

from collections import namedtuple
from typing import List, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

F64 = torch.float64

# constrained_cem_mpc.py:127
Rollouts = namedtuple("Rollouts", "trajectories actions objective_costs constraint_costs")


# ----------------------------------------------------------------------------------------------------------
# GP  (a23-a27 in SURVEY.md 8a)
# ----------------------------------------------------------------------------------------------------------
def rbf_ard(xa: Tensor, xb: Tensor, lengthscale: Tensor, outputscale: Tensor) -> Tensor:
    """k_o(x, x') = s_o exp(-1/2 sum_d ((x_d - x'_d) / l_od)^2) for every output o.

    gp_ssm_cem.py:45-57: ScaleKernel(RBFKernel(batch_size=n_s, ard_num_dims=d), batch_size=n_s).
    :param xa: [A, d], xb: [B, d], lengthscale: [n_s, d], outputscale: [n_s]
    :returns: [n_s, A, B]
    """
    sa = xa.unsqueeze(0) / lengthscale.unsqueeze(1)  # [n_s, A, d]
    sb = xb.unsqueeze(0) / lengthscale.unsqueeze(1)  # [n_s, B, d]
    diff = sa.unsqueeze(2) - sb.unsqueeze(1)  # [n_s, A, B, d]
    return outputscale.view(-1, 1, 1) * torch.exp(-0.5 * (diff * diff).sum(-1))


class OracleGpCemSSM:
    """Restatement of `GpCemSSM` (rbf kernel) -- gp_ssm_cem.py:17-133 -- with explicit hyper-parameters.

    Duck-types the `CemSSM` interface (ssm_cem.py:11-131): num_states, num_actions, x_train, y_train,
    predict_with_jacobians, predict_without_jacobians, predict_raw, update_model.
    """

    parametric = False

    def __init__(self, state_dimen: int, action_dimen: int, lengthscale, outputscale, noise):
        self.num_states = state_dimen
        self.num_actions = action_dimen
        d = state_dimen + action_dimen
        self.lengthscale = torch.as_tensor(lengthscale, dtype=F64).reshape(state_dimen, d)
        self.outputscale = torch.as_tensor(outputscale, dtype=F64).reshape(state_dimen)
        self.noise = torch.as_tensor(noise, dtype=F64).reshape(state_dimen)
        self._x_train = None
        self._y_train = None
        self.alpha = None  # [n_s, n]
        self.chol = None  # [n_s, n, n] lower, K + noise I = L L^T
        self.linv = None  # [n_s, n, n] lower, L^{-1}

    @property
    def x_train(self):
        return self._x_train

    @property
    def y_train(self):
        return self._y_train

    def update_model(self, train_x: Tensor, train_y: Tensor, opt_hyp=False, replace_old=False) -> None:
        """ssm_cem.py:72-99 (replace or concatenate) + gp_ssm_cem.py:96-101 (targets stored as [n_s, n])."""
        train_x = torch.as_tensor(train_x, dtype=F64)
        train_y = torch.as_tensor(train_y, dtype=F64)
        n = train_x.size(0)
        if train_x.shape != (n, self.num_states + self.num_actions):
            raise ValueError(f"Wanted shape {(n, self.num_states + self.num_actions)}, got {train_x.shape}")
        if train_y.shape != (n, self.num_states):
            raise ValueError(f"Wanted shape {(n, self.num_states)}, got {train_y.shape}")
        if replace_old or self._x_train is None or self._y_train is None:
            x_new, y_new = train_x, train_y
        else:
            x_new = torch.cat((self._x_train, train_x), dim=0)
            y_new = torch.cat((self._y_train, train_y), dim=0)
        self._x_train, self._y_train = x_new, y_new
        k = rbf_ard(x_new, x_new, self.lengthscale, self.outputscale)
        k = k + torch.diag_embed(self.noise.view(-1, 1).expand(-1, x_new.size(0)))
        self.chol = torch.linalg.cholesky(k)
        self.alpha = torch.cholesky_solve(y_new.t().unsqueeze(-1), self.chol).squeeze(-1)
        eye = torch.eye(x_new.size(0), dtype=F64).expand(self.num_states, -1, -1)
        self.linv = torch.linalg.solve_triangular(self.chol, eye, upper=False)

    def _posterior(self, z: Tensor, want_jac: bool):
        z = torch.as_tensor(z, dtype=F64)
        kstar = rbf_ard(z, self._x_train, self.lengthscale, self.outputscale)  # [n_s, N, n]
        mean = torch.einsum("onk,ok->on", kstar, self.alpha)
        v = torch.einsum("oij,onj->oni", self.linv, kstar)  # L^{-1} k*
        # gp_ssm_cem.py:93: likelihood(model(z)) => predictive variance includes the noise (quirk Q7)
        var = self.outputscale.view(-1, 1) - (v * v).sum(-1) + self.noise.view(-1, 1)
        jac = None
        if want_jac:
            # d mean_o / d z_d = -sum_i alpha_oi k_o(z, x_i) (z_d - x_id) / l_od^2   (utilities.py:54-85)
            diff = z.unsqueeze(1) - self._x_train.unsqueeze(0)  # [N, n, d]
            w = kstar * self.alpha.unsqueeze(1)  # [n_s, N, n]
            jac = -torch.einsum("onk,nkd->nod", w, diff) / (self.lengthscale ** 2).unsqueeze(0)
        return mean, var, jac

    def predict_raw(self, z: Tensor) -> Tuple[Tensor, Tensor]:
        """gp_ssm_cem.py:90-94 -- returns [n_s, N] mean and variance."""
        mean, var, _ = self._posterior(z, False)
        return mean, var

    def predict_without_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        """gp_ssm_cem.py:75-88 -- [N, n_s] mean and variance."""
        mean, var, _ = self._posterior(torch.cat((states, actions), dim=1), False)
        return mean.t().contiguous(), var.t().contiguous()

    def predict_with_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """gp_ssm_cem.py:59-73 -- adds d mean / d z as [N, n_s, n_s + n_u]."""
        mean, var, jac = self._posterior(torch.cat((states, actions), dim=1), True)
        return mean.t().contiguous(), var.t().contiguous(), jac

    def collect_metrics(self):
        return {}


def exact_mll(x: Tensor, y: Tensor, lengthscale: Tensor, outputscale: Tensor, noise: Tensor) -> Tensor:
    """[n_s] exact marginal log likelihoods  -1/2 y^T (K + noise I)^-1 y - 1/2 log|K + noise I| - n/2 log 2 pi  of the n_s
    independent GPs (what gpytorch's ExactMarginalLogLikelihood evaluates before it divides by n; call site
    gp_ssm_cem.py:110, 120).  PARITY UNPINNED like the posterior: the arithmetic is gpytorch's, restated from the textbook.
    :param x: [n, d], y: [n, n_s]"""
    n = x.size(0)
    k = rbf_ard(x, x, lengthscale, outputscale) + torch.diag_embed(noise.view(-1, 1).expand(-1, n))
    chol = torch.linalg.cholesky(k)
    yt = y.t()
    alpha = torch.cholesky_solve(yt.unsqueeze(-1), chol).squeeze(-1)
    return -0.5 * (yt * alpha).sum(-1) - chol.diagonal(dim1=1, dim2=2).log().sum(-1) - 0.5 * n * np.log(2 * np.pi)


def exact_mll_and_grad(x, y, lengthscale, outputscale, noise) -> Tuple[Tensor, Tensor]:
    """mll [n_s] and d mll_o / d (lengthscale_o[0..d), outputscale_o, noise_o) as [n_s, d + 2], by autograd through the
    Cholesky factorisation -- the route loss.backward() takes in gp_ssm_cem.py:121."""
    x, y = torch.as_tensor(x, dtype=F64), torch.as_tensor(y, dtype=F64)
    ls = torch.as_tensor(lengthscale, dtype=F64).clone().requires_grad_(True)
    os_ = torch.as_tensor(outputscale, dtype=F64).clone().requires_grad_(True)
    nz = torch.as_tensor(noise, dtype=F64).clone().requires_grad_(True)
    mll = exact_mll(x, y, ls, os_, nz)
    mll.sum().backward()  # outputs are independent: the gradient of the sum is the per-output gradient
    return mll.detach(), torch.cat((ls.grad, os_.grad.unsqueeze(1), nz.grad.unsqueeze(1)), dim=1)


def train_hyperparameters(x, y, lengthscale, outputscale, noise, iterations: int, lr: float = 0.01):
    """GpCemSSM._train_model (gp_ssm_cem.py:103-129): Adam(lr=0.01) for `iterations` steps on
    loss = -sum_o mll_o / n, raw parameters through softplus (gpytorch's default positive constraint in 0.3.x).
    :returns: (lengthscale [n_s, d], outputscale [n_s], noise [n_s], losses list)"""
    x, y = torch.as_tensor(x, dtype=F64), torch.as_tensor(y, dtype=F64)
    inv_softplus = lambda v: torch.log(torch.expm1(torch.as_tensor(v, dtype=F64)))
    raw = [inv_softplus(t).clone().requires_grad_(True) for t in (lengthscale, outputscale, noise)]
    opt = torch.optim.Adam(raw, lr=lr)
    losses = []
    for _ in range(iterations):
        opt.zero_grad()
        ls, os_, nz = (torch.nn.functional.softplus(r) for r in raw)
        loss = -(exact_mll(x, y, ls, os_, nz) / x.size(0)).sum()
        loss.backward()
        losses.append(loss.item())
        opt.step()
    with torch.no_grad():
        ls, os_, nz = (torch.nn.functional.softplus(r).clone() for r in raw)
    return ls, os_, nz, losses


# ----------------------------------------------------------------------------------------------------------
# Ellipsoid algebra (a14-a20)
# ----------------------------------------------------------------------------------------------------------
def trace_batch(xs: Tensor, loops: bool = False) -> Tensor:
    """utils.py:668-679.  `loops=True` keeps the reference's per-item Python loop (timing fidelity)."""
    if loops:
        out = torch.empty((xs.shape[0],), dtype=xs.dtype)
        for i in range(xs.shape[0]):
            out[i] = torch.trace(xs[i])
        return out
    return xs.diagonal(dim1=1, dim2=2).sum(-1)


def max_real_eigenvalue_batch(xs: Tensor, loops: bool = False) -> Tensor:
    """utils.py:651-665 + :181-185: general (non-symmetric) eigenvalues, max of the real parts.

    The reference asserts that all imaginary parts vanish (utils.py:182).
    """
    if loops:
        out = torch.empty((xs.shape[0],), dtype=xs.dtype)
        for i in range(xs.shape[0]):
            ev = torch.linalg.eigvals(xs[i])
            assert not (ev.imag != 0).any(), "All imaginary parts should be 0"
            out[i] = ev.real.max()
        return out
    ev = torch.linalg.eigvals(xs)
    assert not (ev.imag != 0).any(), "All imaginary parts should be 0"
    return ev.real.max(dim=1)[0]


def compute_remainder_overapproximations(q: Tensor, k_fb: Tensor, l_mu: Tensor, l_sigma: Tensor,
                                         loops: bool = False) -> Tuple[Tensor, Tensor]:
    """utils.py:152-194: S = [I, k_fb^T], B = S S^T, r^2 = lambda_max(q B), u_mu = l_mu r^2, u_sigma = l_sigma r.

    :param q: [N, n_s, n_s]; k_fb: [n_u, n_s] (shared) or [N, n_u, n_s]; l_mu, l_sigma: [n_s] or [N, n_s]
    """
    n_s = q.size(1)
    if k_fb.dim() == 2:
        k_fb = k_fb.unsqueeze(0)
    s = torch.cat((torch.eye(n_s, dtype=q.dtype).expand(k_fb.size(0), -1, -1), k_fb.transpose(1, 2)), dim=2)
    b = torch.matmul(s, s.transpose(1, 2))
    r_sqr = max_real_eigenvalue_batch(torch.matmul(q, b), loops)
    u_mu = l_mu * r_sqr.unsqueeze(1)
    u_sigma = l_sigma * torch.sqrt(r_sqr).unsqueeze(1)
    return u_mu, u_sigma


def ellipsoid_from_rectangle(u_b: Tensor) -> Tensor:
    """utils_ellipsoid.py:282-309: diag(n u_b^2); asserts u_b > 0."""
    assert u_b.dim() == 2
    assert torch.all(u_b > 0), f"All elements of u_b must be >0: {u_b}"
    return torch.diag_embed(u_b.shape[1] * u_b ** 2)


def sum_two_ellipsoids(p_1: Tensor, q_1: Tensor, p_2: Tensor, q_2: Tensor, loops: bool = False):
    """utils_ellipsoid.py:102-140: c = sqrt(tr q1 / tr q2); q = (1 + 1/c) q1 + (1 + c) q2 (no guard, quirk Q8)."""
    c = torch.sqrt(trace_batch(q_1, loops) / trace_batch(q_2, loops)).view(-1, 1, 1)
    return p_1 + p_2, (1 + (1. / c)) * q_1 + (1 + c) * q_2


def fix_zeros_nans(x: Tensor) -> Tuple[Tensor, bool, bool]:
    """gp_reachability_pytorch.py:234-243.  Returns (x, ok, found_zero): NaN => not ok; any exact zero =>
    every x <= 0 is set to 1e-5 (note: negatives survive when no exact zero is present)."""
    if torch.isnan(x).any():
        return torch.empty_like(x), False, False
    if (x == 0).any():
        x = x.clone()
        x[x <= 0] = 1e-5
        return x, True, True
    return x, True, False


def onestep_reachability(p_center: Tensor, ssm, k_ff: Tensor, l_mu: Tensor, l_sigma: Tensor,
                         q_shape: Optional[Tensor] = None, k_fb: Optional[Tensor] = None, c_safety: float = 1.,
                         a: Optional[Tensor] = None, b: Optional[Tensor] = None, loops: bool = False):
    """gp_reachability_pytorch.py:18-181 (a13): returns (p_1 [N,n_s], q_1 [N,n_s,n_s], sigma [N,n_s])."""
    n_s, n_u = ssm.num_states, ssm.num_actions
    if a is None:  # :60-62
        a = torch.eye(n_s, dtype=F64)
        b = torch.zeros((n_s, n_u), dtype=F64)

    if q_shape is None:  # point branch :65-99
        mu_0, sigm_0 = ssm.predict_without_jacobians(p_center, k_ff)
        sigm_0, ok, _ = fix_zeros_nans(sigm_0)
        if not ok:
            raise ValueError("nan in sigm_0")
        rkhs_bounds, ok, _ = fix_zeros_nans(c_safety * torch.sqrt(sigm_0))
        if not ok:
            raise ValueError("nan/zero in rkhs_bounds")
        q_1 = ellipsoid_from_rectangle(rkhs_bounds)
        p_1 = p_center @ a.t() + k_ff @ b.t() + mu_0
        return p_1, q_1, sigm_0

    mu_0, sigm_0, jac_mu = ssm.predict_with_jacobians(p_center, k_ff)  # :114
    sigm_0, ok, _ = fix_zeros_nans(sigm_0)
    if not ok:
        raise ValueError("nan in sigm_0")
    a_mu = jac_mu[:, :, :n_s]
    b_mu = jac_mu[:, :, n_s:]
    h = a + a_mu + torch.matmul(b_mu + b, k_fb)  # :131
    p_0 = mu_0 + p_center @ a.t() + k_ff @ b.t()  # :132
    q_0 = torch.bmm(h, torch.bmm(q_shape, h.transpose(1, 2)))  # :134
    ub_mean, ub_sigma = compute_remainder_overapproximations(q_shape, k_fb, l_mu, l_sigma, loops)  # :142-145
    b_sigma_eps, ok, _ = fix_zeros_nans(c_safety * (torch.sqrt(sigm_0) + ub_sigma))  # :146-153
    if not ok:
        raise ValueError("nan in b_sigma_eps")
    q_lagrange_sigm = ellipsoid_from_rectangle(b_sigma_eps)  # :155
    q_lagrange_mu = ellipsoid_from_rectangle(ub_mean)  # :162
    zeros = torch.zeros_like(p_0)
    p_sum, q_sum = sum_two_ellipsoids(zeros, q_lagrange_sigm, zeros, q_lagrange_mu, loops)  # :169
    p_1, q_1 = sum_two_ellipsoids(p_sum, q_sum, p_0, q_0, loops)  # :172
    return p_1, q_1, sigm_0


def multistep_reachability(p_0: Tensor, ssm, k_fb: Tensor, k_ff: Tensor, l_mu: Tensor, l_sigma: Tensor,
                           q_0: Optional[Tensor] = None, c_safety: float = 1., a=None, b=None,
                           k_fb_init: Optional[Tensor] = None):
    """Batched restatement of the NumPy `multistep_reachability` (gp_reachability.py:168-221, a28).

    :param p_0: [N, n_s]; q_0: [N, n_s, n_s] or None; k_ff: [N, T, n_u];
                k_fb: [T-1, n_u, n_s] per-step feedback gains (or [n_u, n_s], used for every step)
    :returns: p_T, q_T, p_all [N, T, n_s], q_all [N, T, n_s, n_s], sigma_all [N, T, n_s]
    """
    t_len = k_ff.size(1)
    if k_fb.dim() == 2:
        k_fb = k_fb.unsqueeze(0).expand(max(t_len - 1, 1), -1, -1)
    p_all, q_all, s_all = [], [], []
    p, q, sig = onestep_reachability(p_0, ssm, k_ff[:, 0], l_mu, l_sigma, q_0, k_fb_init, c_safety, a, b)
    p_all.append(p), q_all.append(q), s_all.append(sig)
    for i in range(1, t_len):
        p, q, sig = onestep_reachability(p, ssm, k_ff[:, i], l_mu, l_sigma, q, k_fb[i - 1], c_safety, a, b)
        p_all.append(p), q_all.append(q), s_all.append(sig)
    return p, q, torch.stack(p_all, 1), torch.stack(q_all, 1), torch.stack(s_all, 1)


# ----------------------------------------------------------------------------------------------------------
# Polytope checks (a21, a22, a7, a8)
# ----------------------------------------------------------------------------------------------------------
def lin_ellipsoid_safety_distance(p_center: Tensor, q_shape: Tensor, h_mat: Tensor, h_vec: Tensor,
                                  c_safety: float = 1.0) -> Tensor:
    """gp_reachability_pytorch.py:184-215: d = h_mat p + c sqrt(diag(h_mat q h_mat^T)) - h_vec  -> [N, m]."""
    d_center = p_center @ h_mat.t()
    d_shape = c_safety * torch.sqrt(torch.einsum("mi,nij,mj->nm", h_mat, q_shape, h_mat))
    return d_center + d_shape - h_vec.reshape(1, -1)


def is_ellipsoid_inside_polytope(p_center, q_shape, h_mat, h_vec) -> Tensor:
    """gp_reachability_pytorch.py:218-231: inside iff no face has d >= 0 (touching is unsafe, quirk Q2)."""
    d = lin_ellipsoid_safety_distance(p_center, q_shape, h_mat, h_vec)
    return (d >= 0).sum(dim=1) == 0


def action_constraint_costs(actions: Tensor, a_mat: Tensor, b_vec: Tensor, penalty: float = 3.) -> Tensor:
    """constrained_cem_mpc.py:100-110 + :29-41: penalty * #{t : some face has A u_t - b > 0} per rollout.

    :param actions: [N, T, n_u]; a_mat: [m_u, n_u]; b_vec: [m_u]
    """
    viol = (torch.einsum("mj,ntj->ntm", a_mat, actions) - b_vec.view(1, 1, -1)) > 0
    return penalty * viol.any(dim=2).sum(dim=1).to(F64)


def terminal_constraint_costs(p_last: Tensor, q_last: Tensor, h_mat: Tensor, h_vec: Tensor) -> Tensor:
    """safempc_cem.py:102-132: 10 * (1 - inside) for the LAST ellipsoid only (quirk Q1: the "state" constraint
    is terminal-only as well)."""
    inside = is_ellipsoid_inside_polytope(p_last, q_last, h_mat, h_vec)
    return 10. * (1 - inside.to(F64))


# ----------------------------------------------------------------------------------------------------------
# CEM (a1-a6, a9-a12)
# ----------------------------------------------------------------------------------------------------------
def flatten_pq(p: Tensor, q: Optional[Tensor]) -> Tensor:
    """safempc_cem.py:40-54."""
    n = p.size(0)
    if q is None:
        q = torch.zeros((n, p.size(1), p.size(1)), dtype=p.dtype)
    return torch.cat((p.reshape(n, -1), q.reshape(n, -1)), dim=1)


def unflatten_pq(flat: Tensor, n_s: int):
    """safempc_cem.py:56-73: an all-zero q for the WHOLE batch means "point states" (q = None)."""
    n = flat.size(0)
    p = flat[:, :n_s]
    q = flat[:, n_s:].reshape(n, n_s, n_s)
    if len(q.nonzero()) == 0:
        q = None
    return p, q


class SafeCemProblem:
    """Everything `CemSafeMPC._dynamics_func` + `construct_constraints` close over
    (safempc_cem.py:135-146, 162-200, 288-312), as plain tensors."""

    def __init__(self, ssm, l_mu, l_sigma, k_fb, a, b, beta_safety, h_mat, h_vec, act_a, act_b,
                 objective: str = "neg_variance", objective_theta: float = 0.0, use_prior_model: bool = True,
                 h_mat_obs=None, h_vec_obs=None):
        t = lambda x: torch.as_tensor(x, dtype=F64)
        self.ssm = ssm
        self.n_s, self.n_u = ssm.num_states, ssm.num_actions
        self.l_mu, self.l_sigma = t(l_mu), t(l_sigma)
        self.k_fb = t(k_fb).reshape(self.n_u, self.n_s)
        self.a = t(a).reshape(self.n_s, self.n_s)
        self.b = t(b).reshape(self.n_s, self.n_u)
        if not use_prior_model:  # safempc_cem.py:291-296: zeros, not identity (quirk Q9)
            self.a = torch.zeros_like(self.a)
            self.b = torch.zeros_like(self.b)
        self.beta = float(beta_safety)
        self.h_mat, self.h_vec = t(h_mat), t(h_vec).reshape(-1)
        self.act_a, self.act_b = t(act_a), t(act_b).reshape(-1)
        self.objective = objective
        self.objective_theta = float(objective_theta)
        # Opt-in obstacle polytope for the intermediate ellipsoids.  The reference's CEM path never applies it
        # (safempc_cem.py:137 reads h_mat_obs, h_obs and drops them); the constraint itself is the one its CasADi
        # solver poses (safempc_simple.py:373-383), charged 10 per violating state like safempc_cem.py:132.
        self.h_mat_obs = None if h_mat_obs is None else t(h_mat_obs)
        self.h_vec_obs = None if h_vec_obs is None else t(h_vec_obs).reshape(-1)

    def objective_cost(self, p_next: Tensor, sigma: Tensor) -> Tensor:
        """safempc_cem.py:304-312; environments.py:505-510 (pendulum objective |theta* - p[:,1]|);
        lunarlander.py:111-113 (-p[:,-1])."""
        if self.objective == "neg_variance":
            return -torch.sum(sigma, dim=1)
        if self.objective == "abs_state1":
            return torch.abs(self.objective_theta - p_next[:, 1])
        if self.objective == "neg_last_state":
            return -p_next[:, -1]
        raise ValueError(self.objective)


def perform_rollouts(prob: SafeCemProblem, p0: Tensor, actions: Tensor, loops: bool = False) -> Rollouts:
    """constrained_cem_mpc.py:145-192 with the dynamics of safempc_cem.py:288-302 for GIVEN actions.

    :param p0: [n_s] initial (point) state; actions: [N, T, n_u]
    """
    n, t_len, _ = actions.shape
    n_s = prob.n_s
    p = p0.reshape(1, n_s).repeat(n, 1)
    q = None
    traj = torch.empty((n, t_len + 1, n_s + n_s * n_s), dtype=F64)
    traj[:, 0] = flatten_pq(p, None)
    objective = torch.zeros((n,), dtype=F64)
    obstacle = torch.zeros((n,), dtype=F64)
    for t in range(t_len):
        p, q, sigma = onestep_reachability(p, prob.ssm, actions[:, t], prob.l_mu, prob.l_sigma, q, prob.k_fb,
                                           prob.beta, prob.a, prob.b, loops)
        traj[:, t + 1] = flatten_pq(p, q)
        objective = objective + prob.objective_cost(p, sigma)
        if prob.h_mat_obs is not None and t < t_len - 1:
            obstacle = obstacle + terminal_constraint_costs(p, q, prob.h_mat_obs, prob.h_vec_obs)
        # unflatten's "all-zero q => None" rule (safempc_cem.py:70-71) cannot trigger here: q_1 > 0 on its diagonal.
    if loops:
        # constrained_cem_mpc.py:187-192: per-rollout Python loop over the constraint objects
        cc = torch.empty((n,), dtype=F64)
        for i in range(n):
            cc[i] = (action_constraint_costs(actions[i:i + 1], prob.act_a, prob.act_b)[0]
                     + terminal_constraint_costs(p[i:i + 1], q[i:i + 1], prob.h_mat, prob.h_vec)[0])
    else:
        cc = (action_constraint_costs(actions, prob.act_a, prob.act_b)
              + terminal_constraint_costs(p, q, prob.h_mat, prob.h_vec))
    return Rollouts(traj, actions, objective, cc + obstacle)


def select_elites(objective_costs: Tensor, constraint_costs: Tensor, num_elites: int) -> Tensor:
    """constrained_cem_mpc.py:241-257, with a STABLE index tie-break (quirk Q4: the reference's sorts are
    unstable and constraint costs are small integers, so ties are the norm).  Returns elite ids [E] (int64)."""
    feasible = (constraint_costs == 0).nonzero().reshape(-1)
    if feasible.numel() >= num_elites:
        order = torch.sort(objective_costs[feasible], stable=True)[1]
        return feasible[order[:num_elites]]
    order = torch.sort(constraint_costs, stable=True)[1]
    return order[:num_elites]


def refit(actions: Tensor, elite_ids: Tensor) -> Tuple[Tensor, Tensor]:
    """constrained_cem_mpc.py:234-235: mean and Bessel-corrected std over the elites (quirk Q5: no floor)."""
    el = actions[elite_ids]
    return el.mean(dim=0), el.std(dim=0)


def pick_first_feasible(constraint_costs: Tensor) -> int:
    """constrained_cem_mpc.py:272-280 (quirk Q3): the expression sorts along a size-1 dimension, so the result is
    the FIRST feasible rollout of the last iteration, not the arg-min objective.  -1 if none."""
    feasible = (constraint_costs == 0).nonzero().reshape(-1)
    return int(feasible[0]) if feasible.numel() > 0 else -1


def pick_best_feasible(objective_costs: Tensor, constraint_costs: Tensor) -> int:
    """'Intended semantics' switch for Q3: arg-min objective among feasible, index tie-break.  -1 if none."""
    feasible = (constraint_costs == 0).nonzero().reshape(-1)
    if feasible.numel() == 0:
        return -1
    return int(feasible[torch.sort(objective_costs[feasible], stable=True)[1][0]])


def cem_solve(prob: SafeCemProblem, p0: Tensor, noise: Tensor, num_elites: int, loops: bool = False,
              pick: str = "first"):
    """constrained_cem_mpc.py:222-239 + :259-280 for a FIXED standard-normal noise tensor.

    actions = means + stds * noise[i] reproduces Normal(means, stds).sample((N,)) (:167) for RNG-fixed
    candidates.  :param noise: [I, N, T, n_u].
    :returns: dict(actions [T,n_u] | None, best_id, rollouts: List[Rollouts], means [I,T,n_u], stds [I,T,n_u],
              elite_ids [I,E])
    """
    num_iter, n, t_len, n_u = noise.shape
    means = torch.zeros((t_len, n_u), dtype=F64)
    stds = torch.ones((t_len, n_u), dtype=F64)
    rollouts_by_time: List[Rollouts] = []
    means_hist, stds_hist, elite_hist = [], [], []
    for i in range(num_iter):
        actions = means.unsqueeze(0) + stds.unsqueeze(0) * noise[i]
        rollouts = perform_rollouts(prob, p0, actions, loops)
        elite_ids = select_elites(rollouts.objective_costs, rollouts.constraint_costs, num_elites)
        means, stds = refit(actions, elite_ids)
        rollouts_by_time.append(rollouts)
        means_hist.append(means), stds_hist.append(stds), elite_hist.append(elite_ids)
    last = rollouts_by_time[-1]
    best = (pick_first_feasible(last.constraint_costs) if pick == "first"
            else pick_best_feasible(last.objective_costs, last.constraint_costs))
    return dict(actions=None if best < 0 else last.actions[best], best_id=best, rollouts=rollouts_by_time,
                means=torch.stack(means_hist), stds=torch.stack(stds_hist), elite_ids=torch.stack(elite_hist))


# ----------------------------------------------------------------------------------------------------------
# Counter-based RNG restated from the CUDA side (Philox4x32-10 + Box-Muller), integer part bit-exact
# ----------------------------------------------------------------------------------------------------------
_PHILOX_M0 = np.uint64(0xD2511F53)
_PHILOX_M1 = np.uint64(0xCD9E8D57)
_PHILOX_W0 = 0x9E3779B9
_PHILOX_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter: np.ndarray, key: np.ndarray) -> np.ndarray:
    """Philox4x32-10 (Salmon et al., SC'11).  counter: [..., 4] uint32, key: [2] uint32 -> [..., 4] uint32."""
    c = [counter[..., i].astype(np.uint64) for i in range(4)]
    k0, k1 = int(key[0]), int(key[1])
    for _ in range(10):
        p0 = _PHILOX_M0 * c[0]
        p1 = _PHILOX_M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        c = [(hi1 ^ c[1] ^ np.uint64(k0)) & _MASK32, lo1, (hi0 ^ c[3] ^ np.uint64(k1)) & _MASK32, lo0]
        k0 = (k0 + _PHILOX_W0) & 0xFFFFFFFF
        k1 = (k1 + _PHILOX_W1) & 0xFFFFFFFF
    return np.stack(c, axis=-1).astype(np.uint32)


def philox_normal_noise(seed: int, iteration: int, first_candidate: int, num_candidates: int, t_len: int,
                        n_u: int) -> np.ndarray:
    """Standard-normal noise exactly as the CUDA sampler draws it (csrc/cem_kernels.cuh, `se_cem_noise`).

    For global candidate c, iteration i and pair index m (elements 2m and 2m+1 of the flattened [T*n_u] action
    sequence): counter = (c, m, i, 0x5eCE3), key = (seed & 0xffffffff, seed >> 32);
    u1 = ((r0 << 21 | r1 >> 11) + 0.5) 2^-53, u2 likewise from (r2, r3);
    z0 = sqrt(-2 ln u1) cos(2 pi u2), z1 = sqrt(-2 ln u1) sin(2 pi u2).
    :returns: [num_candidates, T, n_u] float64
    """
    tot = t_len * n_u
    pairs = (tot + 1) // 2
    cand = np.arange(first_candidate, first_candidate + num_candidates, dtype=np.uint32)
    ctr = np.zeros((num_candidates, pairs, 4), dtype=np.uint32)
    ctr[..., 0] = cand[:, None]
    ctr[..., 1] = np.arange(pairs, dtype=np.uint32)[None, :]
    ctr[..., 2] = np.uint32(iteration)
    ctr[..., 3] = np.uint32(0x5ECE3)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint32)
    r = philox4x32_10(ctr, key).astype(np.uint64)
    m1 = (r[..., 0] << np.uint64(21)) | (r[..., 1] >> np.uint64(11))
    m2 = (r[..., 2] << np.uint64(21)) | (r[..., 3] >> np.uint64(11))
    u1 = (m1.astype(np.float64) + 0.5) * 2.0 ** -53
    u2 = (m2.astype(np.float64) + 0.5) * 2.0 ** -53
    rad = np.sqrt(-2.0 * np.log(u1))
    z = np.stack((rad * np.cos(2.0 * np.pi * u2), rad * np.sin(2.0 * np.pi * u2)), axis=-1)
    return z.reshape(num_candidates, 2 * pairs)[:, :tot].reshape(num_candidates, t_len, n_u)
