"""Solver wiring: the CEM branch of the reference's `create_solver` (safe_exploration/utils_config.py:35-49, 58-135) with
the B200 classes plugged in, so that an experiment config selects the accelerated path by name.

    conf.solver_type == 'safempc_cem'  and  conf.cem_ssm in ('exact_gp', 'exact_gp_b200')

Everything else the reference's `create_solver` can build (CasADi solvers, GPy models, MC-dropout SSMs, junk dimensions)
is outside the hot path and raises ValueError here, like an unknown value does in the reference.
"""
from typing import Callable, Tuple

import numpy as np

from .safempc import B200CemSafeMPC, construct_constraints, dlqr
from .ssm import B200GpCemSSM, CemSSM

CEM_SSM_NAMES = ('exact_gp', 'exact_gp_b200')


def create_cem_ssm(conf, env, **ssm_kwargs) -> CemSSM:
    """utils_config.py:35-49 (`_create_cem_ssm`).  `ssm_kwargs` go to B200GpCemSSM (device, dtype, hyper-parameters)."""
    if conf.cem_ssm not in CEM_SSM_NAMES:
        raise ValueError(f'Unknown value for cem_ssm, {conf.cem_ssm} (the B200 path implements {CEM_SSM_NAMES})')
    if getattr(conf, 'junk_state_dimen', 0) > 0 or getattr(conf, 'junk_action_dimen', 0) > 0:
        raise ValueError('junk dimensions (JunkDimensionsSSM, ssm_cem.py:134-187) are not part of the B200 path')
    return B200GpCemSSM(conf, state_dimen=env.n_s, action_dimen=env.n_u, **ssm_kwargs)


def get_prior_model(conf, env, prior_env=None) -> Tuple[np.ndarray, np.ndarray]:
    """utils_config.py:70-73, 162-178: linearisation of the prior environment.  The reference builds that environment
    itself from conf.prior_model; environments are outside this package, so the caller passes it (default: the true
    environment, which is what identical prior options produce).  Without conf.lin_prior the reference hands
    lin_model = None to CemSafeMPC, which cannot unpack it (safempc_cem.py:183-186): a ValueError here."""
    if not getattr(conf, 'lin_prior', False):
        raise ValueError('safempc_cem needs conf.lin_prior: CemSafeMPC derives its LQR feedback from the prior linear model')
    return (prior_env if prior_env is not None else env).linearize_discretize()


def create_solver(conf, env, prior_env=None, **ssm_kwargs) -> Tuple[B200CemSafeMPC, Callable[[np.ndarray], np.ndarray]]:
    """utils_config.py:58-135 for solver_type 'safempc_cem'.  Returns (solver, safe_policy) like the reference.

    `env` needs: n_s, n_u, dt, l_mu, l_sigm, u_min_norm, u_max_norm, get_safety_constraints(normalize=True),
    linearize_discretize(), objective_cost_function."""
    if conf.solver_type != 'safempc_cem':
        raise ValueError(f'Unknown solver type: {conf.solver_type} (the B200 path implements safempc_cem)')
    h_mat_safe, h_safe, h_mat_obs, h_obs = env.get_safety_constraints(normalize=True)
    a_true, b_true = env.linearize_discretize()
    lin_model = get_prior_model(conf, env, prior_env)
    wx_cost, wu_cost = conf.lqr_wx_cost, conf.lqr_wu_cost
    k_lqr, _, _ = dlqr(a_true, b_true, wx_cost, wu_cost)
    k_fb = -k_lqr
    safe_policy = lambda x: np.dot(k_fb, x)
    ctrl_bounds = np.hstack((np.reshape(env.u_min_norm, (-1, 1)), np.reshape(env.u_max_norm, (-1, 1))))
    opt_env = dict(h_mat_safe=h_mat_safe, h_safe=h_safe, lin_model=lin_model, ctrl_bounds=ctrl_bounds,
                   safe_policy=safe_policy, h_mat_obs=h_mat_obs, h_obs=h_obs, dt=env.dt,
                   lin_trafo_gp_input=getattr(conf, 'lin_trafo_gp_input', None))
    ssm = create_cem_ssm(conf, env, **ssm_kwargs)
    constraints = construct_constraints(conf, env)
    solver = B200CemSafeMPC(ssm, constraints=constraints, env=env, conf=conf, opt_env=opt_env,
                            wx_feedback_cost=wx_cost, wu_feedback_cost=wu_cost, beta_safety=conf.cem_beta_safety,
                            safe_policy=safe_policy, verbose=getattr(conf, 'verbose', 0) > 0)
    return solver, safe_policy
