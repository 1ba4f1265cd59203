"""GP ellipsoid reachability on the B200 -- same function names and argument meaning as the reference's
safe_exploration/gp_reachability_pytorch.py (onestep_reachability, lin_ellipsoid_safety_distance,
is_ellipsoid_inside_polytope) plus the batched form of the NumPy multistep_reachability
(safe_exploration/gp_reachability.py:168-221).  All numerics run in libsecem.so.
"""
import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from . import _lib
from .ssm import B200GpCemSSM, CemSSM, assert_shape


class SafeProblem:
    """Host-side owner of an `se_problem` (include/secem.h): the constants CemSafeMPC._dynamics_func and
    construct_constraints close over (safempc_cem.py:135-146, 162-200, 288-312)."""

    def __init__(self, n_s: int, n_u: int, l_mu, l_sigma, k_fb=None, a=None, b=None, c_safety: float = 1.0,
                 h_mat=None, h_vec=None, act_a=None, act_b=None, act_penalty: float = 3.0,
                 state_penalty: float = 10.0, objective: int = _lib.SE_OBJ_NEG_VARIANCE,
                 objective_theta: float = 0.0, all_states_constraint: bool = False, h_mat_obs=None,
                 h_vec_obs=None, clamp_nonpositive: bool = False):
        def arr(x, shape):
            if x is None:
                return None
            if isinstance(x, Tensor):
                x = x.detach().cpu().numpy()
            x = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
            if -1 in shape:
                if x.size % max(1, -int(np.prod(shape))) != 0:
                    raise ValueError(f'Wanted shape {shape}, got {x.shape}')
            elif x.size != int(np.prod(shape)):
                raise ValueError(f'Wanted shape {shape}, got {x.shape}')
            return np.ascontiguousarray(x.reshape(shape))

        self.n_s, self.n_u = n_s, n_u
        self.k_fb = arr(k_fb, (n_u, n_s))
        self.a = arr(a, (n_s, n_s))
        self.b = arr(b, (n_s, n_u))
        self.l_mu = arr(l_mu, (n_s,))
        self.l_sigma = arr(l_sigma, (n_s,))
        self.c_safety = float(c_safety)
        self.h_mat = arr(h_mat, (-1, n_s))
        self.h_vec = None if h_vec is None else arr(h_vec, (self.h_mat.shape[0],))
        self.act_a = arr(act_a, (-1, n_u))
        self.act_b = None if act_b is None else arr(act_b, (self.act_a.shape[0],))
        self.act_penalty, self.state_penalty = float(act_penalty), float(state_penalty)
        self.objective, self.objective_theta = int(objective), float(objective_theta)
        self.all_states_constraint = bool(all_states_constraint)
        # obstacle polytope of the intermediate ellipsoids (safempc_simple.py:373-383); None = reference CEM behaviour
        self.h_mat_obs = arr(h_mat_obs, (-1, n_s))
        self.h_vec_obs = None if h_vec_obs is None else arr(h_vec_obs, (self.h_mat_obs.shape[0],))
        # _fix_zeros_nans inside se_cem_solve: False = only exact zeros become 1e-5 (a negative variance is reported as
        # NaN, the reference's behaviour for a batch without an exact zero); True = every variance <= 0 becomes 1e-5
        self.clamp_nonpositive = bool(clamp_nonpositive)

    def with_(self, **kw) -> 'SafeProblem':
        cur = dict(n_s=self.n_s, n_u=self.n_u, l_mu=self.l_mu, l_sigma=self.l_sigma, k_fb=self.k_fb, a=self.a,
                   b=self.b, c_safety=self.c_safety, h_mat=self.h_mat, h_vec=self.h_vec, act_a=self.act_a,
                   act_b=self.act_b, act_penalty=self.act_penalty, state_penalty=self.state_penalty,
                   objective=self.objective, objective_theta=self.objective_theta,
                   all_states_constraint=self.all_states_constraint, h_mat_obs=self.h_mat_obs,
                   h_vec_obs=self.h_vec_obs, clamp_nonpositive=self.clamp_nonpositive)
        cur.update(kw)
        return SafeProblem(**cur)

    def c_struct(self) -> _lib.SeProblem:
        dp = lambda x: None if x is None else x.ctypes.data_as(C.POINTER(C.c_double))
        s = _lib.SeProblem()
        s.n_s, s.n_u = self.n_s, self.n_u
        s.k_fb, s.a, s.b = dp(self.k_fb), dp(self.a), dp(self.b)
        s.l_mu, s.l_sigma = dp(self.l_mu), dp(self.l_sigma)
        s.c_safety = self.c_safety
        s.m_safe = 0 if self.h_mat is None else self.h_mat.shape[0]
        s.h_mat, s.h_vec = dp(self.h_mat), dp(self.h_vec)
        s.m_act = 0 if self.act_a is None else self.act_a.shape[0]
        s.act_a, s.act_b = dp(self.act_a), dp(self.act_b)
        s.act_penalty, s.state_penalty = self.act_penalty, self.state_penalty
        s.objective, s.objective_theta = self.objective, self.objective_theta
        s.all_states_constraint = 1 if self.all_states_constraint else 0
        s.m_obs = 0 if self.h_mat_obs is None else self.h_mat_obs.shape[0]
        s.h_mat_obs, s.h_vec_obs = dp(self.h_mat_obs), dp(self.h_vec_obs)
        s.clamp_nonpositive = 1 if self.clamp_nonpositive else 0
        return s  # the numpy arrays it points into are owned by `self`


class ReachabilityNumericsError(ValueError):
    """Raised where the reference raises ValueError('nan in sigm_0 ...') (gp_reachability_pytorch.py:76-81, 117-121,
    149-153)."""


def _save_model(ssm, state: Optional[Tensor], action: Optional[Tensor]) -> None:
    """gp_reachability_pytorch.py:256-266: before raising on a NaN variance the reference dumps the model and the inputs
    that produced it to `negative_variance_state.pt` in the working directory.  Same file, same keys where they exist
    here ('state', 'action', 'x_train', 'y_train'); the GPyTorch state dicts become the materialised hyper-parameters."""
    try:
        cpu = lambda t: None if t is None else t.detach().cpu()
        torch.save({'gp_model': {k: cpu(getattr(ssm, k, None)) for k in ('lengthscale', 'outputscale')},
                    'gp_likelihood': {'noise': cpu(getattr(ssm, 'noise', None))},
                    'state': cpu(state), 'action': cpu(action),
                    'x_train': cpu(getattr(ssm, 'x_train', None)), 'y_train': cpu(getattr(ssm, 'y_train', None))},
                   'negative_variance_state.pt')
    except Exception as exc:  # the dump is a debugging aid: never mask the numeric error with an I/O error
        print(f'WARNING: could not write negative_variance_state.pt: {exc}')


def check_flags(flags: Tensor, where: str, dump=None) -> None:
    """Host side of _fix_zeros_nans (gp_reachability_pytorch.py:234-243): one sync per call instead of 2-3 per step.
    `dump` = (ssm, state, action): what the reference's _save_model writes before the ValueError."""
    n_nan, n_zero = int(flags[0]), int(flags[1])
    if n_nan > 0:
        if dump is not None:
            _save_model(*dump)
        raise ReachabilityNumericsError(f'nan in sigm_0 / rkhs_bounds / b_sigma_eps ({n_nan} entries) in {where}')
    if n_zero > 0:
        print(f'WARNING: found 0 in {where} but carried on ({n_zero} entries set to 1e-5)')


def _gp_of(ssm) -> B200GpCemSSM:
    if not isinstance(ssm, B200GpCemSSM):
        raise TypeError(f'{type(ssm).__name__} is not a B200GpCemSSM: this entry point evaluates the GP inside the '
                        f'kernels.  onestep_reachability accepts any CemSSM (it calls the model\'s predict_* methods '
                        f'and runs the ellipsoid algebra of the step on the device: se_reach_step)')
    return ssm


def _onestep_generic_ssm(p_center, ssm, k_ff, prob: SafeProblem, q_shape):
    """Interface fallback for a CemSSM that is not the B200 GP, e.g. the MC-dropout models of
    ssm_cem/dropout_ssm_cem.py: the model's own predict_with(out)_jacobians runs on the caller's device tensors
    (gp_reachability_pytorch.py:73, 114), the ellipsoid algebra of the step runs in libsecem.so (`se_reach_step`)."""
    _lib.require_cuda(p_center, 'p_center')
    lib = _lib.load()
    N = p_center.shape[0]
    n_s = prob.n_s
    dt, dev = p_center.dtype, p_center.device
    code = _lib.dtype_code(dt)
    if q_shape is None:
        mean, var = ssm.predict_without_jacobians(p_center, k_ff)
        jac = None
    else:
        mean, var, jac = ssm.predict_with_jacobians(p_center, k_ff)
    prep = lambda t: None if t is None else t.detach().to(device=dev, dtype=dt).contiguous()
    p, u, q, mean, var, jac = (prep(t) for t in (p_center, k_ff, q_shape, mean, var, jac))
    assert_shape(mean, (N, n_s))
    assert_shape(var, (N, n_s))
    assert_shape(jac, (N, n_s, n_s + prob.n_u), ignore_if_none=True)
    p1 = torch.empty((N, n_s), dtype=dt, device=dev)
    q1 = torch.empty((N, n_s, n_s), dtype=dt, device=dev)
    sig = torch.empty((N, n_s), dtype=dt, device=dev)
    flags = torch.zeros(4, dtype=torch.int32, device=dev)
    if N > 0:
        cs = prob.c_struct()
        with torch.cuda.device(dev):
            _lib.check(lib.se_reach_step(code, C.byref(cs), _lib.ptr(p), _lib.ptr(q), _lib.ptr(u), _lib.ptr(mean),
                                         _lib.ptr(var), _lib.ptr(jac), N, _lib.ptr(p1), _lib.ptr(q1), _lib.ptr(sig),
                                         _lib.ptr(flags), _lib.stream_ptr(dev)))
        check_flags(flags.cpu(), 'onestep_reachability', dump=(ssm, p_center, k_ff))
    return p1, q1, sig


def onestep_reachability(p_center: Tensor, ssm: CemSSM, k_ff: Tensor, l_mu: Tensor, l_sigma: Tensor,
                         q_shape: Optional[Tensor] = None, k_fb: Optional[Tensor] = None, c_safety: float = 1.,
                         verbose: int = 1, a: Optional[Tensor] = None, b: Optional[Tensor] = None
                         ) -> Tuple[Tensor, Tensor, Tensor]:
    """Drop-in for gp_reachability_pytorch.onestep_reachability (:18-181): same arguments, same returns
    (p_1 [N, n_s], q_1 [N, n_s, n_s], sigma [N, n_s]); `verbose` is accepted and ignored.  `ssm` may be any CemSSM:
    a B200GpCemSSM is evaluated inside the kernels, any other model through its own predict_* methods."""
    gp = ssm
    N = p_center.shape[0]
    n_s, n_u = gp.num_states, gp.num_actions
    assert_shape(p_center, (N, n_s))
    assert_shape(k_ff, (N, n_u))
    assert_shape(l_mu, (n_s,))
    assert_shape(l_sigma, (n_s,))
    assert_shape(q_shape, (N, n_s, n_s), ignore_if_none=True)
    assert_shape(k_fb, (n_u, n_s), ignore_if_none=True)
    assert_shape(a, (n_s, n_s), ignore_if_none=True)
    assert_shape(b, (n_s, n_u), ignore_if_none=True)
    prob = SafeProblem(n_s, n_u, l_mu, l_sigma, k_fb=k_fb, a=a, b=b if a is not None else None, c_safety=c_safety)
    if not isinstance(ssm, B200GpCemSSM):
        return _onestep_generic_ssm(p_center, ssm, k_ff, prob, q_shape)
    dev, dt = p_center.device, gp.dtype
    p = gp._prep(p_center, 'p_center')
    u = gp._prep(k_ff, 'k_ff')
    q = gp._prep(q_shape, 'q_shape') if q_shape is not None else None
    p1 = torch.empty((N, n_s), dtype=dt, device=dev)
    q1 = torch.empty((N, n_s, n_s), dtype=dt, device=dev)
    sig = torch.empty((N, n_s), dtype=dt, device=dev)
    flags = torch.zeros(4, dtype=torch.int32, device=dev)
    if N > 0:
        cs = prob.c_struct()
        with torch.cuda.device(dev):
            _lib.check(gp._lib.se_reach_onestep(gp.handle, C.byref(cs), _lib.ptr(p), _lib.ptr(q), _lib.ptr(u), N,
                                                _lib.ptr(p1), _lib.ptr(q1), _lib.ptr(sig), _lib.ptr(flags),
                                                _lib.stream_ptr(dev)))
        check_flags(flags.cpu(), 'onestep_reachability', dump=(ssm, p_center, k_ff))
    return p1, q1, sig


def multistep_reachability(p_0: Tensor, ssm: CemSSM, k_fb, k_ff: Tensor, l_mu, l_sigma, q_0: Optional[Tensor] = None,
                           c_safety: float = 1., a=None, b=None, k_fb_init=None, return_all: bool = True):
    """Batched multistep_reachability (gp_reachability.py:168-221): T ellipsoid steps for N ellipsoids in one call.

    :param p_0: [N, n_s]; q_0: [N, n_s, n_s] or None; k_ff: [N, T, n_u]
    :param k_fb: [n_u, n_s] (same gain every step, the CEM path) or [T-1, n_u, n_s] per-step gains
    :param k_fb_init: gain of the first step when q_0 is given (defaults to k_fb / k_fb[0])
    :returns: p_T, q_T, p_all [N, T, n_s], q_all [N, T, n_s, n_s], sigma_all [N, T, n_s] (the *_all are None if
              return_all is False)
    """
    gp = _gp_of(ssm)
    n_s, n_u = gp.num_states, gp.num_actions
    N, T = k_ff.shape[0], k_ff.shape[1]
    assert_shape(p_0, (N, n_s))
    assert_shape(k_ff, (N, T, n_u))
    assert_shape(q_0, (N, n_s, n_s), ignore_if_none=True)
    k_fb_np = np.asarray(k_fb.detach().cpu().numpy() if isinstance(k_fb, Tensor) else k_fb, dtype=np.float64)
    steps_buf, steps_ptr = None, None
    if k_fb_np.ndim == 3:
        if k_fb_np.shape != (max(T - 1, 0), n_u, n_s) and not (T == 1 and k_fb_np.shape[0] >= 0):
            raise ValueError(f'Wanted shape {(T - 1, n_u, n_s)}, got {k_fb_np.shape}')
        steps_buf, steps_ptr = _lib.host_doubles(k_fb_np)
        first = k_fb_init if k_fb_init is not None else (k_fb_np[0] if k_fb_np.shape[0] > 0 else np.zeros((n_u, n_s)))
    else:
        first = k_fb_init if k_fb_init is not None else k_fb_np
    prob = SafeProblem(n_s, n_u, l_mu, l_sigma, k_fb=first, a=a, b=b if a is not None else None, c_safety=c_safety)
    dev, dt = p_0.device, gp.dtype
    p0 = gp._prep(p_0, 'p_0')
    u = gp._prep(k_ff, 'k_ff')
    q0 = gp._prep(q_0, 'q_0') if q_0 is not None else None
    pT = torch.empty((N, n_s), dtype=dt, device=dev)
    qT = torch.empty((N, n_s, n_s), dtype=dt, device=dev)
    p_all = torch.empty((N, T, n_s), dtype=dt, device=dev) if return_all else None
    q_all = torch.empty((N, T, n_s, n_s), dtype=dt, device=dev) if return_all else None
    s_all = torch.empty((N, T, n_s), dtype=dt, device=dev) if return_all else None
    flags = torch.zeros(4, dtype=torch.int32, device=dev)
    if N > 0:
        cs = prob.c_struct()
        with torch.cuda.device(dev):
            _lib.check(gp._lib.se_reach_multistep(gp.handle, C.byref(cs), _lib.ptr(p0), _lib.ptr(q0), _lib.ptr(u), N,
                                                  T, steps_ptr, _lib.ptr(p_all), _lib.ptr(q_all), _lib.ptr(s_all),
                                                  _lib.ptr(pT), _lib.ptr(qT), _lib.ptr(flags), _lib.stream_ptr(dev)))
        check_flags(flags.cpu(), 'multistep_reachability', dump=(ssm, p_0, k_ff))
    return pT, qT, p_all, q_all, s_all


def _safety(p_center: Tensor, q_shape: Tensor, h_mat: Tensor, h_vec: Tensor, c_safety: float, want_d: bool,
            want_inside: bool):
    N = p_center.size(0)
    m, n_s = h_mat.shape
    assert_shape(p_center, (N, n_s))
    assert_shape(q_shape, (N, n_s, n_s))
    assert_shape(h_vec, (m, 1))
    _lib.require_cuda(p_center, 'p_center')
    lib = _lib.load()
    dt = p_center.dtype
    code = _lib.dtype_code(dt)
    p = p_center.detach().contiguous()
    q = q_shape.detach().to(dt).contiguous()
    _hm, hm_p = _lib.host_doubles(h_mat.detach().cpu().numpy())
    _hv, hv_p = _lib.host_doubles(h_vec.detach().cpu().numpy().reshape(-1))
    d = torch.empty((N, m), dtype=dt, device=p.device) if want_d else None
    inside = torch.empty((N,), dtype=torch.uint8, device=p.device) if want_inside else None
    if N > 0:
        with torch.cuda.device(p.device):
            _lib.check(lib.se_safety_distance(code, n_s, m, hm_p, hv_p, float(c_safety), _lib.ptr(p), _lib.ptr(q), N,
                                              _lib.ptr(d), _lib.ptr(inside), _lib.stream_ptr(p.device)))
    return d, inside


def lin_ellipsoid_safety_distance(p_center: Tensor, q_shape: Tensor, h_mat: Tensor, h_vec: Tensor,
                                  c_safety: float = 1.0) -> Tensor:
    """Drop-in for gp_reachability_pytorch.lin_ellipsoid_safety_distance (:184-215) -> d_safety [N, m]."""
    return _safety(p_center, q_shape, h_mat, h_vec, c_safety, True, False)[0]


def is_ellipsoid_inside_polytope(p_center: Tensor, q_shape: Tensor, h_mat: Tensor, h_vec: Tensor) -> Tensor:
    """Drop-in for gp_reachability_pytorch.is_ellipsoid_inside_polytope (:218-231) -> bool [N]."""
    return _safety(p_center, q_shape, h_mat, h_vec, 1.0, False, True)[1].bool()
