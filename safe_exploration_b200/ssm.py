"""State-space models behind the reference's `CemSSM` interface (safe_exploration/ssm_cem/ssm_cem.py:11-131).

`B200GpCemSSM` is the drop-in for `GpCemSSM` (safe_exploration/ssm_cem/gp_ssm_cem.py:17-133, rbf kernel): same five
methods, same shapes, same error behaviour (`ValueError` on a shape mismatch); the posterior itself is evaluated by
libsecem.so (csrc/gp_kernels.cuh).  PyTorch only owns device memory here and, at `update_model` time, factorises the
training covariance (torch.linalg -> cuSOLVER), which is off the hot path.
"""
import ctypes as C
import math
from abc import ABC, abstractmethod
from typing import Any, Dict, Optional, Tuple

import torch
from torch import Tensor

from . import _lib


def assert_shape(x, shape: tuple, ignore_if_none=False) -> None:
    """safe_exploration/utils.py:640-648."""
    if x is None:
        if ignore_if_none:
            return
        raise ValueError(f'Wanted shape {shape}, got None')
    if tuple(x.shape) != tuple(shape):
        raise ValueError(f'Wanted shape {shape}, got {x.shape}')


class CemSSM(ABC):
    """Mirror of the reference ABC (ssm_cem.py:11-131): same attributes, same method contract."""

    def __init__(self, state_dimen: int, action_dimen: int):
        self.num_states = state_dimen
        self.num_actions = action_dimen
        self._x_train = None
        self._y_train = None

    @property
    def x_train(self) -> Optional[Tensor]:
        return self._x_train

    @property
    def y_train(self) -> Optional[Tensor]:
        return self._y_train

    @abstractmethod
    def predict_with_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """-> mean [N, n_s], variance [N, n_s], d mean / d (state, action) [N, n_s, n_s + n_u]."""

    @abstractmethod
    def predict_without_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        """-> mean [N, n_s], variance [N, n_s]."""

    @abstractmethod
    def predict_raw(self, z: Tensor) -> Tuple[Tensor, Tensor]:
        """Predict from stacked [state | action] inputs."""

    def update_model(self, train_x: Tensor, train_y: Tensor, opt_hyp=False, replace_old=False) -> None:
        """Contract of ssm_cem.py:72-99: the new samples replace the stored training set (`replace_old`, or nothing stored
        yet) or are appended to it; the model is rebuilt on the result and, for parametric models or on request
        (`opt_hyp`), trained."""
        rows = train_x.size(0)
        assert_shape(train_x, (rows, self.num_states + self.num_actions))
        assert_shape(train_y, (rows, self.num_states))
        keep_old = not replace_old and self._x_train is not None and self._y_train is not None
        self._x_train = torch.cat((self._x_train, train_x), dim=0) if keep_old else train_x
        self._y_train = torch.cat((self._y_train, train_y), dim=0) if keep_old else train_y
        self._update_model(self._x_train, self._y_train)
        if opt_hyp or self.parametric:
            self._train_model(self._x_train, self._y_train)

    @abstractmethod
    def _update_model(self, x_train: Tensor, y_train: Tensor) -> None:
        pass

    @abstractmethod
    def _train_model(self, x_train: Tensor, y_train: Tensor) -> None:
        pass

    def _join_states_actions(self, states: Tensor, actions: Tensor) -> Tensor:
        """z = [state | action] per row, after checking that both have the model's dimensions."""
        assert_shape(states, (states.size(0), self.num_states))
        assert_shape(actions, (states.size(0), self.num_actions))
        return torch.cat((states, actions), dim=1)

    @abstractmethod
    def collect_metrics(self) -> Dict[str, Any]:
        pass

    @property
    @abstractmethod
    def parametric(self) -> bool:
        pass


def _rbf_ard(xa: Tensor, xb: Tensor, lengthscale: Tensor, outputscale: Tensor) -> Tensor:
    """[n_s, A, B] covariance of ScaleKernel(RBFKernel(ard)) (gp_ssm_cem.py:45-57); fp64, used for the factors."""
    sa = xa.unsqueeze(0) / lengthscale.unsqueeze(1)  # [n_s, A, d]
    sb = xb.unsqueeze(0) / lengthscale.unsqueeze(1)
    d2 = None
    for dd in range(xa.size(1)):  # explicit differences (no a^2 + b^2 - 2ab cancellation), O(n_s A B) memory
        diff = sa[:, :, dd].unsqueeze(2) - sb[:, :, dd].unsqueeze(1)
        d2 = diff * diff if d2 is None else d2.addcmul_(diff, diff)
    return outputscale.view(-1, 1, 1) * torch.exp(-0.5 * d2)


class B200GpCemSSM(CemSSM):
    """Exact-GP state-space model evaluated by hand-written sm_100a kernels.

    Hyper-parameters cross the boundary as materialised values (lengthscale [n_s, d], outputscale [n_s],
    noise [n_s]), never as GPyTorch raw parameters.  Defaults are GPyTorch's untrained values softplus(0) = ln 2,
    which is what the reference's tests run with (test_gp_reachability_pytorch.py:96-97).

    :param conf: optional config object; `exact_gp_training_iterations` and `exact_gp_kernel` are read if present
                 (only 'rbf' is accelerated; the reference's 'linear' / 'nn' kernels stay on GPyTorch).
    :param dtype: torch.float64 (reference precision, default) or torch.float32 (the whole path in float32 FFMA
                  arithmetic; kept for parity studies -- it is SLOWER than the default, whose variance contraction runs
                  on the tensor cores).  For float32-class tolerance at higher speed keep float64 and pass tc_digits=4
    :param tc_digits: digit planes of the tensor-core variance GEMM: 0 = float64-class, chosen from outputscale/noise
                  (never fewer than 6); 7 / 8 = float64-equivalent; 4 / 3 = float32 tolerance (see set_tc_digits)
    :param incremental: keep the float64 inverse Cholesky factor on the device (W and W^T, 2 n_s n_pad^2 doubles with
                        the native factor engine) so that `update_model(..., replace_old=False)` only computes the block
                        rows that the appended samples touch
    :param factor_impl: 'native' (default): L^-1 and (K + noise I)^-1 come from this library's blocked factorisation
                        (`se_gp_factor`, `se_gp_kinv`: FP64 DMMA products, no cuSOLVER / cuBLAS), which also extends a
                        factor whose training set only grew; 'cusolver': the round-1 path through torch.linalg
    """

    def __init__(self, conf, state_dimen: int, action_dimen: int, lengthscale=None, outputscale=None, noise=None,
                 dtype: torch.dtype = torch.float64, device: Optional[str] = None, gemm_impl: int = _lib.SE_GEMM_AUTO,
                 workspace_bytes: Optional[int] = None, tc_digits: int = 0, incremental: bool = True,
                 factor_impl: str = 'native'):
        super().__init__(state_dimen, action_dimen)
        kernel = getattr(conf, 'exact_gp_kernel', 'rbf') if conf is not None else 'rbf'
        if kernel != 'rbf':
            raise ValueError(f'Unknown kernel {kernel}: the B200 path implements the rbf (ARD) kernel')
        self._training_iterations = getattr(conf, 'exact_gp_training_iterations', 1000) if conf is not None else 1000
        self._lib = _lib.load()
        self.dtype = dtype
        self._code = _lib.dtype_code(dtype)
        self.device = torch.device(device if device is not None else 'cuda:0')
        d = state_dimen + action_dimen
        ln2 = math.log(2.0)
        f64 = dict(dtype=torch.float64, device='cpu')
        self.lengthscale = torch.full((state_dimen, d), ln2, **f64) if lengthscale is None else \
            torch.as_tensor(lengthscale, **f64).reshape(state_dimen, d).clone()
        self.outputscale = torch.full((state_dimen,), ln2, **f64) if outputscale is None else \
            torch.as_tensor(outputscale, **f64).reshape(state_dimen).clone()
        self.noise = torch.full((state_dimen,), ln2, **f64) if noise is None else \
            torch.as_tensor(noise, **f64).reshape(state_dimen).clone()
        self._gemm_impl = gemm_impl
        self._incremental = incremental  # keep the float64 inverse factor so appended samples cost O(m n^2)
        if factor_impl not in ('native', 'cusolver'):
            raise ValueError(f"factor_impl must be 'native' or 'cusolver', got {factor_impl!r}")
        self._factor_impl = factor_impl
        self._factor_pad = None          # native path: (x, W_pad, W^T_pad, hyper-parameters) of the current handle
        self._factor = None              # (x [n, d], L^-1 [n_s, n, n], hyper-parameters) of the current handle
        self._tc_digits = tc_digits
        self._workspace_bytes = workspace_bytes
        self._handle = C.c_void_p(None)
        self.generation = 0  # bumped by every _update_model: dependents (CEM handles) key on it, never on the address
        self._last_training_losses = []

    # ---- handle management -------------------------------------------------------------------------------------
    @property
    def handle(self) -> C.c_void_p:
        if not self._handle:
            raise _lib.SecemError('the model holds no training data yet: call update_model first')
        return self._handle

    def _destroy(self):
        if getattr(self, '_handle', None):
            self._lib.se_gp_destroy(self._handle)
            self._handle = C.c_void_p(None)

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def set_gemm_impl(self, impl: int) -> None:
        self._gemm_impl = impl
        if self._handle:
            _lib.check(self._lib.se_gp_set_gemm_impl(self._handle, impl))

    def set_tc_digits(self, digits: int) -> None:
        """Digit planes of the tcgen05 variance GEMM: 0 = chosen from outputscale / noise (default, float64-class), 5 .. 8,
        or 4 / 3 for float32-class tolerance (1e-7 / 3e-5 of the prior scale) at 2.1x / 3.5x fewer tensor-core MMAs."""
        self._tc_digits = digits
        if self._handle:
            _lib.check(self._lib.se_gp_set_tc_digits(self._handle, digits))

    def gemm_info(self) -> Tuple[str, int]:
        """(implementation the variance GEMM resolved to, digit planes in use -- 0 before the first evaluation)."""
        impl, digits = C.c_int(0), C.c_int(0)
        _lib.check(self._lib.se_gp_get_gemm_info(self.handle, C.byref(impl), C.byref(digits)))
        return {_lib.SE_GEMM_SIMT: 'simt', _lib.SE_GEMM_MMA: 'dmma', _lib.SE_GEMM_TC: 'tcgen05-i8'}[impl.value], digits.value

    # ---- CemSSM ------------------------------------------------------------------------------------------------
    def _update_model(self, x_train: Tensor, y_train: Tensor) -> None:
        """gp_ssm_cem.py:96-101.  Builds alpha = (K + noise I)^-1 y and L^-1 in fp64 and uploads them."""
        x = x_train.detach().to(device=self.device, dtype=torch.float64)
        y = y_train.detach().to(device=self.device, dtype=torch.float64)
        n = x.size(0)
        hyp = torch.cat((self.lengthscale.reshape(-1), self.outputscale, self.noise))
        if self._factor_impl == 'native':
            w_pad, first_block = self._factor_native(x, hyp)
            self.last_update = 'full' if first_block == 0 else 'incremental'
            linv = w_pad[:, :n, :n]
            alpha = torch.matmul(linv.transpose(1, 2), torch.matmul(linv, y.t().unsqueeze(-1))).squeeze(-1)
        else:
            linv = self._append_rows(x, hyp) if self._incremental else None
            self.last_update = 'full' if linv is None else 'incremental'
            if linv is None:
                k = self.kernel_matrix(x)
                chol = torch.linalg.cholesky(k)
                del k
                eye = torch.eye(n, dtype=torch.float64, device=self.device).expand(self.num_states, -1, -1)
                linv = torch.linalg.solve_triangular(chol, eye, upper=False).tril_()
                # alpha = (L L^T)^-1 y by two triangular solves (cholesky_solve's batched cuSOLVER path rejects n = 16384)
                alpha = torch.linalg.solve_triangular(chol, y.t().unsqueeze(-1), upper=False)
                alpha = torch.linalg.solve_triangular(chol.transpose(1, 2), alpha, upper=True).squeeze(-1)
                del chol
            else:
                alpha = torch.matmul(linv.transpose(1, 2), torch.matmul(linv, y.t().unsqueeze(-1))).squeeze(-1)
        if self._incremental and self._factor_impl != 'native':
            self._factor = (x, linv, hyp)
        xs = x.to(self.dtype).contiguous()
        al = alpha.to(self.dtype).contiguous()
        li = linv.to(self.dtype).contiguous()
        _, p_ls = _lib.host_doubles(self.lengthscale.numpy())
        _, p_os = _lib.host_doubles(self.outputscale.numpy())
        _, p_nz = _lib.host_doubles(self.noise.numpy())
        self._destroy()
        handle = C.c_void_p(None)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.se_gp_create(C.byref(handle), self.num_states, self.num_actions, n, self._code,
                                              _lib.ptr(xs), _lib.ptr(al), _lib.ptr(li), p_ls, p_os, p_nz,
                                              _lib.stream_ptr(self.device)))
            torch.cuda.current_stream(self.device).synchronize()  # xs / al / li may be freed after this
        self._handle = handle
        self.generation += 1
        if self._gemm_impl != _lib.SE_GEMM_AUTO:
            _lib.check(self._lib.se_gp_set_gemm_impl(handle, self._gemm_impl))
        if self._tc_digits:
            _lib.check(self._lib.se_gp_set_tc_digits(handle, self._tc_digits))
        if self._workspace_bytes is not None:
            _lib.check(self._lib.se_gp_set_workspace_limit(handle, int(self._workspace_bytes)))

    def _padded_covariance(self, x: Tensor, n_pad: int, ls=None, os_=None, nz=None) -> Tensor:
        """[n_s, n_pad, n_pad]: K + noise I on the training inputs, the identity on the padding."""
        n = x.size(0)
        k_pad = torch.zeros((self.num_states, n_pad, n_pad), dtype=torch.float64, device=x.device)
        k_pad[:, :n, :n] = self.kernel_matrix(x, ls, os_, nz)
        if n_pad > n:
            k_pad.diagonal(dim1=1, dim2=2)[:, n:] = 1.0
        return k_pad

    def _run_factor(self, k_pad: Tensor, w_pad: Tensor, wt_pad: Tensor, first_block: int) -> Tensor:
        """`se_gp_factor` on padded buffers -> diag(L) [n_s, n_pad]; raises like torch.linalg.cholesky on a bad pivot."""
        n_pad = k_pad.size(1)
        dev = k_pad.device
        diag = torch.ones((self.num_states, n_pad), dtype=torch.float64, device=dev)
        ws = torch.empty(max(1, self._lib.se_gp_factor_workspace(self.num_states, n_pad) // 8), dtype=torch.float64,
                         device=dev)
        info = torch.zeros((self.num_states,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(self._lib.se_gp_factor(self.num_states, n_pad, first_block, _lib.ptr(k_pad), _lib.ptr(w_pad),
                                              _lib.ptr(wt_pad), _lib.ptr(diag), _lib.ptr(ws), _lib.ptr(info),
                                              _lib.stream_ptr(dev)))
        bad = info.cpu()
        if int(bad.max()) != 0:
            o = int(torch.nonzero(bad)[0])
            raise torch.linalg.LinAlgError(f'se_gp_factor: the covariance of output {o} is not positive definite '
                                           f'(pivot of row {int(bad[o]) - 1})')
        return diag

    def _factor_native(self, x: Tensor, hyp: Tensor) -> Tuple[Tensor, int]:
        """W = L^-1 on the device through `se_gp_factor`, padded to a multiple of 128 rows.  When the previous factor was
        built on a prefix of `x` with the same hyper-parameters (update_model(replace_old=False), ssm_cem.py:72-99; the
        idea of ssm_pytorch/utilities.py:122-165), the block rows that lie entirely inside the old training set are kept
        and only the later ones are computed: O((n - n_kept) n^2) instead of O(n^3).  -> (W_pad, first block row computed)"""
        n = x.size(0)
        n_pad = -(-n // 128) * 128
        first_block = 0
        w_pad = torch.zeros((self.num_states, n_pad, n_pad), dtype=torch.float64, device=x.device)
        wt_pad = torch.zeros_like(w_pad)
        prev = self._factor_pad if self._incremental else None
        if prev is not None:
            x_old, w_old, wt_old, hyp_old = prev
            n_old = x_old.size(0)
            if (n_old < n and hyp_old.shape == hyp.shape and torch.equal(hyp_old, hyp) and
                    torch.equal(x[:n_old], x_old)):
                first_block = n_old // 128  # the old factor's last, partly padded block row is recomputed
                keep = first_block * 128
                w_pad[:, :keep, :keep] = w_old[:, :keep, :keep]
                wt_pad[:, :keep, :keep] = wt_old[:, :keep, :keep]
        k_pad = self._padded_covariance(x, n_pad)
        self._run_factor(k_pad, w_pad, wt_pad, first_block)
        del k_pad
        self._factor_pad = (x, w_pad, wt_pad, hyp) if self._incremental else None
        return w_pad, first_block

    def _append_rows(self, x: Tensor, hyp: Tensor) -> Optional[Tensor]:
        """Incremental factor update (the idea of ssm_pytorch/utilities.py:122-165 `update_cholesky`):
        when `update_model` appended m samples to the n the factor was built on, with unchanged hyper-parameters,

            L' = [[L, 0], [B, C]],  B = K(new, old) L^-T,  C C^T = K(new, new) + noise I - B B^T
            L'^-1 = [[L^-1, 0], [-C^-1 B L^-1, C^-1]]

        so the first n rows of L^-1 are kept and only m new rows are computed: O(m n^2) instead of O(n^3).
        Returns the new [n_s, n + m, n + m] inverse factor, or None when a full factorisation is needed."""
        if self._factor is None:
            return None
        x_old, linv_old, hyp_old = self._factor
        n_old, n = x_old.size(0), x.size(0)
        if not (n_old < n and hyp_old.shape == hyp.shape and torch.equal(hyp_old, hyp) and
                torch.equal(x[:n_old], x_old)):
            return None
        ls, os_, nz = (t.to(self.device) for t in (self.lengthscale, self.outputscale, self.noise))
        x_add = x[n_old:]
        m = n - n_old
        k_cross = _rbf_ard(x_add, x_old, ls, os_)                       # [n_s, m, n]
        b = torch.matmul(k_cross, linv_old.transpose(1, 2))             # B = K(new, old) L^-T
        s = _rbf_ard(x_add, x_add, ls, os_) - torch.matmul(b, b.transpose(1, 2))
        s.diagonal(dim1=1, dim2=2).add_(nz.view(-1, 1))
        c = torch.linalg.cholesky(s)
        c_inv = torch.linalg.solve_triangular(c, torch.eye(m, dtype=torch.float64, device=self.device)
                                              .expand(self.num_states, -1, -1), upper=False).tril_()
        low = -torch.matmul(c_inv, torch.matmul(b, linv_old))           # -C^-1 B L^-1
        linv = torch.zeros((self.num_states, n, n), dtype=torch.float64, device=self.device)
        linv[:, :n_old, :n_old] = linv_old
        linv[:, n_old:, :n_old] = low
        linv[:, n_old:, n_old:] = c_inv
        return linv

    # ---- model refresh kernels -----------------------------------------------------
    def _hyp_ptrs(self, ls: Tensor, os_: Tensor, nz: Tensor):
        keep = [_lib.host_doubles(t.detach().cpu().numpy()) for t in (ls, os_, nz)]
        return keep, [k[1] for k in keep]

    def kernel_matrix(self, x: Tensor, lengthscale=None, outputscale=None, noise=None) -> Tensor:
        """[n_s, n, n] float64 covariance K_o + noise_o I of the training inputs (`se_gp_kernel_matrix`): what GPyTorch
        builds from ScaleKernel(RBFKernel(ard)) + GaussianLikelihood (gp_ssm_cem.py:33, 45-57)."""
        _lib.require_cuda(x, 'x')
        x = x.detach().to(torch.float64).contiguous()
        n, d = x.shape
        ls, os_, nz = (self.lengthscale if lengthscale is None else lengthscale,
                       self.outputscale if outputscale is None else outputscale,
                       self.noise if noise is None else noise)
        keep, (p_ls, p_os, p_nz) = self._hyp_ptrs(ls, os_, nz)
        k = torch.empty((self.num_states, n, n), dtype=torch.float64, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(self._lib.se_gp_kernel_matrix(self.num_states, d, n, _lib.ptr(x), p_ls, p_os, p_nz, _lib.ptr(k),
                                                     _lib.stream_ptr(x.device)))
        return k

    def mll_and_grad(self, x: Tensor, y: Tensor, lengthscale=None, outputscale=None, noise=None):
        """Exact marginal log likelihood of each output's GP and its gradient w.r.t. the materialised hyper-parameters.

        :param x: [n, d] training inputs (CUDA), :param y: [n, n_s] training targets
        :returns: mll [n_s] (not divided by n), grad [n_s, d + 2] = d mll_o / d (lengthscale_o[0..d), outputscale_o,
                  noise_o), alpha [n_s, n].  Covariance assembly and gradient are kernels of libsecem.so
                  (`se_gp_kernel_matrix`, `se_gp_mll_grad`); the Cholesky factorisation and inverse in between are
                  cuSOLVER calls through torch.linalg."""
        _lib.require_cuda(x, 'x')
        x = x.detach().to(torch.float64).contiguous()
        yt = y.detach().to(device=x.device, dtype=torch.float64).t().contiguous()
        n, d = x.shape
        ls, os_, nz = (self.lengthscale if lengthscale is None else lengthscale,
                       self.outputscale if outputscale is None else outputscale,
                       self.noise if noise is None else noise)
        keep, (p_ls, p_os, p_nz) = self._hyp_ptrs(ls, os_, nz)
        if self._factor_impl == 'native':
            n_pad = -(-n // 128) * 128
            k_pad = self._padded_covariance(x, n_pad, ls, os_, nz)
            w_pad = torch.zeros_like(k_pad)
            wt_pad = torch.zeros_like(k_pad)
            diag = self._run_factor(k_pad, w_pad, wt_pad, 0)
            logdet_half = diag[:, :n].log().sum(-1)
            kinv_pad = k_pad  # reuse the buffer: se_gp_kinv writes the lower triangle, the gradient kernel reads only that
            with torch.cuda.device(x.device):
                _lib.check(self._lib.se_gp_kinv(self.num_states, n_pad, _lib.ptr(wt_pad), _lib.ptr(kinv_pad),
                                                _lib.stream_ptr(x.device)))
            linv = w_pad[:, :n, :n]
            alpha = torch.matmul(linv.transpose(1, 2), torch.matmul(linv, yt.unsqueeze(-1))).squeeze(-1).contiguous()
            kinv = kinv_pad[:, :n, :n].contiguous()
            del w_pad, wt_pad, kinv_pad, k_pad, linv
        else:
            k = self.kernel_matrix(x, ls, os_, nz)
            chol = torch.linalg.cholesky(k)
            del k
            logdet_half = chol.diagonal(dim1=1, dim2=2).log().sum(-1)
            # K^-1 = L^-T L^-1 from one triangular solve and one DGEMM (potri is 3x slower at n = 2000 .. 5000)
            eye = torch.eye(n, dtype=torch.float64, device=x.device).expand(self.num_states, -1, -1)
            linv = torch.linalg.solve_triangular(chol, eye, upper=False)
            del chol
            kinv = torch.matmul(linv.transpose(1, 2), linv).contiguous()
            alpha = torch.matmul(linv.transpose(1, 2), torch.matmul(linv, yt.unsqueeze(-1))).squeeze(-1).contiguous()
            del linv
        mll = -0.5 * (yt * alpha).sum(-1) - logdet_half - 0.5 * n * math.log(2 * math.pi)
        ws = torch.empty(max(1, self._lib.se_gp_mll_grad_workspace(self.num_states, d, n) // 8), dtype=torch.float64,
                         device=x.device)
        grad = torch.empty((self.num_states, d + 2), dtype=torch.float64, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(self._lib.se_gp_mll_grad(self.num_states, d, n, _lib.ptr(x), _lib.ptr(alpha), _lib.ptr(kinv),
                                                p_ls, p_os, p_nz, _lib.ptr(ws), _lib.ptr(grad),
                                                _lib.stream_ptr(x.device)))
        return mll, grad, alpha

    def _train_model(self, x_train: Tensor, y_train: Tensor) -> None:
        """gp_ssm_cem.py:103-129: Adam(lr=0.01) on loss = -sum_o MLL_o / n (GPyTorch's ExactMarginalLogLikelihood
        divides by the number of data points), raw parameters mapped through softplus like GPyTorch's.  The gradient
        comes from `mll_and_grad` (analytic, one pass over K^-1) instead of autograd through the factorisation; the
        chain rule through softplus and torch.optim.Adam's update on the n_s (d + 2) raw scalars stay in torch."""
        x = x_train.detach().to(device=self.device, dtype=torch.float64).contiguous()
        y = y_train.detach().to(device=self.device, dtype=torch.float64)
        n, d = x.shape
        inv_softplus = lambda v: torch.log(torch.expm1(v))
        raw = torch.cat((inv_softplus(self.lengthscale), inv_softplus(self.outputscale).unsqueeze(1),
                         inv_softplus(self.noise).unsqueeze(1)), dim=1).to(self.device).requires_grad_(True)
        opt = torch.optim.Adam([raw], lr=0.01)
        losses = []
        for _ in range(self._training_iterations):
            host = torch.nn.functional.softplus(raw.detach()).cpu()
            mll, grad, _ = self.mll_and_grad(x, y, host[:, :d], host[:, d], host[:, d + 1])
            raw.grad = -(grad * torch.sigmoid(raw.detach())) / n
            losses.append(-(mll / n).sum())
            opt.step()
        self._last_training_losses = torch.stack(losses).cpu().tolist() if losses else []
        host = torch.nn.functional.softplus(raw.detach()).cpu()
        self.lengthscale, self.outputscale, self.noise = host[:, :d].clone(), host[:, d].clone(), host[:, d + 1].clone()
        self._update_model(x_train, y_train)

    def _prep(self, t: Tensor, name: str) -> Tensor:
        _lib.require_cuda(t, name)
        return t.detach().to(self.dtype).contiguous()

    def _predict_z(self, z: Tensor, want_jac: bool):
        N = z.size(0)
        z = self._prep(z, 'z')
        mean = torch.empty((N, self.num_states), dtype=self.dtype, device=z.device)
        var = torch.empty_like(mean)
        jac = torch.empty((N, self.num_states, self.num_states + self.num_actions), dtype=self.dtype,
                          device=z.device) if want_jac else None
        if N > 0:
            with torch.cuda.device(z.device):
                _lib.check(self._lib.se_gp_predict(self.handle, _lib.ptr(z), N, _lib.ptr(mean), _lib.ptr(var),
                                                   _lib.ptr(jac), _lib.stream_ptr(z.device)))
        return mean, var, jac

    def predict_with_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """gp_ssm_cem.py:59-73 (one fused pass instead of two forwards + one backward)."""
        z = self._join_states_actions(states, actions)
        return self._predict_z(z, True)

    def predict_without_jacobians(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        """gp_ssm_cem.py:75-77."""
        z = self._join_states_actions(states, actions)
        mean, var, _ = self._predict_z(z, False)
        return mean, var

    def predict_raw(self, z: Tensor) -> Tuple[Tensor, Tensor]:
        """gp_ssm_cem.py:90-94: [n_s, N] mean and variance (the GPyTorch batch layout)."""
        N = z.size(0)
        assert_shape(z, (N, self.num_states + self.num_actions))
        mean, var, _ = self._predict_z(z, False)
        return mean.t(), var.t()

    def collect_metrics(self) -> Dict[str, Any]:
        return {'losses': self._last_training_losses}

    @property
    def parametric(self) -> bool:
        return False
