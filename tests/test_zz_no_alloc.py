"""GPU test of `se_gp_mll_grad_static` (include/secem.h): the marginal-log-likelihood gradient with its per-tile
partial sums in the library's file-scope static device array instead of a caller workspace, an entry point that calls
no cudaMalloc.  It has no counterpart in the reference; the check is
against the existing entry point `se_gp_mll_grad` through the C ABI.

Tolerance: the two entry points launch the same two kernels with the same fixed reduction order, so the results must be
bit-identical; the assertion still allows the 1e-8 (relative to the output's largest gradient entry) that
tests/test_gpu_train.py uses for this gradient and additionally reports exact equality.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from safe_exploration_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _problem(n, n_s, d, seed):
    """Training inputs, hyper-parameters and the (alpha, K^-1) the gradient kernels read, built with torch on the
    device (set-up only; the code under test is the two C-ABI calls)."""
    g = torch.Generator().manual_seed(seed)
    x = (torch.rand((n, d), generator=g, dtype=torch.float64) * 2 - 1).to(DEV)
    y = torch.randn((n_s, n), generator=g, dtype=torch.float64).to(DEV) * 0.3
    ls = 0.5 + torch.rand((n_s, d), generator=g, dtype=torch.float64)
    os_ = 0.1 + torch.rand((n_s,), generator=g, dtype=torch.float64)
    nz = 1e-2 * (1 + torch.rand((n_s,), generator=g, dtype=torch.float64))
    lib = _lib.load()
    k = torch.empty((n_s, n, n), dtype=torch.float64, device=DEV)
    keep = [_lib.host_doubles(v) for v in (ls, os_, nz)]
    hyp = [p for _, p in keep]
    _lib.check(lib.se_gp_kernel_matrix(n_s, d, n, _lib.ptr(x), *hyp, _lib.ptr(k), _lib.stream_ptr(DEV)))
    kinv = torch.linalg.inv(k).contiguous()
    kinv = (0.5 * (kinv + kinv.transpose(1, 2))).contiguous()
    alpha = torch.matmul(kinv, y.unsqueeze(-1)).squeeze(-1).contiguous()
    return x, alpha, kinv, keep, hyp


def _grad_with_workspace(lib, n_s, d, n, x, alpha, kinv, hyp):
    ws = torch.empty(max(1, lib.se_gp_mll_grad_workspace(n_s, d, n) // 8), dtype=torch.float64, device=DEV)
    grad = torch.full((n_s, d + 2), float('nan'), dtype=torch.float64, device=DEV)
    _lib.check(lib.se_gp_mll_grad(n_s, d, n, _lib.ptr(x), _lib.ptr(alpha), _lib.ptr(kinv), *hyp, _lib.ptr(ws),
                                  _lib.ptr(grad), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    return grad.cpu()


def _grad_static(lib, n_s, d, n, x, alpha, kinv, hyp, grad):
    """`grad` is preallocated by the caller: between here and the synchronize nothing allocates."""
    _lib.check(lib.se_gp_mll_grad_static(n_s, d, n, _lib.ptr(x), _lib.ptr(alpha), _lib.ptr(kinv), *hyp,
                                         _lib.ptr(grad), _lib.stream_ptr(DEV)))
    torch.cuda.synchronize()
    return grad.cpu()


def test_static_scratch_gradient_equals_workspace_gradient_on_two_calls_in_a_row():
    lib = _lib.load()
    assert lib.se_gp_mll_grad_static_max_bytes() == 4 << 20
    # two different problems back to back: the second, smaller one runs on scratch still holding the first one's
    # partial sums (stale contents must not leak into its result), then the first again
    cases = [(700, 4, 5, 11), (130, 2, 3, 12), (700, 4, 5, 11)]
    prepared = [(c, _problem(c[0], c[1], c[2], c[3])) for c in cases[:2]]
    prepared.append(prepared[0])
    want = [_grad_with_workspace(lib, n_s, d, n, x, alpha, kinv, hyp)
            for (n, n_s, d, _), (x, alpha, kinv, keep, hyp) in prepared]
    outs = [torch.full((n_s, d + 2), float('nan'), dtype=torch.float64, device=DEV) for (n, n_s, d, _), _ in prepared]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(DEV)
    got = [_grad_static(lib, n_s, d, n, x, alpha, kinv, hyp, out)
           for ((n, n_s, d, _), (x, alpha, kinv, keep, hyp)), out in zip(prepared, outs)]
    assert torch.cuda.memory_allocated(DEV) == before
    for g, w in zip(got, want):
        assert torch.isfinite(g).all() and torch.isfinite(w).all()
        scale = w.abs().amax(dim=1, keepdim=True)
        assert float(((g - w).abs() / scale).max()) <= 1e-8
        assert torch.equal(g, w), 'same kernels, same reduction order: the two entry points must agree bit for bit'


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    lib = _lib.load()
    n_s, d = 6, 9
    n = 8000
    assert lib.se_gp_mll_grad_workspace(n_s, d, n) <= lib.se_gp_mll_grad_static_max_bytes()
    n = 9000
    assert lib.se_gp_mll_grad_workspace(n_s, d, n) > lib.se_gp_mll_grad_static_max_bytes()
    # the size check comes before any device access: dummy (never dereferenced) buffers are enough
    dummy = torch.zeros(16, dtype=torch.float64, device=DEV)
    keep = [_lib.host_doubles(np.ones((n_s, d))), _lib.host_doubles(np.ones(n_s)), _lib.host_doubles(np.ones(n_s))]
    hyp = [p for _, p in keep]
    rc = lib.se_gp_mll_grad_static(n_s, d, n, _lib.ptr(dummy), _lib.ptr(dummy), _lib.ptr(dummy), *hyp,
                                   _lib.ptr(dummy), _lib.stream_ptr(DEV))
    assert rc != 0
    msg = lib.se_last_error().decode()
    assert 'SE_MLL_GRAD_STATIC_MAX_BYTES' in msg and 'se_gp_mll_grad' in msg
    torch.cuda.synchronize()
