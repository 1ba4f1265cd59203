#!/usr/bin/env python
"""Model-refresh measurements on one GPU: the covariance-assembly and MLL-gradient kernels
against their roofline, and one hyper-parameter training iteration (kernel path) next to torch autograd through the
Cholesky factorisation on the same GPU (what `_train_model` did before) and on the host cores.

    python tools/train_bench.py [--configs C2,C3] > profiles/r01_train_bench.jsonl
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import safe_exploration_b200 as se  # noqa: E402
from safe_exploration_b200 import _lib, workloads  # noqa: E402


def cuda_ms(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--configs', default='C2,C3')
    ap.add_argument('--cpu', action='store_true', help='also time one oracle (host, autograd) iteration')
    args = ap.parse_args()
    lib = _lib.load()
    peaks = {}
    if os.path.exists('MEASURED_PEAKS.json'):
        peaks = json.load(open('MEASURED_PEAKS.json'))
    hbm = peaks.get('hbm_gbs', 6540.0)
    import ctypes as C
    f64 = C.c_double(0)
    _lib.check(lib.se_measure_peak(1, C.byref(f64)))
    for name in args.configs.split(','):
        w = workloads.cartpole_workload(num_rollouts=256) if name == 'C3' else workloads.pendulum_workload(num_rollouts=256)
        n_s, d = w.y_train.shape[1], w.x_train.shape[1]
        x = torch.as_tensor(w.x_train, dtype=torch.float64, device='cuda:0')
        y = torch.as_tensor(w.y_train, dtype=torch.float64, device='cuda:0')
        n = x.size(0)
        ls = torch.as_tensor(w.lengthscale, dtype=torch.float64).reshape(n_s, d)
        os_ = torch.as_tensor(w.outputscale, dtype=torch.float64).reshape(n_s)
        nz = torch.as_tensor(w.noise, dtype=torch.float64).reshape(n_s)
        gp = se.B200GpCemSSM(None, n_s, d - n_s, ls, os_, nz, device='cuda:0')
        keep, (p_ls, p_os, p_nz) = gp._hyp_ptrs(ls, os_, nz)
        k = torch.empty((n_s, n, n), dtype=torch.float64, device='cuda:0')
        st = _lib.stream_ptr('cuda:0')
        t_k = cuda_ms(lambda: _lib.check(lib.se_gp_kernel_matrix(n_s, d, n, _lib.ptr(x), p_ls, p_os, p_nz, _lib.ptr(k), st)))
        chol = torch.linalg.cholesky(k)
        kinv = torch.cholesky_inverse(chol)
        kinv = kinv if kinv.is_contiguous() else kinv.transpose(1, 2).contiguous()
        alpha = torch.cholesky_solve(y.t().unsqueeze(-1), chol).squeeze(-1).contiguous()
        ws = torch.empty(lib.se_gp_mll_grad_workspace(n_s, d, n) // 8, dtype=torch.float64, device='cuda:0')
        grad = torch.empty((n_s, d + 2), dtype=torch.float64, device='cuda:0')
        t_g = cuda_ms(lambda: _lib.check(lib.se_gp_mll_grad(n_s, d, n, _lib.ptr(x), _lib.ptr(alpha), _lib.ptr(kinv), p_ls,
                                                             p_os, p_nz, _lib.ptr(ws), _lib.ptr(grad), st)))
        t_chol = cuda_ms(lambda: torch.linalg.cholesky(k))
        t_inv = cuda_ms(lambda: torch.cholesky_inverse(chol))
        t_iter = cuda_ms(lambda: gp.mll_and_grad(x, y, ls, os_, nz), reps=3)

        # the previous route: autograd through the factorisation, on the same GPU (and, with --cpu, on the host cores)
        def autograd_iter(dev='cuda:0'):
            from safe_exploration_b200.ssm import _rbf_ard
            xd, yd = x.to(dev), y.to(dev)
            pl = [t.to(dev).clone().requires_grad_(True) for t in (ls, os_, nz)]
            kk = _rbf_ard(xd, xd, pl[0], pl[1]) + torch.diag_embed(pl[2].view(-1, 1).expand(-1, n))
            ch = torch.linalg.cholesky(kk)
            a = torch.cholesky_solve(yd.t().unsqueeze(-1), ch).squeeze(-1)
            mll = -0.5 * (yd.t() * a).sum(-1) - ch.diagonal(dim1=1, dim2=2).log().sum(-1)
            (-(mll / n).sum()).backward()
        t_auto = cuda_ms(autograd_iter, reps=3)
        rec = {
            'config': name, 'n_train': n, 'n_s': n_s, 'd': d,
            'kernel_matrix': {'ms': t_k, 'bytes_written': 8.0 * n_s * n * n, 'gbs': 8e-6 * n_s * n * n / t_k,
                              'frac_hbm': 8e-6 * n_s * n * n / t_k / hbm,
                              'fp64_flop_per_element': 3 * d + 27, 'tflops': (3 * d + 27) * n_s * n * n / t_k * 1e-9,
                              'fp64_peak_tflops': f64.value * 1e-12},
            'mll_grad': {'ms': t_g, 'bytes_read': 4.0 * n_s * n * n, 'gbs': 4e-6 * n_s * n * n / t_g,
                         'frac_hbm': 4e-6 * n_s * n * n / t_g / hbm,
                         'fp64_flop_per_element': 5 * d + 31, 'tflops': (5 * d + 31) * n_s * n * n / 2 / t_g * 1e-9},
            'cholesky_ms (cuSOLVER via torch)': t_chol, 'cholesky_inverse_ms (cuSOLVER via torch)': t_inv,
            'training_iteration_ms': t_iter, 'training_iteration_autograd_same_gpu_ms': t_auto,
            'hbm_peak_gbs': hbm,
        }
        if args.cpu:
            t0 = time.perf_counter()
            autograd_iter('cpu')
            rec['training_iteration_autograd_cpu_ms'] = (time.perf_counter() - t0) * 1e3
            rec['cpu_threads'] = torch.get_num_threads()
        print(json.dumps(rec), flush=True)


if __name__ == '__main__':
    main()
