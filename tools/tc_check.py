"""GPU check of the tcgen05 digit-sliced variance GEMM (SE_GEMM_TC) against the DMMA path on the bench workloads."""
import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from safe_exploration_b200 import _lib, workloads
from safe_exploration_b200.ssm import B200GpCemSSM

def run(w, N, seed=0):
    lib = _lib.load()
    ssm = B200GpCemSSM(None, w.n_s, w.n_u, lengthscale=w.lengthscale, outputscale=w.outputscale, noise=w.noise)
    ssm.update_model(torch.as_tensor(w.x_train, device='cuda'), torch.as_tensor(w.y_train, device='cuda'), replace_old=True)
    g = torch.Generator(device='cpu').manual_seed(seed)
    d = w.n_s + w.n_u
    lo, hi = w.x_train.min(0), w.x_train.max(0)
    z = torch.rand(N, d, generator=g, dtype=torch.float64) * torch.as_tensor(hi - lo) + torch.as_tensor(lo)
    z[: N // 2] *= 0.2
    z = z.cuda()
    ssm.set_gemm_impl(_lib.SE_GEMM_MMA)
    m0, v0, j0 = ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
    torch.cuda.synchronize()
    for S in (8, 7, 6, 5):
        ssm.set_gemm_impl(_lib.SE_GEMM_TC)
        ssm.set_tc_digits(S)
        m1, v1, j1 = ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
        torch.cuda.synchronize()
        rel = ((v1 - v0).abs() / v0.abs())
        print(f'{w.name} n={w.n_train} N={N} digits={S}: var rel err max {rel.max().item():.3e} median {rel.median().item():.3e}; '
              f'mean max abs diff {(m1-m0).abs().max().item():.2e}; jac max abs diff {(j1-j0).abs().max().item():.2e}; '
              f'var range [{v0.min().item():.3e}, {v0.max().item():.3e}]', flush=True)
        t0 = time.perf_counter()
        for _ in range(3):
            ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
        torch.cuda.synchronize()
        print(f'   tc predict: {(time.perf_counter()-t0)/3*1e3:.2f} ms', flush=True)
    ssm.set_gemm_impl(_lib.SE_GEMM_MMA)
    t0 = time.perf_counter()
    for _ in range(3):
        ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
    torch.cuda.synchronize()
    print(f'   dmma predict: {(time.perf_counter()-t0)/3*1e3:.2f} ms', flush=True)

run(workloads.pendulum_workload(n_train=100, name='C1'), 500)
run(workloads.pendulum_workload(n_train=300, name='Cx'), 1000)
run(workloads.pendulum_workload(), 16384)
if len(sys.argv) > 1:
    run(workloads.cartpole_workload(), 16384)
