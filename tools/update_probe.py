"""Where update_model spends its time (one GPU): covariance, Cholesky, triangular inverse, alpha, pack + digit image."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import safe_exploration_b200 as se
from safe_exploration_b200 import workloads

def ms(fn, reps=3):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts)[len(ts) // 2]

for name in sys.argv[1:] or ['C2', 'C3']:
    w = workloads.cartpole_workload(num_rollouts=256) if name == 'C3' else workloads.pendulum_workload(num_rollouts=256)
    n_s, d = w.y_train.shape[1], w.x_train.shape[1]
    x = torch.as_tensor(w.x_train, device='cuda:0'); y = torch.as_tensor(w.y_train, device='cuda:0')
    n = x.size(0)
    gp = se.B200GpCemSSM(None, n_s, d - n_s, w.lengthscale, w.outputscale, w.noise, device='cuda:0', incremental=False)
    rec = {'config': name, 'n': n, 'n_s': n_s}
    rec['update_model_ms'] = ms(lambda: gp.update_model(x, y, replace_old=True))
    gc = se.B200GpCemSSM(None, n_s, d - n_s, w.lengthscale, w.outputscale, w.noise, device='cuda:0', incremental=False,
                         factor_impl='cusolver')
    rec['update_model_cusolver_ms'] = ms(lambda: gc.update_model(x, y, replace_old=True))
    rec['train_iteration_native_ms'] = ms(lambda: gp.mll_and_grad(x, y))
    rec['train_iteration_cusolver_ms'] = ms(lambda: gc.mll_and_grad(x, y))
    n_pad = -(-n // 128) * 128
    k_pad = gp._padded_covariance(x, n_pad)
    w_pad = torch.zeros_like(k_pad); wt_pad = torch.zeros_like(k_pad)
    def fac():
        w_pad.zero_(); wt_pad.zero_()
        gp._run_factor(k_pad, w_pad, wt_pad, 0)
    rec['se_gp_factor_ms (incl. zeroing W, W^T)'] = ms(fac)
    del k_pad, w_pad, wt_pad
    gi = se.B200GpCemSSM(None, n_s, d - n_s, w.lengthscale, w.outputscale, w.noise, device='cuda:0', incremental=True)
    def append():
        gi.update_model(x[: n - 100], y[: n - 100], replace_old=True)
    append(); torch.cuda.synchronize()
    t_base = ms(append)
    def append2():
        gi.update_model(x[: n - 100], y[: n - 100], replace_old=True)
        gi.update_model(x[n - 100:], y[n - 100:])
    rec['append_100_rows_ms'] = ms(append2) - t_base
    k = gp.kernel_matrix(x)
    rec['kernel_matrix_ms'] = ms(lambda: gp.kernel_matrix(x))
    rec['cholesky_ms'] = ms(lambda: torch.linalg.cholesky(k))
    chol = torch.linalg.cholesky(k)
    eye = torch.eye(n, dtype=torch.float64, device='cuda:0').expand(n_s, -1, -1)
    rec['solve_triangular_eye_ms'] = ms(lambda: torch.linalg.solve_triangular(chol, eye, upper=False))
    rec['first_predict_ms (digit image)'] = None
    z = x[:256].contiguous()
    gp.update_model(x, y, replace_old=True); torch.cuda.synchronize(); t0 = time.perf_counter()
    gp.predict_raw(z); torch.cuda.synchronize(); rec['first_predict_ms (digit image)'] = (time.perf_counter() - t0) * 1e3
    print(json.dumps(rec), flush=True)
