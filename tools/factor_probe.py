"""One se_gp_factor + se_gp_kinv call on the C2 / C3 model sizes (run under `ncu --metrics gpu__time_duration.sum` for the per-kernel times)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_exploration_b200 as se
from safe_exploration_b200 import workloads, _lib
name = sys.argv[1] if len(sys.argv) > 1 else 'C2'
w = workloads.cartpole_workload(num_rollouts=256) if name == 'C3' else workloads.pendulum_workload(num_rollouts=256)
n_s, d = w.y_train.shape[1], w.x_train.shape[1]
x = torch.as_tensor(w.x_train, device='cuda:0')
gp = se.B200GpCemSSM(None, n_s, d - n_s, w.lengthscale, w.outputscale, w.noise, device='cuda:0', incremental=False)
n = x.size(0); n_pad = -(-n // 128) * 128
k_pad = gp._padded_covariance(x, n_pad)
w_pad = torch.zeros_like(k_pad); wt_pad = torch.zeros_like(k_pad)
for rep in range(2):
    w_pad.zero_(); wt_pad.zero_(); torch.cuda.synchronize(); t0 = time.perf_counter()
    gp._run_factor(k_pad, w_pad, wt_pad, 0); torch.cuda.synchronize()
    print(name, 'se_gp_factor', round((time.perf_counter() - t0) * 1e3, 2), 'ms')
kinv = torch.empty_like(k_pad)
lib = _lib.load()
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    _lib.check(lib.se_gp_kinv(n_s, n_pad, _lib.ptr(wt_pad), _lib.ptr(kinv), _lib.stream_ptr('cuda:0'))); torch.cuda.synchronize()
    print(name, 'se_gp_kinv', round((time.perf_counter() - t0) * 1e3, 2), 'ms')
