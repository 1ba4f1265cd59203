"""Times se_gp_predict on the C2 shapes under the current SE_OZ_* environment (library-internal event timing)."""
import sys, os, ctypes as C
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from safe_exploration_b200 import _lib, workloads
from safe_exploration_b200.ssm import B200GpCemSSM
lib = _lib.load()
cfg = sys.argv[1] if len(sys.argv) > 1 else 'C2'
digits = int(sys.argv[2]) if len(sys.argv) > 2 else 7
N = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
w = workloads.pendulum_workload() if cfg == 'C2' else workloads.cartpole_workload()
ssm = B200GpCemSSM(None, w.n_s, w.n_u, lengthscale=w.lengthscale, outputscale=w.outputscale, noise=w.noise, gemm_impl=_lib.SE_GEMM_TC)
ssm.update_model(torch.as_tensor(w.x_train, device='cuda'), torch.as_tensor(w.y_train, device='cuda'), replace_old=True)
_lib.check(lib.se_gp_set_tc_digits(ssm.handle, digits))
z = (torch.rand(N, w.n_s + w.n_u, dtype=torch.float64, device='cuda') - 0.5) * 0.4
for _ in range(3):
    ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
_lib.check(lib.se_gp_timing_enable(ssm.handle, 1))
for _ in range(10):
    ssm.predict_with_jacobians(z[:, :w.n_s], z[:, w.n_s:])
g, k, gn, kn = C.c_double(0), C.c_double(0), C.c_int64(0), C.c_int64(0)
_lib.check(lib.se_gp_timing_read(ssm.handle, C.byref(g), C.byref(gn), C.byref(k), C.byref(kn), _lib.stream_ptr('cuda:0')))
print(f'{cfg} digits={digits} N={N} env={ {k_: v for k_, v in os.environ.items() if k_.startswith("SE_OZ")} }: gemm {g.value/gn.value:.3f} ms, kstar {k.value/kn.value:.3f} ms', flush=True)
