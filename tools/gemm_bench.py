"""Time the variance GEMM alone (se_gp_predict on one candidate batch) for the DMMA variants / orderings selected by the
SE_DMMA_* environment variables.  `python tools/gemm_bench.py sweep [n_train] [n_s]` runs every combination in a
subprocess (the library reads the variables once per process)."""
import ctypes as C
import json
import os
import subprocess
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(n_train=2000, n_cand=16384, n_s=2, reps=5, ref_path=None):
    import numpy as np
    import torch
    import safe_exploration_b200 as se
    from safe_exploration_b200 import _lib, workloads
    dev = 'cuda:0'
    if n_s == 2:
        w = workloads.pendulum_workload(n_train=n_train, num_rollouts=64)
    else:
        w = workloads.cartpole_workload(n_train=n_train, num_rollouts=64)
    gp = se.B200GpCemSSM(None, w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise, device=dev)
    gp.update_model(torch.as_tensor(w.x_train, device=dev), torch.as_tensor(w.y_train, device=dev), replace_old=True)
    rng = np.random.default_rng(0)
    lo, hi = w.x_train.min(axis=0), w.x_train.max(axis=0)
    z = torch.as_tensor(rng.uniform(lo, hi, (n_cand, w.n_s + w.n_u)), device=dev)
    st, ac = z[:, :w.n_s].contiguous(), z[:, w.n_s:].contiguous()
    lib = se.load_library()
    for _ in range(2):
        m, v, j = gp.predict_with_jacobians(st, ac)
    torch.cuda.synchronize()
    _lib.check(lib.se_gp_timing_enable(gp.handle, 1))
    for _ in range(reps):
        m, v, j = gp.predict_with_jacobians(st, ac)
    g_ms, k_ms, g_n, k_n = C.c_double(0), C.c_double(0), C.c_int64(0), C.c_int64(0)
    _lib.check(lib.se_gp_timing_read(gp.handle, C.byref(g_ms), C.byref(g_n), C.byref(k_ms), C.byref(k_n),
                                     _lib.stream_ptr(dev)))
    gemm = g_ms.value / g_n.value
    flop = w.n_s * float(n_cand) * float(n_train) ** 2
    out = dict(variant=os.environ.get('SE_DMMA_VARIANT', 'default'), skip=os.environ.get('SE_DMMA_SKIP', 'default'),
               l2mb=os.environ.get('SE_DMMA_L2MB', 'default'), n_train=n_train, n_cand=n_cand, n_s=w.n_s,
               gemm_ms=round(gemm, 4), kstar_ms=round(k_ms.value / k_n.value, 4), tflops=round(flop / gemm / 1e9, 3))
    if ref_path:
        if os.path.exists(ref_path):
            ref = torch.load(ref_path)
            out['max_rel_diff_vs_ref'] = float(((v.cpu() - ref).abs() / ref.abs()).max())
        else:
            torch.save(v.cpu(), ref_path)
    print(json.dumps(out), flush=True)


if __name__ == '__main__':
    mode = sys.argv[1] if len(sys.argv) > 1 else 'one'
    n_train = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    n_s = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    if mode == 'sweep':
        ref = f'/tmp/gemm_ref_{n_train}_{n_s}.pt'
        combos = [(0, 0, 0)] + [(v, s, 0) for v in (2, 5, 6) for s in (0, 1)]
        for v, s, l in combos:
            env = dict(os.environ, SE_DMMA_VARIANT=str(v), SE_DMMA_SKIP=str(s), SE_DMMA_L2MB=str(l))
            subprocess.run([sys.executable, os.path.abspath(__file__), 'one', str(n_train), str(n_s), ref], env=env)
    else:
        one(n_train=n_train, n_s=n_s, n_cand=16384 if n_s == 2 else 8192,
            ref_path=sys.argv[4] if len(sys.argv) > 4 else None)
