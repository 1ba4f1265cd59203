"""Measure how far the float32 GP variance / one-step ellipsoids are from the float64 ones on the bench workload, for
several noise levels (conditioning of K + noise I).  Evidence for the precision choice in DESIGN.md."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import safe_exploration_b200 as se
from safe_exploration_b200 import workloads

dev = 'cuda:0'
out = []
for n_train in (2000,):
    for noise in (1e-5, 1e-4, 1e-3):
        w = workloads.pendulum_workload(n_train=n_train, num_rollouts=4096)
        w.noise = np.full(2, noise)
        rng = np.random.default_rng(0)
        z = torch.as_tensor(rng.uniform([-1.0, -0.3, -1.0], [1.0, 0.3, 1.0], (4096, 3)), device=dev)
        res = {}
        for dt in (torch.float64, torch.float32):
            gp = se.B200GpCemSSM(None, 2, 1, w.lengthscale, w.outputscale, w.noise, dtype=dt, device=dev)
            gp.update_model(torch.as_tensor(w.x_train, device=dev), torch.as_tensor(w.y_train, device=dev), replace_old=True)
            m, v, j = gp.predict_with_jacobians(z[:, :2].to(dt).contiguous(), z[:, 2:].to(dt).contiguous())
            res[dt] = (m.double().cpu().numpy(), v.double().cpu().numpy(), j.double().cpu().numpy())
        m64, v64, j64 = res[torch.float64]
        m32, v32, j32 = res[torch.float32]
        prior = (w.outputscale + w.noise)[None, :]
        out.append(dict(n_train=n_train, noise=noise, outputscale=w.outputscale.tolist(),
                        var_min=float(v64.min()), var_median=float(np.median(v64)), var_max=float(v64.max()),
                        var_abs_err_max=float(np.abs(v32 - v64).max()), var_abs_err_median=float(np.median(np.abs(v32 - v64))),
                        var_rel_err_max=float((np.abs(v32 - v64) / v64).max()),
                        var_rel_err_median=float(np.median(np.abs(v32 - v64) / v64)),
                        var_err_rel_prior_max=float((np.abs(v32 - v64) / prior).max()),
                        mean_rel_err_max=float(np.abs(m32 - m64).max() / np.abs(m64).max()),
                        jac_rel_err_max=float(np.abs(j32 - j64).max() / np.abs(j64).max())))
        print(json.dumps(out[-1]), flush=True)
