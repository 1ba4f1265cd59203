"""How far apart are the means of predict_with_jacobians and predict_without_jacobians on the golden 'trained' model, per factor engine and variance engine?"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_exploration_b200 as se
from safe_exploration_b200 import _lib
g = dict(np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'onestep_invpend.npz')))
G = lambda a: torch.as_tensor(np.asarray(a, dtype=np.float64)).to('cuda:0')
for fi in ('native', 'cusolver'):
    for impl in (_lib.SE_GEMM_SIMT, _lib.SE_GEMM_MMA):
        gp = se.B200GpCemSSM(None, 2, 1, g['trained_lengthscale'], g['trained_outputscale'], g['trained_noise'], device='cuda:0', gemm_impl=impl, factor_impl=fi)
        gp.update_model(G(g['x_train']), G(g['y_train']), replace_old=True)
        mean, var, jac = gp.predict_with_jacobians(G(g['p']), G(g['u']))
        m2, v2 = gp.predict_without_jacobians(G(g['p']), G(g['u']))
        print(fi, impl, 'max rel diff', float(((m2 - mean).abs() / mean.abs()).max()), 'vs golden', float(np.abs(mean.cpu().numpy() - g['trained_mean']).max() / np.abs(g['trained_mean']).max()))
