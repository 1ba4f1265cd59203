"""Times the elite selection + refit kernel (se_cem_select_refit) alone at the global population sizes an 8-rank
weak-scaling solve reaches (16384 per rank): what every rank runs on the all-gathered costs once per CEM iteration."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from safe_exploration_b200 import cem

dev = 'cuda:0'
T, n_u = 10, 1
for N, E in ((4096, 64), (16384, 256), (32768, 512), (65536, 1024), (131072, 2048), (262144, 4096)):
    g = torch.Generator(device=dev).manual_seed(N)
    obj = torch.randn(N, dtype=torch.float64, device=dev, generator=g)
    for frac_feasible in (0.9, 0.001):
        con = torch.where(torch.rand(N, device=dev, generator=g) < frac_feasible, 0.0, 10.0).to(torch.float64)
        act = torch.randn(N, T, n_u, dtype=torch.float64, device=dev, generator=g)
        for _ in range(3):
            cem.select_elites_refit(obj, con, act, E)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            cem.select_elites_refit(obj, con, act, E)
        e1.record()
        torch.cuda.synchronize()
        print(f'select_refit N={N} E={E} feasible={frac_feasible}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per call '
              f'(includes the host wrapper: 3 small allocations)', flush=True)
