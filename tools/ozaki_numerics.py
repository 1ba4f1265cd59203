"""CPU emulation of int8-slice (Ozaki scheme I) evaluation of the variance GEMM v = L^-1 k*, on the C2 workload.
Reports the error of var = s + noise - ||v||^2 against fp64 for several slice counts / level cut-offs."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from safe_exploration_b200 import workloads
import scipy.linalg

def slices(a, nsl, bits=7):
    """row-scaled signed-digit decomposition: a ~= 2^e * sum_t d_t 2^{-(bits-1) - bits*t}, |d_t| <= 2^(bits-1)"""
    amax = np.abs(a).max(axis=1, keepdims=True)
    e = np.ceil(np.log2(np.where(amax > 0, amax, 1.0)))
    r = a / 2.0 ** e  # in [-1, 1]
    ds = []
    scale = 2.0 ** (bits - 1)
    for t in range(nsl):
        r = r * scale
        d = np.rint(r)
        r = r - d
        ds.append(d)
        scale = 2.0 ** bits
    return ds, e

def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    nc = 256
    w = workloads.pendulum_workload(n_train=n)
    rng = np.random.default_rng(0)
    x = w.x_train
    o = 0
    ls, s, nz = w.lengthscale[o], w.outputscale[o], w.noise[o]
    xs = x / ls
    d2 = ((xs[:, None, :] - xs[None, :, :]) ** 2).sum(-1)
    K = s * np.exp(-0.5 * d2) + nz * np.eye(n)
    L = np.linalg.cholesky(K)
    W = scipy.linalg.solve_triangular(L, np.eye(n), lower=True)
    W = np.tril(W)
    # candidates: half near the origin (typical rollout states), half anywhere in the data box
    z = np.concatenate([rng.normal(0, 0.1, size=(nc // 2, 3)), rng.uniform(-1, 1, size=(nc // 2, 3)) * np.array([1.2, .4, 1.])])
    zs = z / ls
    ks = s * np.exp(-0.5 * ((zs[:, None, :] - xs[None, :, :]) ** 2).sum(-1))  # [nc, n]
    v = ks @ W.T
    q = (v * v).sum(1)
    var = s + nz - q
    # long double reference on a few rows
    vl = (ks[:16].astype(np.longdouble)) @ W.T.astype(np.longdouble)
    ql = (vl * vl).sum(1)
    print('fp64 vs longdouble: rel err of var', np.abs((s + nz - ql) - var[:16]).max() / var[:16].min(), 'var range', var.min(), var.max(), 's/noise', s / nz)
    print('row scales W: max|W| per row range', np.abs(W).max(1).min(), np.abs(W).max(1).max(), ' ||W_i||_1 max', np.abs(W).sum(1).max())
    for bits in (7,):
        for S in (4, 5, 6, 7, 8):
            da, ea = slices(ks, S, bits)
            db, eb = slices(W, S, bits)
            for L_cut in sorted({S - 1, S, S + 1}):
                acc = np.zeros((nc, n))
                npairs = 0
                for t in range(S):
                    for u in range(S):
                        if t + u <= L_cut - 1:
                            acc += (da[t] @ db[u].T) * 2.0 ** (-2 * (bits - 1) - bits * (t + u))
                            npairs += 1
                vo = acc * 2.0 ** ea * 2.0 ** eb.T
                qo = (vo * vo).sum(1)
                varo = s + nz - qo
                rel = np.abs(varo - var) / var
                print(f'bits {bits} slices {S} levels {L_cut} pairs {npairs}: var rel err max {rel.max():.3e} median {np.median(rel):.3e}; '
                      f'|dq|/s max {np.abs(qo - q).max() / s:.3e}')

main()
