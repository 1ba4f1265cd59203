"""GPU check of the tcgen05 kind::i8 tile primitive (se_debug_tc_probe) against integer matmuls; tries both stride
conventions of the no-swizzle K-major UMMA descriptor."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from safe_exploration_b200 import _lib


def image(planes):  # planes [nk][S][rows][32] int8 -> bytes in core-matrix order [row/8][k/16][row%8][k%16]
    nk, S, R, K = planes.shape
    return np.ascontiguousarray(planes.reshape(nk, S, R // 8, 8, 2, 16).transpose(0, 1, 2, 4, 3, 5)).reshape(-1)


def main():
    lib = _lib.load()
    rng = np.random.default_rng(0)
    ok_all = True
    for S in (8, 7, 6, 5):
        for nk in (1, 3):
            a = rng.integers(-128, 128, size=(nk, S, 128, 32), dtype=np.int8)
            b = rng.integers(-128, 128, size=(nk, S, 64, 32), dtype=np.int8)
            want = np.zeros((S, 128, 64), dtype=np.int64)
            for t in range(S):
                for u in range(S - t):
                    for k in range(nk):
                        want[t + u] += a[k, t].astype(np.int64) @ b[k, u].astype(np.int64).T
            da = torch.from_numpy(image(a)).cuda()
            db = torch.from_numpy(image(b)).cuda()
            for lbo, sbo in ((128, 256), (256, 128)):
                out = torch.full((S, 128, 64), -1, dtype=torch.int32, device='cuda')
                _lib.check(lib.se_debug_tc_probe(_lib.ptr(da), _lib.ptr(db), nk, S, lbo, sbo, _lib.ptr(out),
                                                 _lib.stream_ptr('cuda:0')))
                torch.cuda.synchronize()
                got = out.cpu().numpy().astype(np.int64)
                ok = np.array_equal(got, want)
                print(f'S={S} nk={nk} lbo={lbo} sbo={sbo}: {"MATCH" if ok else "mismatch"}'
                      f' (equal entries {np.mean(got == want):.3f}; got[0,0,:4]={got[0,0,:4]} want={want[0,0,:4]})', flush=True)
                if (lbo, sbo) == (128, 256):
                    ok_all &= ok
    print('PROBE', 'OK' if ok_all else 'FAILED')


main()
