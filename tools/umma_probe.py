"""Explore tcgen05 descriptor hypotheses on a B200 (run under gpurun).  Prints the max error of each candidate."""
import ctypes as C
import itertools
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mappo_b200 import _lib

L = _lib.lib()
L.mappo_selftest_umma.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_int] + [C.c_uint32] * 6 + [C.c_void_p, C.c_void_p]
L.mappo_selftest_umma.restype = C.c_int
rng = np.random.RandomState(0)


def run(A, B, N, K, a_mn, b_mn, a, b):
    At = torch.as_tensor(A, dtype=torch.float32).cuda().contiguous()
    Bt = torch.as_tensor(B, dtype=torch.float32).cuda().contiguous()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.mappo_selftest_umma(At.data_ptr(), A.shape[0], A.shape[1], Bt.data_ptr(), B.shape[0], B.shape[1], 128, N, K,
                               a_mn, b_mn, a[0], a[1], a[2], b[0], b[1], b[2], D.data_ptr(), None)
    assert rc == 0, L.mappo_last_error()
    torch.cuda.synchronize()
    return D.cpu().numpy()


def cands(R, kmajor):
    big, small = R * 16, 128
    if kmajor:   # (lbo, sbo, kstep)
        return {"H1 lbo=R16 sbo=128": (big, small, 2 * big)}
    return {"H1 sbo=R16 lbo=128": (small, big, small)}


ints = lambda *s: rng.randint(-4, 5, size=s).astype(np.float32)
# mode 0: K-major x K-major
A, B = ints(128, 64), ints(64, 64)
for (na, a), (nb, b) in itertools.product(cands(128, True).items(), cands(64, True).items()):
    D = run(A, B, 64, 64, 0, 0, a, b)
    print("mode0 KK  A:%s B:%s  maxerr %.3g" % (na, nb, np.nanmax(np.abs(D - A @ B.T)) if not np.isnan(D).all() else -1))
# mode 1: MN-major x MN-major
A, B = ints(128, 128), ints(128, 64)
for (na, a), (nb, b) in itertools.product(cands(128, False).items(), cands(128, False).items()):
    D = run(A, B, 64, 128, 1, 1, a, b)
    print("mode1 MM  A:%s B:%s  maxerr %.3g" % (na, nb, np.nanmax(np.abs(D - A.T @ B))))
# mode 2: K-major A x MN-major B
A, B = ints(128, 64), ints(64, 64)
for (na, a), (nb, b) in itertools.product(cands(128, True).items(), cands(64, False).items()):
    D = run(A, B, 64, 64, 0, 1, a, b)
    print("mode2 KM  A:%s B:%s  maxerr %.3g" % (na, nb, np.nanmax(np.abs(D - A @ B))))
# precision: 3xTF32 split vs fp32 and single tf32
A, B = rng.randn(128, 64).astype(np.float32), rng.randn(64, 64).astype(np.float32)
D = run(A, B, 64, 64, 0, 0, cands(128, True)["H1 lbo=R16 sbo=128"], cands(64, True)["H1 lbo=R16 sbo=128"])
ref = A.astype(np.float64) @ B.astype(np.float64).T
print("raw fp32 inputs through kind::tf32: max rel err %.3g (vs |ref| max)" % (np.abs(D - ref).max() / np.abs(ref).max()))
