"""bf16 (kind::f16) operand-layout probe on a B200: K-major / MN-major, no swizzle, chunk-major tiles."""
import ctypes as C
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mappo_b200 import _lib

L = _lib.lib()
L.mappo_selftest_umma.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_int] + [C.c_uint32] * 6 + [C.c_void_p, C.c_int, C.c_void_p]
L.mappo_selftest_umma.restype = C.c_int
rng = np.random.RandomState(0)


def run(A, B, N, K, a_mn, b_mn, a, b):
    At = torch.as_tensor(A, dtype=torch.float32).cuda().contiguous()
    Bt = torch.as_tensor(B, dtype=torch.float32).cuda().contiguous()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.mappo_selftest_umma(At.data_ptr(), A.shape[0], A.shape[1], Bt.data_ptr(), B.shape[0], B.shape[1], 128, N, K,
                               a_mn, b_mn, a[0], a[1], a[2], b[0], b[1], b[2], D.data_ptr(), 1, None)
    assert rc == 0, L.mappo_last_error()
    torch.cuda.synchronize()
    return D.cpu().numpy()


kmaj = lambda R: (R * 16, 128, 2 * R * 16)     # (lbo, sbo, kstep) K-major, K = 16 per instruction
mnmaj = lambda R: (128, R * 16, 256)           # MN-major: sbo = MN-group stride, lbo = K-group stride
ints = lambda *s: rng.randint(-4, 5, size=s).astype(np.float32)

A, B = ints(128, 64), ints(64, 64)
D = run(A, B, 64, 64, 0, 0, kmaj(128), kmaj(64)); print("KK  maxerr", np.abs(D - A @ B.T).max())
A, B = ints(128, 128), ints(128, 64)
D = run(A, B, 64, 128, 1, 1, mnmaj(128), mnmaj(128)); print("MM  maxerr", np.abs(D - A.T @ B).max())
A, B = ints(128, 64), ints(64, 64)
D = run(A, B, 64, 64, 0, 1, kmaj(128), mnmaj(64)); print("KM  maxerr", np.abs(D - A @ B).max())
A, B = ints(128, 128), ints(128, 128)
D = run(A, B, 128, 128, 1, 1, mnmaj(128), mnmaj(128)); print("MM N=128 maxerr", np.abs(D - A.T @ B).max())
# alternative hypothesis for MN-major if the above fails: swap lbo/sbo
A, B = ints(128, 64), ints(64, 64)
D = run(A, B, 64, 64, 0, 1, kmaj(128), (64 * 16, 128, 256)); print("KM swapped maxerr", np.abs(D - A @ B).max())
