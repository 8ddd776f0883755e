"""Recover the shared-memory address map of an MN-major tcgen05 operand: fill the A region with its own word
index, multiply by an identity B (known-good K-major) and read which word each (m, k) came from."""
import ctypes as C
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mappo_b200 import _lib

L = _lib.lib()
L.mappo_selftest_umma.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_int, C.c_int] + [C.c_uint32] * 6 + [C.c_void_p, C.c_void_p]
L.mappo_selftest_umma.restype = C.c_int


def tile_off(R, r, c):
    return ((c >> 2) * R + r) * 16 + (c & 3) * 4


def run(A, B, N, K, a_mn, b_mn, a, b):
    At = torch.as_tensor(A, dtype=torch.float32).cuda().contiguous()
    Bt = torch.as_tensor(B, dtype=torch.float32).cuda().contiguous()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.mappo_selftest_umma(At.data_ptr(), A.shape[0], A.shape[1], Bt.data_ptr(), B.shape[0], B.shape[1], 128, N, K,
                               a_mn, b_mn, a[0], a[1], a[2], b[0], b[1], b[2], D.data_ptr(), None)
    assert rc == 0, L.mappo_last_error()
    torch.cuda.synchronize()
    return D.cpu().numpy()


R, Ccols = 128, 128
w = np.zeros((R, Ccols), np.int64)
for r in range(R):
    for c in range(Ccols):
        w[r, c] = tile_off(R, r, c) // 4
B = np.zeros((16, 8), np.float32)
B[:8, :8] = np.eye(8)
bdesc = (16 * 16, 128, 0)
for (lbo, sbo) in ((128, 2048), (2048, 128), (4096, 256)):
    lo = run((w % 1024).astype(np.float32), B, 16, 8, 1, 0, (lbo, sbo, 0), bdesc)
    hi = run((w // 1024).astype(np.float32), B, 16, 8, 1, 0, (lbo, sbo, 0), bdesc)
    word = (hi * 1024 + lo).astype(np.int64)[:, :8]
    print("A MN-major, LBO=%d SBO=%d: byte offset read for (m, k):" % (lbo, sbo))
    for m in list(range(0, 10)) + [16, 32, 33, 64, 127]:
        print("  m=%3d: %s" % (m, " ".join("%6d" % (4 * x) for x in word[m])))

print("---- diagnostics")
ones = np.ones((R, Ccols), np.float32)
Bones = np.ones((16, 8), np.float32)
d = run(ones, B, 16, 8, 1, 0, (128, 2048, 0), bdesc); print("A(MN)=1, B=eye:", d[0, :10], d[77, :10])
d = run(ones, Bones, 16, 8, 1, 0, (128, 2048, 0), bdesc); print("A(MN)=1, B=1:", d[0, :10], d[77, :10])
d = run((w % 1024).astype(np.float32), Bones, 16, 8, 1, 0, (128, 2048, 0), bdesc); print("A(MN)=w, B=1:", d[0, :4], d[1, :4], d[4,:4], d[77, :4])
Ak = np.zeros((128, 8), np.float32)
for r in range(128):
    for c in range(8):
        Ak[r, c] = (tile_off(128, r, c) // 4) % 1024
d = run(Ak, B, 16, 8, 0, 0, (2048, 128, 0), bdesc); print("A(K)=w, B=eye: m=0", d[0, :10], "m=1", d[1, :10], "m=9", d[9, :10])
d = run(ones[:, :8].copy(), B, 16, 8, 0, 0, (2048, 128, 0), bdesc); print("A(K)=1, B=eye:", d[0, :10])
