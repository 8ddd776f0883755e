"""GPU: the tcgen05 operand layouts the tensor-core policy kernels rely on (mappo_b200/csrc/umma.cuh), checked on
the hardware against numpy: bf16 tiles in the chunk-major no-swizzle layout read as K-major and as MN-major operands
(the same bytes serve both), M = 128, fp32 accumulation in TMEM.  Integer-valued inputs make every product exact."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(A, B, N, K, a_mn, b_mn, a, b, bf16=1):
    from mappo_b200 import _lib
    L = _lib.lib()
    L.mappo_selftest_umma.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.c_int] + [C.c_uint32] * 6 + [C.c_void_p, C.c_int, C.c_void_p]
    L.mappo_selftest_umma.restype = C.c_int
    At = torch.as_tensor(A, dtype=torch.float32).cuda().contiguous()
    Bt = torch.as_tensor(B, dtype=torch.float32).cuda().contiguous()
    D = torch.full((128, N), float("nan"), device="cuda")
    _lib.check(L.mappo_selftest_umma(At.data_ptr(), A.shape[0], A.shape[1], Bt.data_ptr(), B.shape[0], B.shape[1], 128,
                                     N, K, a_mn, b_mn, a[0], a[1], a[2], b[0], b[1], b[2], D.data_ptr(), bf16, None),
               "selftest_umma")
    torch.cuda.synchronize()
    return D.cpu().numpy()


kmaj = lambda R: (R * 16, 128, 2 * R * 16)      # (LBO, SBO, start advance per K = 16) of a K-major [R][C] tile
mnmaj = lambda R: (128, R * 16, 256)            # the same tile read as an MN-major operand (K = rows)


def test_bf16_operand_layouts():
    rng = np.random.RandomState(0)
    ints = lambda *s: rng.randint(-4, 5, size=s).astype(np.float32)
    A, B = ints(128, 64), ints(64, 64)           # forward: activations [rows][k] x weights [j][k]
    np.testing.assert_array_equal(_run(A, B, 64, 64, 0, 0, kmaj(128), kmaj(64)), A @ B.T)
    np.testing.assert_array_equal(_run(A, B, 64, 64, 0, 1, kmaj(128), mnmaj(64)), A @ B)     # dgrad: same W tile
    A, B = ints(128, 128), ints(128, 128)        # wgrad: both operands reduce over the tile rows
    np.testing.assert_array_equal(_run(A, B, 128, 128, 1, 1, mnmaj(128), mnmaj(128)), A.T @ B)
    np.testing.assert_array_equal(_run(A, B[:, :32].copy(), 32, 128, 1, 1, mnmaj(128), mnmaj(128)), A.T @ B[:, :32])


def test_tf32_k_major():
    rng = np.random.RandomState(1)
    A = rng.randint(-4, 5, size=(128, 64)).astype(np.float32)
    B = rng.randint(-4, 5, size=(64, 64)).astype(np.float32)
    got = _run(A, B, 64, 64, 0, 0, (128 * 16, 128, 2 * 128 * 16), (64 * 16, 128, 2 * 64 * 16), bf16=0)
    np.testing.assert_array_equal(got, A @ B.T)
