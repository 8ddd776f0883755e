"""GPU: the allocation-free entry point mappo_moments_static (static __device__ scratch inside the library; no
counterpart in the reference) against the existing pair
mappo_moments + mappo_adv_stats_reduce, through the C ABI, on two calls in a row.  Collected last on purpose.
Passes alone in a fresh process and after the other tests: it shares no state with them
except the library's static array, which every call overwrites before reading."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind():
    from mappo_b200 import _lib
    L = _lib.lib()
    L.mappo_moments_static.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    L.mappo_moments_static.restype = C.c_int
    L.mappo_moments_static_max_count.restype = C.c_longlong
    L.mappo_moments_num_partials.argtypes = [C.c_longlong]
    L.mappo_moments.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    L.mappo_adv_stats_reduce.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    return _lib, L


def _existing(_lib, L, x):
    n = x.numel()
    nparts = L.mappo_moments_num_partials(n)
    partials = torch.empty(nparts, 3, dtype=torch.float64, device=x.device)
    out = torch.empty(3, dtype=torch.float64, device=x.device)
    _lib.check(L.mappo_moments(_lib.ptr(x), n, _lib.ptr(partials), _lib.stream_ptr()), "moments")
    _lib.check(L.mappo_adv_stats_reduce(_lib.ptr(partials), nparts, _lib.ptr(out), _lib.stream_ptr()), "stats_reduce")
    return out


def test_static_scratch_moments_equal_the_existing_entry_point_twice_in_a_row():
    _lib, L = _bind()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(7)
    # two different sizes back to back: the second call uses fewer partial rows than the first left behind
    sizes = (4_915_200, 70_001)
    outs = [torch.full((3,), float("nan"), dtype=torch.float64, device=dev) for _ in sizes]     # allocated up front
    xs = [(-5.0 + 3.0 * torch.randn(n, device=dev, generator=g)).contiguous() for n in sizes]
    wants = [_existing(_lib, L, x) for x in xs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    for x, out in zip(xs, outs):                         # the two calls in a row, same stream
        _lib.check(L.mappo_moments_static(_lib.ptr(x), x.numel(), _lib.ptr(out), _lib.stream_ptr()), "moments_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before    # nothing was allocated through torch on the way
    for x, out, want in zip(xs, outs, wants):
        # same kernels, same grid, same fixed-order reduction: bit-identical (the GAE / ValueNorm tests' bar)
        assert (out == want).all(), (out.tolist(), want.tolist())
        x64 = x.double().cpu().numpy()
        np.testing.assert_allclose(out.cpu().numpy(),
                                   [x64.sum(), (x.cpu().numpy() ** 2).astype(np.float64).sum(), x.numel()], rtol=1e-9)
    # a third call on the first input again: the stale rows of the larger first call do not leak into a smaller one
    again = torch.empty(3, dtype=torch.float64, device=dev)
    _lib.check(L.mappo_moments_static(_lib.ptr(xs[1]), xs[1].numel(), _lib.ptr(again), _lib.stream_ptr()), "moments_static")
    assert (again == wants[1]).all()


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    _lib, L = _bind()
    mx = int(L.mappo_moments_static_max_count())
    assert mx == 1 << 28
    x = torch.zeros(16, device="cuda:0")
    out = torch.zeros(3, dtype=torch.float64, device="cuda:0")
    rc = L.mappo_moments_static(_lib.ptr(x), mx + 1, _lib.ptr(out), _lib.stream_ptr())    # rejected before any launch
    assert rc == -1
    msg = L.mappo_last_error().decode()
    assert "exceeds the maximum" in msg and str(mx) in msg
    assert L.mappo_moments_static(None, 16, _lib.ptr(out), _lib.stream_ptr()) == -1
