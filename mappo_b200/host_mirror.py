"""Host <-> device twins for the reference-facing (host array) boundary.

The reference's runner moves every step's tensors device -> host (`_t2n`) and host -> device again
(`SharedReplayBuffer.insert`, onpolicy/runner/shared/mpe_runner.py:95-139).  When a host array handed to this library
is one the library itself just produced from a device tensor -- the pinned arrays `MPEVecEnv.step` returns in host_io
mode, or arrays made by `t2n` below -- the device copy still exists, and sending the bytes back over PCIe is redundant.
This registry remembers, per host buffer (by address and size), the device tensor it was copied from; `lookup` gives
`insert` / `step` the twin so they copy device -> device instead.

PASS-THROUGH SEMANTICS: the twin is used only if the array handed in is the registered
buffer itself (same address, same byte count, same serial).  A different host array -- including a `.copy()` -- takes
the normal host -> device path.  Code that MODIFIES a registered array in place between receiving it and handing it
back must call `forget(array)` (or set MAPPO_HOST_PASSTHROUGH=0): the device twin would not see the edit."""
import os

import numpy as np
import torch

ENABLED = os.environ.get("MAPPO_HOST_PASSTHROUGH", "1") != "0"
_TWINS = {}          # data pointer -> (nbytes, device tensor)
_PINNED = {}         # (tag, shape, dtype) -> pinned host tensor
_DEVICE_SCRATCH = {}  # (tag, shape) -> device uint8 tensor (one-hot built on the device)


def _key(arr):
    if isinstance(arr, np.ndarray):
        return arr.__array_interface__["data"][0], arr.nbytes
    if torch.is_tensor(arr) and not arr.is_cuda:
        return arr.data_ptr(), arr.numel() * arr.element_size()
    return None, 0


def register(host, device_tensor):
    """host: numpy array or CPU tensor (contiguous); device_tensor: the CUDA tensor holding the same values."""
    ptr, nbytes = _key(host)
    if ptr is not None:
        _TWINS[ptr] = (nbytes, device_tensor)


def forget(host):
    ptr, _ = _key(host)
    _TWINS.pop(ptr, None)


def lookup(host):
    """The device twin of `host`, or None."""
    if not ENABLED:
        return None
    ptr, nbytes = _key(host)
    if ptr is None:
        return None
    hit = _TWINS.get(ptr)
    if hit is None or hit[0] != nbytes:
        return None
    if isinstance(host, np.ndarray) and not host.flags["C_CONTIGUOUS"]:
        return None
    return hit[1]


def pinned(tag, shape, dtype):
    k = (tag, tuple(shape), dtype)
    t = _PINNED.get(k)
    if t is None:
        t = _PINNED[k] = torch.empty(*shape, dtype=dtype).pin_memory()
    return t


def t2n_many(named):
    """Several `_t2n` calls under ONE stream synchronise: named = {tag: cuda tensor} -> {tag: numpy array}."""
    out, dev = {}, None
    for tag, x in named.items():
        x = x.detach()
        h = pinned(tag, x.shape, x.dtype)
        h.copy_(x, non_blocking=True)
        register(h, x)
        out[tag], dev = h.numpy(), x.device
    if dev is not None:
        torch.cuda.current_stream(dev).synchronize()
    return out


def t2n(x, tag="t2n"):
    """Drop-in for the reference's `_t2n(x)` = x.detach().cpu().numpy() (runner/shared/mpe_runner.py:8-9): one
    asynchronous copy into a reused PINNED buffer + one stream synchronise, and the result is registered as the twin
    of `x`, so that handing it back (buffer.insert) costs a device -> device copy.  `tag` separates arrays of equal
    shape that are alive at the same time (values / log-probs)."""
    x = x.detach()
    if not x.is_cuda:
        return x.numpy()
    h = pinned(tag, x.shape, x.dtype)
    h.copy_(x, non_blocking=True)
    torch.cuda.current_stream(x.device).synchronize()
    register(h, x)
    return h.numpy()


def onehot(actions, head_dims, tag="onehot"):
    """Drop-in for the runner's `np.eye(n)[actions]` (runner/shared/mpe_runner.py:110-121; MultiDiscrete: heads
    concatenated): int64 ids [..., heads] on the host -> uint8 one-hot [..., sum(head_dims)] in a reused pinned array
    (one C pass, mappo_host_onehot_u8).  `MPEVecEnv.step` takes it as is (1 byte per logit over PCIe)."""
    import ctypes as C
    from . import _lib
    heads = len(head_dims)
    twin = lookup(actions)
    if twin is not None and twin.dtype in (torch.float32, torch.int64, torch.int32) and twin.shape[-1] == heads:
        # the ids are a host copy this library made of a device tensor: build the one-hot next to them (one small
        # kernel), bring it out as bytes, and remember the device copy -- `MPEVecEnv.step` then needs no upload at all
        from . import gather
        shape = tuple(twin.shape[:-1]) + (int(sum(head_dims)),)
        dev = _DEVICE_SCRATCH.get((tag, shape))
        if dev is None:
            dev = _DEVICE_SCRATCH[(tag, shape)] = torch.empty(shape, dtype=torch.uint8, device=twin.device)
        ids = twin.to(torch.float32).contiguous()        # (the policy returns int64 ids, as torch's Categorical does)
        bad = _DEVICE_SCRATCH.get((tag, "bad"))
        if bad is None:
            bad = _DEVICE_SCRATCH[(tag, "bad")] = torch.zeros(1, dtype=torch.int32, device=twin.device)
        gather.ids_to_onehot(ids.view(-1, heads), head_dims, dev.view(-1, shape[-1]), bad)
        out, flag = pinned(tag, shape, torch.uint8), pinned(tag + ".bad", (1,), torch.int32)
        out.copy_(dev, non_blocking=True)
        flag.copy_(bad, non_blocking=True)
        torch.cuda.current_stream(twin.device).synchronize()
        if int(flag[0]) != 0:                        # the reference's np.eye(n)[actions] raises IndexError here
            bad.zero_()
            forget(out)
            raise IndexError("host_mirror.onehot: an action id lies outside its head %r" % (list(head_dims),))
        register(out, dev)
        return out.numpy()
    a = np.asarray(actions)
    f32 = a.dtype == np.float32                  # what the policy returns (act.py stores ids as floats): no int64 pass
    a = np.ascontiguousarray(a, dtype=np.float32 if f32 else np.int64)
    rows = a.size // heads
    out = pinned(tag, tuple(a.shape[:-1]) + (int(sum(head_dims)),), torch.uint8)
    forget(out)                                  # (a device one-hot registered for this buffer earlier is stale now)
    L = _lib.lib()
    fn = L.mappo_host_onehot_u8_f32 if f32 else L.mappo_host_onehot_u8
    fn.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.POINTER(C.c_int), C.c_void_p]
    hd = (C.c_int * heads)(*[int(d) for d in head_dims])
    _lib.check(fn(C.c_void_p(a.ctypes.data), rows, heads, hd, C.c_void_p(out.data_ptr())), "host_onehot_u8")
    return out.numpy()
