"""ctypes binding of the C-ABI library (include/mappo_b200.h).

The product has no CPU fallback: if the library is missing or a call fails, this raises."""
import ctypes as C
import os

import torch  # noqa: F401  (loads the CUDA runtime the library links against)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmappo_b200.so")

_lib = None


class MappoError(RuntimeError):
    pass


class EnvConfig(C.Structure):
    _fields_ = [("scenario", C.c_int32), ("num_envs", C.c_int32), ("num_agents", C.c_int32),
                ("num_landmarks", C.c_int32), ("episode_length", C.c_int32), ("num_adversaries", C.c_int32),
                ("seed", C.c_uint64), ("env_id_offset", C.c_int64)]


class EnvState(C.Structure):
    _fields_ = [("p_pos", C.c_void_p), ("p_vel", C.c_void_p), ("lm_pos", C.c_void_p), ("comm", C.c_void_p),
                ("goal", C.c_void_p), ("step_count", C.c_void_p), ("episode", C.c_void_p)]


def lib():
    """Load (once) and return the library; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MappoError(
                "mappo_b200: %s not found. Build it with `python -m mappo_b200.build` "
                "(there is no CPU or PyTorch fallback for the hot path)." % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
        if _lib.mappo_abi_version() != 6:
            raise MappoError("mappo_b200: ABI version mismatch")
    return _lib


def _declare(L):
    L.mappo_abi_version.restype = C.c_int
    L.mappo_sm_arch.restype = C.c_int
    L.mappo_last_error.restype = C.c_char_p
    L.mappo_launch_count.restype = C.c_uint64
    L.mappo_env_obs_width.argtypes = [C.c_int, C.c_int, C.c_int]
    L.mappo_env_obs_dim.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    L.mappo_env_obs_dim_cfg.argtypes = [C.POINTER(EnvConfig), C.c_int]
    L.mappo_env_obs_width_cfg.argtypes = [C.POINTER(EnvConfig)]
    L.mappo_env_reset.argtypes = [C.POINTER(EnvConfig), C.POINTER(EnvState), C.c_void_p, C.c_int, C.c_void_p]
    L.mappo_env_step.argtypes = [C.POINTER(EnvConfig), C.POINTER(EnvState), C.c_void_p, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(EnvState), C.c_void_p]
    for name, sig in _EXTRA_SIGS.items():
        fn = getattr(L, name)
        fn.argtypes = sig
        fn.restype = C.c_int


class _Sigs(dict):
    """Table of further entry points, filled by the modules that bind them (kept in one place so the CPU test that
    checks "every symbol in the header is exported" can iterate over it).  Entries added after the library was
    loaded are applied immediately."""

    def update(self, other):
        dict.update(self, other)
        if _lib is not None:
            for name, sig in other.items():
                fn = getattr(_lib, name)
                fn.argtypes = sig
                fn.restype = C.c_int


_EXTRA_SIGS = _Sigs()


def check(rc, what=""):
    if rc != 0:
        msg = lib().mappo_last_error().decode()
        if rc == -2:
            raise NotImplementedError("%s: %s" % (what, msg))
        if rc == -1:
            raise AssertionError("%s: %s" % (what, msg))
        raise MappoError("%s failed (%d): %s" % (what, rc, msg))


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL).  Refuses CPU tensors: no host path exists."""
    if t is None:
        return None
    if not t.is_cuda:
        raise MappoError("mappo_b200: expected a CUDA tensor, got a %s tensor" % t.device)
    if not t.is_contiguous():
        raise MappoError("mappo_b200: expected a contiguous tensor")
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count():
    return int(lib().mappo_launch_count())
