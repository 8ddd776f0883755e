"""CPU: the C-ABI library loads and exports every symbol include/mappo_b200.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "mappo_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mappo_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__
    __graft_entry__.build()
    from mappo_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 8
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert lib.mappo_abi_version() == 6 and lib.mappo_sm_arch() == 100


def test_obs_widths_match_reference_shapes():
    from mappo_b200 import _lib
    L = _lib.lib()
    assert L.mappo_env_obs_width(0, 3, 3) == 54 and L.mappo_env_obs_dim(0, 3, 3, 0) == 18
    assert L.mappo_env_obs_width(0, 10, 3) == 460 and L.mappo_env_obs_width(0, 10, 10) == 600
    assert L.mappo_env_obs_width(1, 2, 3) == 14 and L.mappo_env_obs_dim(1, 2, 3, 0) == 3
    assert L.mappo_env_obs_width(2, 2, 3) == 42


def test_obs_widths_of_the_adversarial_scenarios():
    """simple_adversary.py:116-137, simple_push.py:81-104, simple_tag.py:125-144 (widths probed on the reference when the
    env fixtures were recorded: tests/golden/env_{adversary,push,tag}.npz)."""
    import ctypes as C
    from mappo_b200 import _lib
    L = _lib.lib()
    cfg = lambda scn, M, NL, nadv=0: C.byref(_lib.EnvConfig(scn, 4, M, NL, 25, nadv, 0, 0))
    assert [L.mappo_env_obs_dim_cfg(cfg(3, 3, 2), i) for i in range(3)] == [8, 10, 10]
    assert [L.mappo_env_obs_dim_cfg(cfg(4, 2, 2), i) for i in range(2)] == [8, 19]
    assert [L.mappo_env_obs_dim_cfg(cfg(5, 4, 2, 3), i) for i in range(4)] == [16, 16, 16, 14]
    assert L.mappo_env_obs_width_cfg(cfg(5, 4, 2, 3)) == 62
    # simple_crypto.py:160-171 (tests/golden/env_crypto.npz): adversary 4, listener 8, speaker 8
    assert [L.mappo_env_obs_dim_cfg(cfg(6, 3, 2), i) for i in range(3)] == [4, 8, 8]
    # simple_world_comm.py:281-288, probed on the reference (4 adversaries, 2 good agents, 1 obstacle): 34 / 28
    assert [L.mappo_env_obs_dim_cfg(cfg(7, 6, 1, 4), i) for i in range(6)] == [34, 34, 34, 34, 28, 28]
    assert L.mappo_env_num_landmarks(cfg(7, 6, 1, 4)) == 5 and L.mappo_env_num_landmarks(cfg(0, 3, 3)) == 3


def test_benchmark_networks_take_the_tensor_core_kernels():
    """Host-side decision of the library (mappo_mlp_kernel_path, no device call): the networks of BASELINE.json's
    configurations must be served by the fused tcgen05 kernels -- a shared-memory map that outgrows the SM silently
    sends a network to the fp32 tile kernels (4x slower), which no parity test notices."""
    from mappo_b200 import ops
    path = lambda in_dim, heads, critic, **kw: ops.kernel_path(ops.net_config(
        in_dim, 64, kw.get("layer_N", 1), kw.get("relu", False), True, kw.get("rnn", 0), heads, critic))
    for relu in (False, True):
        assert path(18, [5], False, relu=relu) == 1 and path(54, [1], True, relu=relu) == 1        # C1 / C2 / C3: 3 agents
        assert path(60, [5], False, relu=relu) == 1 and path(600, [1], True, relu=relu) == 2       # C5: 10 agents, 10 landmarks
        assert path(46, [5], False, relu=relu) == 1 and path(460, [1], True, relu=relu) == 2       # 10 agents, 3 landmarks
    assert path(66, [5], False) == 2                                                               # 15 agents: wide actor too
    assert path(11, [3, 5], False) == 1 and path(21, [5, 10], False) == 1                          # speaker_listener / reference heads
    assert path(18, [5], False, rnn=1) == 3 and path(54, [1], True, rnn=2) == 3                    # C4: GRU chain
    assert path(18, [5], False, layer_N=2) == 0                                                    # deep base: fp32 tiles
    assert path(40, [16], False) == 0 and path(40, [8], False) == 1 and path(31, [16], False) == 1  # 64-wide tile + 16 logits: 229 KB
