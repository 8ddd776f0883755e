"""Headless `render('rgb_array')` for the device env.

The reference's module of this name (onpolicy/envs/mpe/rendering.py) wraps pyglet / OpenGL, which cannot exist on
this box; here the same scene -- entity circles, outlines, communication circles (environment.py:297-412) -- is
rasterised by `mappo_env_render` (csrc/render.cu) straight from the SoA world state into a device buffer.  This file
only assembles what the scenarios fix on the host side in the reference: entity sizes, colours (scenario
`reset_world`s, core.py:183-204) and which agents carry communication circles."""
import ctypes as C
import math

import numpy as np
import torch

from ... import _lib

WIDTH, HEIGHT, RES, CAM_RANGE, LINE_WIDTH = 700, 700, 30, 2.0, 2.0   # environment.py:8,297; rendering.py:59,243


class RenderConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("num_images", C.c_int32), ("num_entities", C.c_int32),
                ("res", C.c_int32), ("reserved", C.c_int32), ("cam_range", C.c_double), ("line_width", C.c_double)]


_P = C.c_void_p
_lib._EXTRA_SIGS.update({
    "mappo_env_render": [C.POINTER(_lib.EnvConfig), C.POINTER(_lib.EnvState), C.POINTER(RenderConfig), _P, _P, _P, _P,
                         _P, _P, _P],
})

_GOOD, _ADV, _GREY = (0.25, 0.25, 0.75), (0.75, 0.25, 0.25), (0.25, 0.25, 0.25)


def unit_circle(res=RES):
    """rendering.make_circle's vertices (rendering.py:244-246), float64 [res, 2]."""
    return np.array([[math.cos(2 * math.pi * i / res), math.sin(2 * math.pi * i / res)] for i in range(res)],
                    dtype=np.float64)


def entity_sizes(env):
    """[E] entity.size: agents then landmarks (core.py:53 default 0.05 and the scenarios' make_world)."""
    name, M, L, nadv = env.scenario_name, env.num_agents, env.num_landmarks_total, env.num_adversaries
    if name == "simple_spread":
        a, l = [0.15] * M, [0.05] * L                                  # simple_spread.py:21
    elif name == "simple_speaker_listener":
        a, l = [0.075] * M, [0.04] * L                                 # simple_speaker_listener.py:22,33
    elif name == "simple_adversary":
        a, l = [0.15] * M, [0.08] * L                                  # simple_adversary.py:24,31
    elif name == "simple_tag":
        a = [0.075 if i < nadv else 0.05 for i in range(M)]            # simple_tag.py:21,31
        l = [0.2] * L
    elif name == "simple_world_comm":                                  # simple_world_comm.py:23,31,38,45
        a = [0.075 if i < nadv else 0.045 for i in range(M)]
        l = [0.2] * (L - 4) + [0.03] * 2 + [0.3] * 2                   # obstacles, food, forests
    else:                                                              # simple_reference, simple_push, simple_crypto
        a, l = [0.05] * M, [0.05] * L
    return np.asarray(a + l, dtype=np.float32)


def comm_dims(env):
    """[E] world.dim_c for agents that are not silent (they get communication circles, environment.py:317-330)."""
    name, M, L = env.scenario_name, env.num_agents, env.num_landmarks_total
    d = [0] * (M + L)
    if name == "simple_speaker_listener":
        d[0] = 3                                                       # the speaker; the listener is silent (:26)
    elif name == "simple_reference":
        d[:M] = [10] * M
    elif name in ("simple_crypto", "simple_crypto_display"):
        d[:M] = [4] * M
    elif name == "simple_world_comm":
        d[0] = 4                                                       # the leader
    return np.asarray(d, dtype=np.int32)


def entity_rgba(env, env_ids):
    """float32 [n, E, 4] on the device: entity.color as the scenarios' reset_world leaves it, alpha 0.5 for agents and
    1 for landmarks (environment.py:315,338)."""
    name, M, L, nadv = env.scenario_name, env.num_agents, env.num_landmarks_total, env.num_adversaries
    dev = env.device
    ids = torch.as_tensor(env_ids, dtype=torch.long, device=dev)
    n = ids.numel()
    t = lambda rows: torch.tensor(rows, dtype=torch.float32, device=dev)
    rgb = torch.zeros(n, M + L, 3, dtype=torch.float32, device=dev)
    goal = env.goal[:, ids].t().long()                                 # [n, M] goal landmark ids
    if name in ("simple_spread", "simple_tag", "simple_adversary"):    # assign_agent_colors / assign_landmark_colors
        rgb[:, :M] = t([_ADV if i < nadv else _GOOD for i in range(M)])
        rgb[:, M:] = t(_GREY)
        if name == "simple_adversary":                                 # simple_adversary.py:42-43: the goal landmark
            rgb[torch.arange(n, device=dev), M + goal[:, 0]] = t((0.15, 0.65, 0.15))
    elif name == "simple_speaker_listener":                            # simple_speaker_listener.py:47-55
        lm = t([(0.65, 0.15, 0.15), (0.15, 0.65, 0.15), (0.15, 0.15, 0.65)])[:L]
        rgb[:, M:] = lm
        rgb[:, 0] = t(_GREY)
        rgb[:, 1] = lm[goal[:, 0]] + 0.45                              # goal_a (the listener) takes goal_b's colour + 0.45
    elif name == "simple_reference":                                   # simple_reference.py:45-52
        lm = t([(0.75, 0.25, 0.25), (0.25, 0.75, 0.25), (0.25, 0.25, 0.75)])[:L]
        rgb[:, M:] = lm
        rgb[:, 1] = lm[goal[:, 0]]                                     # agents[0].goal_a (= agent 1) <- agents[0].goal_b
        rgb[:, 0] = lm[goal[:, 1]]
    elif name == "simple_push":                                        # simple_push.py:42-56
        lm = torch.full((L, 3), 0.1, dtype=torch.float32, device=dev)
        for i in range(L):
            lm[i, i + 1] += 0.8
        rgb[:, M:] = lm
        bump = torch.zeros(L, 3, dtype=torch.float32, device=dev)
        for j in range(L):
            bump[j, j + 1] = 0.5
        for i in range(M):
            rgb[:, i] = t(_ADV) if i < nadv else t(_GREY) + bump[goal[:, i]]
    elif name in ("simple_crypto", "simple_crypto_display"):           # simple_crypto.py:50-66
        rgb[:, :M] = t([_ADV, _GREY, _GREY])
        lm = torch.eye(4, dtype=torch.float32, device=dev)[:L, :3]     # colour e_l; only r, g, b reach the rasteriser
        rgb[:, M:] = lm
        rgb[:, 1] = lm[goal[:, 0]]                                     # world.agents[1].color = goal.color
    elif name == "simple_world_comm":                                  # simple_world_comm.py:79-89
        rgb[:, :M] = t([(0.95, 0.45, 0.45) if i < nadv else (0.45, 0.95, 0.45) for i in range(M)])
        rgb[:, 0] -= 0.3                                               # the leader
        rgb[:, M:M + L - 4] = t(_GREY)
        rgb[:, M + L - 4:M + L - 2] = t((0.15, 0.15, 0.65))            # food
        rgb[:, M + L - 2:] = t((0.6, 0.9, 0.6))                        # forests
    else:
        raise NotImplementedError(name)
    alpha = torch.ones(n, M + L, 1, dtype=torch.float32, device=dev)
    alpha[:, :M] = 0.5
    return torch.cat([rgb, alpha], dim=-1).contiguous()


def render_rgb(env, env_ids=None, width=WIDTH, height=HEIGHT, out=None):
    """uint8 device tensor [n, height, width, 3] of the envs `env_ids` (default: all)."""
    N, M, L = env.num_envs, env.num_agents, env.num_landmarks_total
    if env_ids is None:
        env_ids = np.arange(N)
    env_ids = np.asarray(env_ids, dtype=np.int32).reshape(-1)
    assert env_ids.size == 0 or (0 <= env_ids.min() and env_ids.max() < N), "env_ids outside [0, num_envs)"
    n = int(env_ids.size)
    dev = env.device
    if out is None:
        out = torch.empty(n, height, width, 3, dtype=torch.uint8, device=dev)
    assert out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and out.numel() == n * height * width * 3
    ids_dev = torch.as_tensor(env_ids, device=dev)
    size = torch.as_tensor(entity_sizes(env), device=dev)
    rgba = entity_rgba(env, env_ids)
    cdim = torch.as_tensor(comm_dims(env), device=dev)
    unit = torch.as_tensor(unit_circle(RES), device=dev)
    rc = RenderConfig(int(width), int(height), n, M + L, RES, 0, CAM_RANGE, LINE_WIDTH)
    _lib.check(_lib.lib().mappo_env_render(C.byref(env._cfg), C.byref(env._st), C.byref(rc), _lib.ptr(ids_dev),
                                           _lib.ptr(size), _lib.ptr(rgba), _lib.ptr(cdim), _lib.ptr(unit),
                                           _lib.ptr(out), _lib.stream_ptr()), "env_render")
    return out
