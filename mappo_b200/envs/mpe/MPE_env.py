"""Device-resident batched MPE: the mirror of `MPEEnv(args)` (onpolicy/envs/mpe/MPE_env.py:5-31) and of
`SubprocVecEnv`/`DummyVecEnv` (onpolicy/envs/env_wrappers.py:235-290,661-689) in one object.

The reference builds one `MultiAgentEnv` per OS process and fans a step out over pipes; here all envs
of the shard live in HBM (float64 structure-of-arrays, see include/mappo_b200.h) and `step` is ONE kernel
launch over envs x agents with auto-reset inside.  Spaces follow environment.py:54-100."""
import ctypes as C

import numpy as np
import torch

from ... import _lib
from ..spaces import Box, Discrete, MultiDiscrete, head_sizes

SCENARIOS = {"simple_spread": 0, "simple_speaker_listener": 1, "simple_reference": 2,
             "simple_adversary": 3, "simple_push": 4, "simple_tag": 5,
             "simple_crypto": 6, "simple_crypto_display": 6,      # the display variant only adds prints
             "simple_world_comm": 7}


class _Infos:
    """Lazy `infos`: infos[n][m]['individual_reward'] like the reference (environment.py:135),
    materialised from a device tensor only when indexed (it is only read at log time,
    runner/shared/mpe_runner.py:62-70)."""

    def __init__(self, indiv):
        self._indiv = indiv
        self._host = None

    def __len__(self):
        return self._indiv.shape[0]

    def __getitem__(self, n):
        if self._host is None:
            self._host = self._indiv.cpu().numpy()
        return [{"individual_reward": float(r)} for r in self._host[n]]

    def __iter__(self):
        for n in range(len(self)):
            yield self[n]

    @property
    def individual_reward(self):
        """[N, M] float32 device tensor."""
        return self._indiv


class MPEVecEnv:
    """`envs` object of the runners: reset() / step(actions) / step_async / step_wait / close, attributes
    num_envs, observation_space, share_observation_space, action_space (lists of length M)."""

    def __init__(self, args, num_envs=None, device="cuda", seed=None, env_id_offset=0, host_io=False, auto_reset=True):
        name = args.scenario_name
        if name not in SCENARIOS:
            raise NotImplementedError("Can not support the %s scenario on the device env" % name)
        self.scenario_name = name
        self.scenario = SCENARIOS[name]
        self.num_envs = int(num_envs if num_envs is not None else args.n_rollout_threads)
        self.n = self.num_agents = int(args.num_agents)
        self.num_landmarks = int(args.num_landmarks)
        self.episode_length = int(args.episode_length)
        self.num_adversaries = 0
        if name in ("simple_speaker_listener", "simple_reference"):
            assert self.num_agents == 2, "only 2 agents is supported, check the config.py."
        elif name in ("simple_adversary", "simple_push"):
            self.num_adversaries = 1                                   # simple_adversary.py:16, simple_push.py:14
            if name == "simple_adversary":
                self.num_landmarks = self.num_agents - 1               # simple_adversary.py:17 (not args.num_landmarks)
            else:
                assert self.num_landmarks <= 2, "simple_push colours at most 2 landmarks (simple_push.py:42)"
        elif name in ("simple_crypto", "simple_crypto_display"):       # simple_crypto.py:27-31
            assert self.num_agents == 3, "simple_crypto: adversary, listener, speaker"
            self.num_adversaries = 1
        elif name in ("simple_tag", "simple_world_comm"):              # simple_tag.py:10-13, simple_world_comm.py:10-13:
            default_adv = 3 if name == "simple_tag" else 4             # adversaries first
            self.num_adversaries = int(getattr(args, "num_adversaries", 0) or default_adv)
            good = getattr(args, "num_good_agents", None)
            if good is not None:
                self.n = self.num_agents = self.num_adversaries + int(good)
            assert 0 < self.num_adversaries < self.num_agents
        self.device = torch.device(device)
        L = _lib.lib()
        M, NL, N = self.num_agents, self.num_landmarks, self.num_envs
        # auto_reset=False (the single-env facade, envs/mpe/environment.py): the kernel's done test `step >= episode_length`
        # never fires, so the env keeps stepping past the episode end like the reference's bare MultiAgentEnv
        self._auto_reset = bool(auto_reset)
        self._cfg = _lib.EnvConfig(self.scenario, N, M, NL, self.episode_length if auto_reset else 2 ** 31 - 1,
                                   self.num_adversaries,
                                   int(seed if seed is not None else getattr(args, "seed", 1)) & (2 ** 64 - 1),
                                   int(env_id_offset))
        L.mappo_env_num_landmarks.argtypes = [C.POINTER(_lib.EnvConfig)]
        # landmarks held in the state arrays (simple_world_comm: obstacles + 2 food + 2 forests)
        self.num_landmarks_total = NL = int(L.mappo_env_num_landmarks(C.byref(self._cfg)))
        self.obs_dims = [L.mappo_env_obs_dim_cfg(C.byref(self._cfg), i) for i in range(M)]
        self.obs_width = L.mappo_env_obs_width_cfg(C.byref(self._cfg))
        self.homogeneous = len(set(self.obs_dims)) == 1
        # spaces (environment.py:54-100)
        if name in ("simple_spread", "simple_adversary", "simple_push", "simple_tag"):
            acts = [Discrete(5) for _ in range(M)]
        elif name == "simple_speaker_listener":
            acts = [Discrete(3), Discrete(5)]
        elif name in ("simple_crypto", "simple_crypto_display"):
            acts = [Discrete(4) for _ in range(M)]                     # immobile, not silent: the word (dim_c = 4)
        elif name == "simple_world_comm":                              # the leader moves and speaks (environment.py:66-93)
            acts = [MultiDiscrete([[0, 4], [0, 3]])] + [Discrete(5) for _ in range(M - 1)]
        else:
            acts = [MultiDiscrete([[0, 4], [0, 9]]) for _ in range(M)]
        self.action_space = acts
        self.head_sizes = [head_sizes(a) for a in acts]
        self.act_width = max(len(h) for h in self.head_sizes)
        self.observation_space = [Box(-np.inf, np.inf, (d,), np.float32) for d in self.obs_dims]
        self.share_observation_space = [Box(-np.inf, np.inf, (self.obs_width,), np.float32) for _ in range(M)]
        # world state in HBM
        dev = self.device
        self.p_pos = torch.zeros(2, M, N, dtype=torch.float64, device=dev)
        self.p_vel = torch.zeros(2, M, N, dtype=torch.float64, device=dev)
        self.lm_pos = torch.zeros(2, NL, N, dtype=torch.float64, device=dev)
        self.comm = torch.full((M, N), -1, dtype=torch.int32, device=dev)
        self.goal = torch.full((M, N), -1, dtype=torch.int32, device=dev)
        self.step_count = torch.zeros(N, dtype=torch.int32, device=dev)
        self.episode = torch.zeros(N, dtype=torch.int32, device=dev)
        self._st = self._state_struct(self)
        # default outputs (callers may pass buffer slots instead)
        self._obs = torch.empty(N, self.obs_width, dtype=torch.float32, device=dev)
        self._rew = torch.empty(N, M, 1, dtype=torch.float32, device=dev)
        self._masks = torch.empty(N, M, 1, dtype=torch.float32, device=dev)
        self._dones = torch.empty(N, M, dtype=torch.uint8, device=dev)
        self._indiv = torch.empty(N, M, dtype=torch.float32, device=dev)
        self._pending = None
        self._ids_scratch = None
        self.closed = False
        self._host = None
        self.set_host_io(host_io)

    def set_host_io(self, on):
        """host_io=True: `reset` / `step` return HOST numpy arrays like the reference's vec envs
        (env_wrappers.py:257-272) -- views of pinned staging buffers, valid until the next call -- instead of
        device tensors.  Inputs may be host arrays in either mode."""
        self.host_io = bool(on)
        if self.host_io and self._host is None:
            N, M = self.num_envs, self.num_agents
            pin = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt).pin_memory()
            self._host = dict(obs=pin(N, self.obs_width), rew=pin(N, M, 1), dones=pin(N, M, dt=torch.uint8),
                              indiv=pin(N, M))

    def _to_host(self, obs_dev, with_step_outputs):
        h = self._host
        h["obs"].copy_(obs_dev.view(self.num_envs, self.obs_width), non_blocking=True)
        if with_step_outputs:
            h["rew"].copy_(self._rew_last.view(self.num_envs, self.num_agents, 1), non_blocking=True)
            h["dones"].copy_(self._dones, non_blocking=True)
            h["indiv"].copy_(self._indiv, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        # the device copies stay valid until the next step: handing these host arrays back (SharedReplayBuffer.insert)
        # is served device -> device (host_mirror)
        from ... import host_mirror
        host_mirror.register(h["obs"], obs_dev.view(self.num_envs, self.obs_width))
        if with_step_outputs:
            host_mirror.register(h["rew"], self._rew_last.view(self.num_envs, self.num_agents, 1))
        obs = self._shape_obs(h["obs"])
        return obs.numpy() if torch.is_tensor(obs) else [o.numpy() for o in obs]

    @staticmethod
    def _state_struct(o):
        return _lib.EnvState(*[C.c_void_p(t.data_ptr()) for t in
                               (o.p_pos, o.p_vel, o.lm_pos, o.comm, o.goal, o.step_count, o.episode)])

    # ------------------------------------------------------------------ helpers
    def _shape_obs(self, flat):
        """[N, W] -> reference layout: [N, M, D] for homogeneous agents, else list of [N, D_m] views."""
        if self.homogeneous:
            return flat.view(self.num_envs, self.num_agents, self.obs_dims[0])
        outs, off = [], 0
        for d in self.obs_dims:
            outs.append(flat[:, off:off + d])
            off += d
        return outs

    def actions_to_ids(self, actions):
        """Accept what the reference's runners pass (one-hot `actions_env`, shared/mpe_runner.py:110-121) or
        the policy's action ids; return float32 [N, M, act_width] ids on the device."""
        if isinstance(actions, (list, tuple)):  # ragged per-agent list (separated runner)
            cols = [self.actions_to_ids_one(a, i) for i, a in enumerate(actions)]
            return torch.stack(cols, dim=1).contiguous()
        from ... import gather, host_mirror
        N, M = self.num_envs, self.num_agents
        twin = host_mirror.lookup(actions) if not torch.is_tensor(actions) or not actions.is_cuda else None
        if twin is not None:
            a = twin                                                     # ids a host_mirror.t2n call just brought out
        elif torch.is_tensor(actions):
            a = actions.to(self.device, non_blocking=True)
        else:
            a = torch.from_numpy(np.ascontiguousarray(actions)).to(self.device, non_blocking=True)
        a = a.reshape(N, M, -1)
        if a.shape[-1] == self.act_width and all(len(h) == self.act_width for h in self.head_sizes):
            return a.to(torch.float32).contiguous()
        sizes = self.head_sizes[0]
        assert all(h == sizes for h in self.head_sizes) and a.shape[-1] == sum(sizes), \
            "actions must be ids [N,M,%d] or one-hot [N,M,%d]" % (self.act_width, sum(sizes))
        if a.dtype not in (torch.float32, torch.float64, torch.uint8):
            a = a.to(torch.float32)
        if self._ids_scratch is None:
            self._ids_scratch = torch.zeros(N, M, self.act_width, dtype=torch.float32, device=self.device)
        gather.onehot_to_ids(a.reshape(N * M, -1).contiguous(), sizes, self._ids_scratch.view(N * M, -1))   # one launch
        return self._ids_scratch

    def actions_to_ids_one(self, a, agent):
        a = torch.as_tensor(a, device=self.device) if not torch.is_tensor(a) else a.to(self.device)
        a = a.reshape(self.num_envs, -1)
        sizes = self.head_sizes[agent]
        if a.shape[-1] == len(sizes):
            ids = a.to(torch.float32)
        else:
            cols, off = [], 0
            for s in sizes:
                cols.append(a[:, off:off + s].argmax(dim=-1))
                off += s
            ids = torch.stack(cols, dim=-1).to(torch.float32)
        if ids.shape[-1] < self.act_width:
            ids = torch.nn.functional.pad(ids, (0, self.act_width - ids.shape[-1]))
        return ids

    # ------------------------------------------------------------------ vec-env contract
    def seed(self, seed):
        self._cfg.seed = int(seed) & (2 ** 64 - 1)

    def reset(self, out_obs=None, use_given_state=False):
        obs = self._obs if out_obs is None else out_obs
        assert obs.is_contiguous() and obs.numel() == self.num_envs * self.obs_width
        _lib.check(_lib.lib().mappo_env_reset(C.byref(self._cfg), C.byref(self._st), _lib.ptr(obs),
                                              1 if use_given_state else 0, _lib.stream_ptr()), "env_reset")
        if self.host_io:
            return self._to_host(obs, False)
        return self._shape_obs(obs.view(self.num_envs, self.obs_width))

    def set_state(self, p_pos=None, p_vel=None, lm_pos=None, goal=None, comm=None, step_count=None):
        """Install a world state given in the reference's per-env layout ([N,M,2], [N,L,2], [N,M]);
        used by parity tests to inject the reference's own states."""
        def put(dst, src, perm):
            if src is not None:
                dst.copy_(torch.as_tensor(np.asarray(src)).permute(*perm).to(dst.dtype))
        put(self.p_pos, p_pos, (2, 1, 0)); put(self.p_vel, p_vel, (2, 1, 0)); put(self.lm_pos, lm_pos, (2, 1, 0))
        put(self.goal, goal, (1, 0)); put(self.comm, comm, (1, 0))
        if step_count is not None:
            self.step_count.copy_(torch.as_tensor(np.asarray(step_count)).to(torch.int32))

    def get_state(self):
        return dict(p_pos=self.p_pos.permute(2, 1, 0).cpu().numpy(), p_vel=self.p_vel.permute(2, 1, 0).cpu().numpy(),
                    lm_pos=self.lm_pos.permute(2, 1, 0).cpu().numpy(), goal=self.goal.t().cpu().numpy(),
                    comm=self.comm.t().cpu().numpy(), step_count=self.step_count.cpu().numpy())

    def step_ids(self, action_ids, out_obs=None, out_rewards=None, out_masks=None, reset_state=None):
        """Hot-path entry: `action_ids` float32 [N, M, act_width] on the device (a SharedReplayBuffer.actions
        slot); outputs may be buffer slots so the step writes straight into the rollout storage."""
        obs = self._obs if out_obs is None else out_obs
        rew = self._rew if out_rewards is None else out_rewards
        masks = self._masks if out_masks is None else out_masks
        rs = None
        if reset_state is not None:
            keep = []
            class _RS: pass
            o = _RS()
            dev = self.device
            o.p_pos = torch.as_tensor(np.asarray(reset_state["p_pos"])).permute(2, 1, 0).contiguous().to(dev, torch.float64)
            o.lm_pos = torch.as_tensor(np.asarray(reset_state["lm_pos"])).permute(2, 1, 0).contiguous().to(dev, torch.float64)
            o.goal = torch.as_tensor(np.asarray(reset_state["goal"])).t().contiguous().to(dev, torch.int32)
            o.p_vel, o.comm, o.step_count, o.episode = o.p_pos, o.goal, o.goal, o.goal
            self._rs_keep = o
            rs = C.byref(self._state_struct(o))
        from ... import ops
        with ops.TIMERS.span("env_step"):
            _lib.check(_lib.lib().mappo_env_step(
                C.byref(self._cfg), C.byref(self._st), _lib.ptr(action_ids), int(action_ids.shape[-1]),
                _lib.ptr(obs), _lib.ptr(rew), _lib.ptr(masks), _lib.ptr(self._dones), _lib.ptr(self._indiv),
                rs, _lib.stream_ptr()), "env_step")
        self._rew_last = rew
        return obs, rew, masks

    def step_async(self, actions):
        self._pending = self.actions_to_ids(actions)

    def step_wait(self):
        obs, rew, _ = self.step_ids(self._pending)
        self._pending = None
        if self.host_io:
            o = self._to_host(obs, True)
            h = self._host
            return o, h["rew"].numpy(), h["dones"].numpy().astype(bool), _Infos(h["indiv"])
        return (self._shape_obs(obs.view(self.num_envs, self.obs_width)), rew.view(self.num_envs, self.num_agents, 1),
                self._dones.bool(), _Infos(self._indiv))

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.closed = True

    def render(self, mode="rgb_array", env_ids=None, width=None, height=None, as_numpy=True):
        """env_wrappers.py:292-297,695-702: `rgb_array` returns the frames of all envs stacked, each env contributing
        its list of one shared-viewer image -> uint8 [n, 1, 700, 700, 3] (the runner reads `[0][0]`,
        shared/mpe_runner.py:195).  Rasterised on the device (csrc/render.cu); `env_ids` restricts the envs drawn
        (default: all, refused above 1 GiB of pixels), `as_numpy=False` keeps the frames in HBM.  `human` needs a
        window: unavailable on this box."""
        if mode != "rgb_array":
            raise NotImplementedError("render(mode=%r): only 'rgb_array' exists without a display" % (mode,))
        from . import rendering
        w = rendering.WIDTH if width is None else int(width)
        h = rendering.HEIGHT if height is None else int(height)
        n = self.num_envs if env_ids is None else len(env_ids)
        if n * w * h * 3 > (1 << 30):
            raise ValueError("render: %d frames of %dx%d exceed 1 GiB; pass env_ids" % (n, w, h))
        frames = rendering.render_rgb(self, env_ids, w, h).unsqueeze(1)
        return frames.cpu().numpy() if as_numpy else frames


def MPEEnv(args, **kw):
    """onpolicy/envs/mpe/MPE_env.py:5-31: ONE env with the reference's per-agent-list MultiAgentEnv API
    (envs/mpe/environment.py here).  The batched env the runners use is MPEVecEnv."""
    from .environment import MultiAgentEnv
    return MultiAgentEnv(args, **kw)
