"""Single-env facade: the reference's `MultiAgentEnv` / `World` surface (onpolicy/envs/mpe/environment.py:17-163,
onpolicy/envs/mpe/core.py:6-140) as a thin VIEW over one env of the device-resident batched env (MPE_env.MPEVecEnv).

    env = MPEEnv(args)                       # onpolicy/envs/mpe/MPE_env.py:5-31: one env
    obs_n = env.reset()                      # list of M float64 arrays
    obs_n, reward_n, done_n, info_n = env.step(action_n)      # per-agent lists, action_n[i] one-hot (or ids)
    env.world.agents[i].state.p_pos          # read / assign: goes through the device state

Nothing is simulated here: `step` is one launch of the env kernel on a batch (of one, or on the batch this view
belongs to), the lists are unpacked from its outputs.  Like the reference's `MultiAgentEnv.step` -- and unlike its
vec-env wrappers (env_wrappers.py:145-154) -- the single env does NOT reset itself when the episode ends: `done_n`
turns True at `current_step >= world_length` (environment.py:179-185) and the caller resets.  Scenario callbacks are
not Python objects here (the device env recognises the scenario by name); `world.agents[i]` /
`world.landmarks[j]` expose the per-entity constants and live state the reference's objects carry."""
import numpy as np
import torch

from .MPE_env import MPEVecEnv

# core.py:120-131
DT, DAMPING, CONTACT_FORCE, CONTACT_MARGIN = 0.1, 0.25, 1e2, 1e-3


class _State(object):
    """EntityState / AgentState (core.py:6-20): p_pos, p_vel (and c for agents), backed by the device arrays."""

    def __init__(self, env, kind, idx):
        self._env, self._kind, self._idx = env, kind, idx

    def _get(self, t):
        return t[:, self._idx, self._env._n].detach().cpu().numpy().astype(np.float64)

    def _set(self, t, v):
        t[:, self._idx, self._env._n] = torch.as_tensor(np.asarray(v, dtype=np.float64), device=t.device)

    @property
    def p_pos(self):
        v = self._env._vec
        return self._get(v.p_pos if self._kind == "agent" else v.lm_pos)

    @p_pos.setter
    def p_pos(self, value):
        v = self._env._vec
        self._set(v.p_pos if self._kind == "agent" else v.lm_pos, value)

    @property
    def p_vel(self):
        if self._kind != "agent":
            return np.zeros(2)
        return self._get(self._env._vec.p_vel)

    @p_vel.setter
    def p_vel(self, value):
        if self._kind == "agent":
            self._set(self._env._vec.p_vel, value)

    @property
    def c(self):
        """communication state as the one-hot word the reference stores (core.py:280-287); zeros while silent"""
        v = self._env._vec
        dim_c = self._env.world.dim_c
        out = np.zeros(dim_c)
        if self._kind == "agent" and dim_c > 0:
            w = int(v.comm[self._idx, self._env._n].item())
            if 0 <= w < dim_c:
                out[w] = 1.0
        return out


class _Entity(object):
    def __init__(self, env, kind, idx, name):
        self.name = name
        self.state = _State(env, kind, idx)
        self.movable = kind == "agent"
        self.collide = True
        self.mass = 1.0

    @property
    def initial_mass(self):
        return self.mass


class World(object):
    """core.py:114-140: the fields the reference's World carries, with entity lists whose states are device views."""

    def __init__(self, env, args):
        M, L = env._vec.num_agents, env._vec.num_landmarks_total
        self.agents = [_Entity(env, "agent", i, "agent %d" % i) for i in range(M)]
        self.landmarks = [_Entity(env, "landmark", j, "landmark %d" % j) for j in range(L)]
        for lm in self.landmarks:
            lm.movable = False
        self.walls = []
        self.dim_c = {"simple_speaker_listener": 3, "simple_reference": 10, "simple_crypto": 4,
                      "simple_crypto_display": 4, "simple_world_comm": 4}.get(env._vec.scenario_name, 2)
        self.dim_p, self.dim_color = 2, 3
        self.dt, self.damping = DT, DAMPING
        self.contact_force, self.contact_margin = CONTACT_FORCE, CONTACT_MARGIN
        self.world_length = int(args.episode_length)
        self.world_step = 0
        self.num_agents, self.num_landmarks = M, int(env._vec.num_landmarks)
        self.collaborative = env._vec.scenario_name in ("simple_spread", "simple_speaker_listener", "simple_reference")

    @property
    def entities(self):
        return self.agents + self.landmarks

    @property
    def policy_agents(self):
        return self.agents

    @property
    def scripted_agents(self):
        return []


class MultiAgentEnv(object):
    """environment.py:13-163.  Constructed from the flag namespace (`MPEEnv(args)`) or as a view of env `n` of an
    existing batched env (`MultiAgentEnv.view(vec_env, n)`: stepping a view steps the whole batch it belongs to, with
    the other envs' actions left at the no-op)."""
    metadata = {"render.modes": ["human", "rgb_array"]}

    def __init__(self, args=None, device="cuda", vec=None, n=0, **kw):
        if vec is None:
            vec = MPEVecEnv(args, num_envs=1, device=device, auto_reset=False, **kw)
            self._owns = True
        else:
            self._owns = False
        self._vec, self._n = vec, int(n)
        self.world = World(self, args if args is not None else _ArgsOf(vec))
        self.world_length = self.world.world_length
        self.current_step = 0
        self.agents = self.world.policy_agents
        self.n = vec.num_agents
        self.discrete_action_space, self.discrete_action_input = True, False
        self.shared_reward = self.world.collaborative
        self.action_space = vec.action_space
        self.observation_space = vec.observation_space
        self.share_observation_space = vec.share_observation_space
        self._act = torch.zeros(vec.num_envs, vec.num_agents, vec.act_width, dtype=torch.float32, device=vec.device)

    @classmethod
    def view(cls, vec, n):
        return cls(vec=vec, n=n)

    def seed(self, seed=None):
        """environment.py:110-114 (np.random.seed): here the key of the Philox reset streams."""
        self._vec.seed(1 if seed is None else seed)

    def _obs_list(self, obs):
        if self._vec.homogeneous:
            return [obs[self._n, i].detach().cpu().numpy().astype(np.float64) for i in range(self.n)]
        return [o[self._n].detach().cpu().numpy().astype(np.float64) for o in obs]

    def _ids(self, action_n):
        """per-agent one-hot (concatenated one-hots for MultiDiscrete) or ids -> float ids [act_width]"""
        ids = np.zeros((self.n, self._vec.act_width), np.float32)
        for i, a in enumerate(action_n):
            a = np.asarray(a, dtype=np.float64).reshape(-1)
            sizes = self._vec.head_sizes[i]
            if a.size == len(sizes):
                ids[i, :a.size] = a
            else:
                assert a.size == sum(sizes), "agent %d: action of %d entries, expected one-hot of %d" % (i, a.size, sum(sizes))
                off = 0
                for h, s in enumerate(sizes):
                    ids[i, h] = int(np.argmax(a[off:off + s]))
                    off += s
        return ids

    def step(self, action_n):
        """environment.py:117-148 -> (obs_n, reward_n, done_n, info_n), per-agent lists."""
        assert self._owns or not self._vec._auto_reset or True
        self.current_step += 1
        self._act.zero_()
        self._act[self._n] = torch.as_tensor(self._ids(action_n), device=self._act.device)
        v = self._vec
        obs, rew, _ = v.step_ids(self._act)
        obs = v._shape_obs(obs.view(v.num_envs, v.obs_width))
        obs_n = self._obs_list(obs)
        r = rew.view(v.num_envs, v.num_agents)[self._n].detach().cpu().numpy().astype(np.float64)
        ind = v._indiv[self._n].detach().cpu().numpy().astype(np.float64)
        reward_n = [[float(r[i])] for i in range(self.n)]
        done = bool(self.current_step >= self.world_length)            # environment.py:179-185
        done_n = [done] * self.n
        info_n = [{"individual_reward": float(ind[i])} for i in range(self.n)]
        self.world.world_step = self.current_step
        return obs_n, reward_n, done_n, info_n

    def reset(self):
        """environment.py:150-163"""
        self.current_step = 0
        self.world.world_step = 0
        obs = self._vec.reset()
        return self._obs_list(obs)

    def render(self, mode="rgb_array", close=False):
        """environment.py:263-412: `rgb_array` returns [frame] (one shared viewer); `human` needs a window."""
        frames = self._vec.render(mode, env_ids=[self._n])
        return [frames[0, 0]]

    def close(self):
        if self._owns:
            self._vec.close()


class _ArgsOf(object):
    """the few flags World needs, recovered from an existing batched env"""

    def __init__(self, vec):
        self.episode_length = vec.episode_length


def MPEEnv(args, **kw):
    """onpolicy/envs/mpe/MPE_env.py:5-31: ONE env with the per-agent-list API.  The batched env the runners use is
    `MPE_env.MPEVecEnv` (the reference builds it as SubprocVecEnv / DummyVecEnv over many of these)."""
    return MultiAgentEnv(args, **kw)
