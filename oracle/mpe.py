"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's MPE step for the
three target scenarios (plus simple_adversary / simple_push / simple_tag), batched over envs in
float64 numpy.

Follows, operation for operation (so results agree with the reference to the last f64 bit wherever
numpy/libm agree):
  * MultiAgentEnv.step / reset / _set_action / _get_done      onpolicy/envs/mpe/environment.py:117-163,179-185,194-256
  * World.step, apply_action_force, apply_environment_force,
    integrate_state, update_agent_state, get_entity_collision_force   onpolicy/envs/mpe/core.py:207-321
  * simple_spread   make_world/reset_world/reward/observation   scenarios/simple_spread.py:7-45,72-103
  * simple_speaker_listener                                       scenarios/simple_speaker_listener.py:7-98
  * simple_reference                                              scenarios/simple_reference.py:7-97
  * simple_adversary / simple_push / simple_tag                   scenarios/simple_adversary.py:9-137,
                                                                  simple_push.py:8-104, simple_tag.py:7-144
    (per-agent size / accel / max_speed, agent-landmark contacts, individual rewards: core.py:229-278,290-321)
  * auto-reset on done (obs replaced by reset obs, reward/done kept)  onpolicy/envs/env_wrappers.py:140-157,676-682

Reset states come from Philox (oracle/philox.py) instead of the reference's process-global MT19937
(see the note there); parity tests against the reference inject the reference's own reset states.
Pinned against reference-generated fixtures in tests/golden/env_*.npz by tests/test_oracle_env.py.
"""
import numpy as np

from . import philox

DT = 0.1              # core.py:124
DAMPING = 0.25        # core.py:126
CONTACT_FORCE = 1e2   # core.py:128
CONTACT_MARGIN = 1e-3 # core.py:129
SENSITIVITY = 5.0     # environment.py:243

# goal colour tables (scenario-specific): simple_speaker_listener.py:50-52, simple_reference.py:47-49
COLORS_SPEAKER_LISTENER = np.array([[0.65, 0.15, 0.15], [0.15, 0.65, 0.15], [0.15, 0.15, 0.65]])
COLORS_REFERENCE = np.array([[0.75, 0.25, 0.25], [0.25, 0.75, 0.25], [0.25, 0.25, 0.75]])

STREAM_RESET = 0x52455345  # 'RESE': Philox key word 1 for reset draws


def np_sum_small(vals):
    """np.sum over a short 1-D float64 sequence with numpy's pairwise-sum unrolling (n < 128):
    sequential for n < 8, else 8 partial sums combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) then the
    remainder added in order.  `vals` is a list of [N] arrays (summed elementwise with that order).
    Restates what environment.py:141 (`np.sum(reward_n)`) evaluates."""
    n = len(vals)
    if n < 8:
        res = vals[0].copy() if n else 0.0
        for v in vals[1:]:
            res = res + v
        return res
    r = [vals[j].copy() for j in range(8)]
    i = 8
    while i < n - (n % 8):
        for j in range(8):
            r[j] = r[j] + vals[i + j]
        i += 8
    res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
    while i < n:
        res = res + vals[i]
        i += 1
    return res


class ScenarioSpec:
    """Static description of a scenario (what make_world sets)."""

    def __init__(self, name, M, L, num_adversaries=None):
        self.name, self.M, self.L = name, M, L
        if name == "simple_spread":
            self.dim_c = 2
            self.movable = [True] * M
            self.silent = [True] * M
            self.collide = True
            self.agent_size = 0.15
            self.landmark_scale = 0.8
            self.n_goals = 0
            self.obs_dims = [4 + 2 * L + 4 * (M - 1)] * M
            self.act_heads = [[5]] * M            # Discrete(5)
        elif name == "simple_speaker_listener":
            assert M == 2 and L == 3, "simple_speaker_listener: 2 agents, 3 landmarks (scenario asserts / colour table)"
            self.dim_c = 3
            self.movable = [False, True]
            self.silent = [False, True]
            self.collide = False
            self.agent_size = 0.075
            self.landmark_scale = 1.0
            self.n_goals = 1
            self.obs_dims = [3, 2 + 2 * L + 3]
            self.act_heads = [[3], [5]]           # speaker Discrete(3) (comm), listener Discrete(5)
        elif name == "simple_reference":
            assert M == 2 and L == 3, "simple_reference: 2 agents, 3 landmarks (scenario asserts / colour table)"
            self.dim_c = 10
            self.movable = [True, True]
            self.silent = [False, False]
            self.collide = False
            self.agent_size = 0.05
            self.landmark_scale = 0.8
            self.n_goals = 2
            self.obs_dims = [2 + 2 * L + 3 + 10] * M
            self.act_heads = [[5, 10]] * M        # MultiDiscrete([[0,4],[0,9]])
        elif name == "simple_adversary":                                # simple_adversary.py:9-36
            assert L == M - 1, "simple_adversary: num_landmarks = num_agents - 1"
            self.n_adv = 1
            self.collide = False
            self.agent_size = 0.15
            self.landmark_scale = 1.0
            self.n_goals = 1
            self.shared_goal = True
            self.collaborative = False
            self.obs_dims = [2 * L + 2 * (M - 1)] + [2 + 2 * L + 2 * (M - 1)] * (M - 1)
        elif name == "simple_push":                                     # simple_push.py:8-36
            assert L <= 2, "simple_push: landmark.color[i + 1] limits it to 2 landmarks"
            self.n_adv = 1
            self.collide = True
            self.agent_size = 0.05                                      # Entity default, core.py:38
            self.landmark_scale = 0.8
            self.n_goals = 1
            self.shared_goal = True
            self.collaborative = False
            self.obs_dims = [2 + 2 * L + 2 * (M - 1)] + [2 + 2 + 3 + 2 * L + 3 * L + 2 * (M - 1)] * (M - 1)
        elif name == "simple_tag":                                      # simple_tag.py:7-38
            self.n_adv = n_adv = int(num_adversaries if num_adversaries else 3)
            assert 0 < n_adv < M
            self.collide = True
            self.sizes = [0.075] * n_adv + [0.05] * (M - n_adv)
            self.accel = [3.0] * n_adv + [4.0] * (M - n_adv)
            self.max_speed = [1.0] * n_adv + [1.3] * (M - n_adv)
            self.landmark_collide, self.landmark_size = True, 0.2
            self.agent_size = None
            self.landmark_scale = 0.8
            self.n_goals = 0
            self.collaborative = False
            n_good = M - n_adv
            self.obs_dims = [4 + 2 * L + 2 * (M - 1) + 2 * n_good] * n_adv + \
                            [4 + 2 * L + 2 * (M - 1) + 2 * (n_good - 1)] * n_good
        elif name == "simple_world_comm":                               # simple_world_comm.py:7-53
            # world.landmarks = L obstacles (collide, 0.2) + 2 food (0.03) + 2 forests (0.3); agent 0 leads and speaks
            self.n_adv = n_adv = int(num_adversaries if num_adversaries else 4)
            assert 0 < n_adv < M
            self.n_extra = 4
            self.collide = True
            self.sizes = [0.075] * n_adv + [0.045] * (M - n_adv)
            self.accel = [3.0] * n_adv + [4.0] * (M - n_adv)
            self.max_speed = [1.0] * n_adv + [1.3] * (M - n_adv)
            self.landmark_collide, self.landmark_size = True, 0.2
            self.food_size, self.forest_size = 0.03, 0.3
            self.agent_size = None
            self.landmark_scale = 0.8
            self.n_goals = 0
            self.collaborative = False
            self.dim_c = 4
            self.movable = [True] * M
            self.silent = [False] + [True] * (M - 1)
            self.act_heads = [[5, 4]] + [[5]] * (M - 1)                 # leader: MultiDiscrete([[0,4],[0,3]])
            n_good, Lt = M - n_adv, L + 4
            self.obs_dims = [4 + 2 * Lt + 2 * (M - 1) + 2 * n_good + 2 + 4] * n_adv + \
                            [4 + 2 * Lt + 2 * (M - 1) + 2 + 2 * (n_good - 1)] * n_good
        elif name in ("simple_crypto", "simple_crypto_display"):         # simple_crypto.py:21-47 (display: same + prints)
            assert M == 3 and 1 <= L <= 4, "simple_crypto: 3 agents (adversary, listener, speaker), dim_c = 4 colours"
            self.name = "simple_crypto"
            self.dim_c = 4
            self.movable = [False] * M
            self.silent = [False] * M
            self.collide = False
            self.agent_size = 0.05
            self.landmark_scale = 1.0
            self.n_goals = 2                                            # goal landmark, key landmark
            self.n_adv = 1
            self.collaborative = False
            self.obs_dims = [4, 8, 8]
            self.act_heads = [[4]] * M                                  # not movable, not silent: Discrete(dim_c)
        else:
            raise NotImplementedError(name)
        if name in ("simple_adversary", "simple_push", "simple_tag"):
            self.dim_c = 2
            self.movable = [True] * M
            self.silent = [True] * M
            self.act_heads = [[5]] * M
        if not hasattr(self, "sizes"):
            self.sizes = [self.agent_size] * M
            self.accel = [None] * M
            self.max_speed = [None] * M
            self.landmark_collide, self.landmark_size = False, 0.05
        self.shared_goal = getattr(self, "shared_goal", False)
        self.collaborative = getattr(self, "collaborative", True)
        self.n_adv = getattr(self, "n_adv", 0)
        self.n_extra = getattr(self, "n_extra", 0)                      # landmarks behind the L the flag counts
        self.L_obst = L
        self.n_draws = self.n_goals + 2 * M + 2 * (L + self.n_extra)


class OracleMPE:
    def __init__(self, scenario, num_envs, num_agents, num_landmarks, episode_length,
                 seed=0, env_id_offset=0, num_adversaries=None):
        self.spec = ScenarioSpec(scenario, num_agents, num_landmarks, num_adversaries)
        self.N, self.M, self.L = num_envs, num_agents, num_landmarks + self.spec.n_extra   # L: all of world.landmarks
        self.world_length = episode_length
        self.seed = seed
        self.env_ids = np.arange(num_envs, dtype=np.uint64) + np.uint64(env_id_offset)
        N, M, L = self.N, self.M, self.L
        self.p_pos = np.zeros((N, M, 2))
        self.p_vel = np.zeros((N, M, 2))
        self.lm_pos = np.zeros((N, L, 2))
        self.comm = np.zeros((N, M, self.spec.dim_c))
        self.goal = np.full((N, M), -1, dtype=np.int64)
        self.step_count = np.zeros(N, dtype=np.int64)
        self.episode = np.zeros(N, dtype=np.int64)   # index of the NEXT reset draw

    # ---------------------------------------------------------------- reset
    def philox_reset_state(self, idx):
        """Reset state for envs `idx` from Philox: draw order = [goal ids][agent x,y ...][landmark x,y ...]."""
        sp = self.spec
        d = philox.draws(self.env_ids[idx], self.episode[idx], sp.n_draws, self.seed, STREAM_RESET)
        k = 0
        goal = np.full((len(idx), self.M), -1, dtype=np.int64)
        for g in range(sp.n_goals):
            # uniform landmark id: floor(u32 * L / 2**32)
            goal[:, g] = (d[:, k].astype(np.uint64) * np.uint64(self.L)) >> np.uint64(32)
            k += 1
        if sp.shared_goal:                                              # agent.goal_a = goal for every agent
            goal[:, :] = goal[:, :1]
        u = philox.u32_to_sym(d[:, k:])
        p_pos = u[:, :2 * self.M].reshape(-1, self.M, 2).copy()
        lm = u[:, 2 * self.M:].reshape(-1, self.L, 2)
        lm_pos = lm * 0.8 if sp.landmark_scale == 0.8 else lm.copy()
        return dict(p_pos=p_pos, lm_pos=lm_pos, goal=goal)

    def _apply_reset(self, idx, state):
        self.p_pos[idx] = state["p_pos"]
        self.lm_pos[idx] = state["lm_pos"]
        self.p_vel[idx] = 0.0
        self.comm[idx] = 0.0
        if "goal" in state and state["goal"] is not None:
            self.goal[idx] = state["goal"]
        self.step_count[idx] = 0
        self.episode[idx] += 1

    def reset(self, state=None):
        idx = np.arange(self.N)
        self._apply_reset(idx, state if state is not None else self.philox_reset_state(idx))
        return self.observations()

    def set_state(self, p_pos=None, p_vel=None, lm_pos=None, goal=None, comm=None, step_count=None):
        for name, v in (("p_pos", p_pos), ("p_vel", p_vel), ("lm_pos", lm_pos), ("goal", goal),
                        ("comm", comm), ("step_count", step_count)):
            if v is not None:
                getattr(self, name)[...] = v

    # ---------------------------------------------------------------- step
    def step(self, actions, reset_state=None):
        """actions: int array [N, M, n_heads] of action ids (head 0 = move or speaker-comm, head 1 = comm).
        Returns (obs list of M arrays [N, D_m] f64, rewards [N, M, 1] f64, dones [N, M] bool,
        individual_reward [N, M] f64)."""
        sp, N, M = self.spec, self.N, self.M
        actions = np.asarray(actions).reshape(N, M, -1)
        self.step_count += 1                                            # environment.py:118
        # _set_action: one-hot -> u, c (environment.py:194-256)
        u = np.zeros((N, M, 2))
        c_act = np.zeros((N, M, sp.dim_c))
        for i in range(M):
            head = 0
            if sp.movable[i]:
                a = actions[:, i, head]
                onehot = np.eye(5)[a]
                u[:, i, 0] += onehot[:, 1] - onehot[:, 2]
                u[:, i, 1] += onehot[:, 3] - onehot[:, 4]
                u[:, i] *= SENSITIVITY if sp.accel[i] is None else sp.accel[i]   # environment.py:243-246
                head += 1
            if not sp.silent[i]:
                c_act[:, i] = np.eye(sp.dim_c)[actions[:, i, head]]
        # World.step (core.py:207-226)
        force = [None] * M
        for i in range(M):                                              # apply_action_force core.py:229-238
            if sp.movable[i]:
                force[i] = (1.0 if sp.accel[i] is None else 1.0 * sp.accel[i]) * u[:, i] + 0.0
        if sp.collide:                                                  # apply_environment_force core.py:241-263
            for a in range(M):
                for b in range(a + 1, M):
                    dist_min = sp.sizes[a] + sp.sizes[b]
                    delta = self.p_pos[:, a] - self.p_pos[:, b]
                    dist = np.sqrt(np.sum(np.square(delta), axis=-1))
                    k = CONTACT_MARGIN
                    pen = np.logaddexp(0, -(dist - dist_min) / k) * k
                    f = CONTACT_FORCE * delta / dist[:, None] * pen[:, None]
                    fa = 1.0 * f
                    fb = -(1 / 1.0) * f
                    force[a] = fa + force[a]
                    force[b] = fb + force[b]
                if sp.landmark_collide:                                 # entities = agents + landmarks: (agent a, landmark l)
                    for l in range(sp.L_obst):                          # food and forests do not collide
                        dist_min = sp.sizes[a] + sp.landmark_size
                        delta = self.p_pos[:, a] - self.lm_pos[:, l]
                        dist = np.sqrt(np.sum(np.square(delta), axis=-1))
                        k = CONTACT_MARGIN
                        pen = np.logaddexp(0, -(dist - dist_min) / k) * k
                        f = CONTACT_FORCE * delta / dist[:, None] * pen[:, None]
                        force[a] = +f + force[a]                        # core.py:318-320 (landmark not movable)
        for i in range(M):                                              # integrate_state core.py:265-278
            if not sp.movable[i]:
                continue
            self.p_vel[:, i] = self.p_vel[:, i] * (1 - DAMPING)
            if force[i] is not None:
                self.p_vel[:, i] += (force[i] / 1.0) * DT
            if sp.max_speed[i] is not None:                             # core.py:272-277
                v = self.p_vel[:, i]
                speed = np.sqrt(np.square(v[:, 0]) + np.square(v[:, 1]))
                fast = speed > sp.max_speed[i]
                with np.errstate(invalid="ignore", divide="ignore"):
                    clamped = v / np.sqrt(np.square(v[:, 0]) + np.square(v[:, 1]))[:, None] * sp.max_speed[i]
                self.p_vel[:, i] = np.where(fast[:, None], clamped, v)
            self.p_pos[:, i] += self.p_vel[:, i] * DT
        for i in range(M):                                              # update_agent_state core.py:280-287
            self.comm[:, i] = 0.0 if sp.silent[i] else c_act[:, i] + 0.0
        obs = self.observations()
        indiv = self.rewards()                                          # [N, M]
        if sp.collaborative:
            total = np_sum_small([indiv[:, i] for i in range(M)])       # environment.py:141-143 (collaborative)
            rewards = np.repeat(total[:, None, None], M, axis=1)
        else:
            rewards = indiv[:, :, None].copy()
        done_env = self.step_count >= self.world_length                 # environment.py:179-185
        dones = np.repeat(done_env[:, None], M, axis=1)
        idx = np.nonzero(done_env)[0]
        if len(idx):                                                    # env_wrappers.py:145-154
            st = reset_state if reset_state is not None else self.philox_reset_state(idx)
            self._apply_reset(idx, st)
            new_obs = self.observations()
            for i in range(M):
                obs[i][idx] = new_obs[i][idx]
        return obs, rewards, dones, indiv

    # ---------------------------------------------------------------- scenario callbacks
    def observations(self):
        sp, N, M, L = self.spec, self.N, self.M, self.L
        out = []
        for i in range(M):
            pos_i = self.p_pos[:, i]
            ent = [self.lm_pos[:, l] - pos_i for l in range(L)]
            if sp.name == "simple_spread":                              # simple_spread.py:86-103
                other = [self.p_pos[:, j] - pos_i for j in range(M) if j != i]
                comm = [self.comm[:, j] for j in range(M) if j != i]
                o = np.concatenate([self.p_vel[:, i], pos_i] + ent + other + comm, axis=-1)
            elif sp.name == "simple_speaker_listener":                  # simple_speaker_listener.py:75-98
                if i == 0:
                    o = COLORS_SPEAKER_LISTENER[self.goal[:, 0]].copy()
                else:
                    o = np.concatenate([self.p_vel[:, 1]] + ent + [self.comm[:, 0]], axis=-1)
            elif sp.name == "simple_reference":                         # simple_reference.py:69-97
                color = COLORS_REFERENCE[self.goal[:, i]]
                o = np.concatenate([self.p_vel[:, i]] + ent + [color, self.comm[:, 1 - i]], axis=-1)
            elif sp.name == "simple_adversary":                         # simple_adversary.py:116-137
                other = [self.p_pos[:, j] - pos_i for j in range(M) if j != i]
                goal_pos = self.lm_pos[np.arange(N), self.goal[:, i]]
                o = np.concatenate(([goal_pos - pos_i] if i >= sp.n_adv else []) + ent + other, axis=-1)
            elif sp.name == "simple_push":                              # simple_push.py:81-104
                other = [self.p_pos[:, j] - pos_i for j in range(M) if j != i]
                if i < sp.n_adv:
                    o = np.concatenate([self.p_vel[:, i]] + ent + other, axis=-1)
                else:
                    g = self.goal[:, i]
                    goal_pos = self.lm_pos[np.arange(N), g]
                    color = np.full((N, 3), 0.25)                       # simple_push.py:48-54
                    color[np.arange(N), g + 1] += 0.5
                    lm_colors = []
                    for l in range(L):                                  # simple_push.py:41-44
                        c = np.full((N, 3), 0.1)
                        c[:, l + 1] += 0.8
                        lm_colors.append(c)
                    o = np.concatenate([self.p_vel[:, i], goal_pos - pos_i, color] + ent + lm_colors + other, axis=-1)
            elif sp.name == "simple_world_comm":                        # simple_world_comm.py:214-288
                f1, f2 = L - 2, L - 1                                   # the two forests close world.landmarks
                def in_f(a, f):
                    d = np.sqrt(np.sum(np.square(self.p_pos[:, a] - self.lm_pos[:, f]), axis=-1))
                    return d < sp.sizes[a] + sp.forest_size
                inf1, inf2 = in_f(i, f1), in_f(i, f2)
                in_forest = np.stack([np.where(inf1, 1.0, -1.0), np.where(inf2, 1.0, -1.0)], axis=-1)
                other, other_vel = [], []
                for j in range(M):
                    if j == i:
                        continue
                    o1, o2 = in_f(j, f1), in_f(j, f2)
                    vis = (inf1 & o1) | (inf2 & o2) | (~inf1 & ~o1 & ~inf2 & ~o2) | (i == 0)
                    other.append(np.where(vis[:, None], self.p_pos[:, j] - pos_i, 0.0))
                    if j >= sp.n_adv:
                        other_vel.append(np.where(vis[:, None], self.p_vel[:, j], 0.0))
                if i < sp.n_adv:
                    o = np.concatenate([self.p_vel[:, i], pos_i] + ent + other + other_vel + [in_forest, self.comm[:, 0]],
                                       axis=-1)
                else:
                    o = np.concatenate([self.p_vel[:, i], pos_i] + ent + other + [in_forest] + other_vel, axis=-1)
            elif sp.name == "simple_crypto":                            # simple_crypto.py:127-171
                goal_color = np.eye(4)[self.goal[:, 0]]                 # landmark i's colour is e_i (:54-58)
                key = np.eye(4)[self.goal[:, 1]]
                if i == 2:
                    o = np.concatenate([goal_color, key], axis=-1)      # speaker
                elif i == 1:
                    o = np.concatenate([key, self.comm[:, 2]], axis=-1)  # good listener
                else:
                    o = self.comm[:, 2].copy()                          # adversary: the speaker's utterance only
            else:                                                       # simple_tag.py:125-144
                other = [self.p_pos[:, j] - pos_i for j in range(M) if j != i]
                other_vel = [self.p_vel[:, j] for j in range(M) if j != i and j >= sp.n_adv]
                o = np.concatenate([self.p_vel[:, i], pos_i] + ent + other + other_vel, axis=-1)
            out.append(o)
        return out

    def rewards(self):
        """Per-agent (individual) reward, [N, M] f64, before the collaborative sum."""
        sp, N, M, L = self.spec, self.N, self.M, self.L
        ar = np.arange(N)
        r = np.zeros((N, M))
        if sp.name == "simple_spread":                                  # simple_spread.py:72-84
            base = np.zeros(N)
            for l in range(L):
                d = np.stack([np.sqrt(np.sum(np.square(self.p_pos[:, a] - self.lm_pos[:, l]), axis=-1))
                              for a in range(M)], axis=0)
                base = base - d.min(axis=0)
            dist_min = sp.agent_size + sp.agent_size
            for i in range(M):
                ri = base.copy()
                for a in range(M):
                    delta = self.p_pos[:, a] - self.p_pos[:, i]
                    dist = np.sqrt(np.sum(np.square(delta), axis=-1))
                    ri = ri - (dist < dist_min).astype(np.float64)
                r[:, i] = ri
        elif sp.name == "simple_speaker_listener":                      # simple_speaker_listener.py:69-73
            d = self.p_pos[:, 1] - self.lm_pos[ar, self.goal[:, 0]]
            r[:] = -np.sum(np.square(d), axis=-1)[:, None]
        elif sp.name == "simple_reference":                             # simple_reference.py:62-67
            for i in range(M):
                d = self.p_pos[:, 1 - i] - self.lm_pos[ar, self.goal[:, i]]
                r[:, i] = -np.sum(np.square(d), axis=-1)
        elif sp.name in ("simple_adversary", "simple_push"):
            goal_pos = self.lm_pos[ar, self.goal[:, 0]]
            dist = [np.sqrt(np.sum(np.square(self.p_pos[:, a] - goal_pos), axis=-1)) for a in range(M)]
            good_min = np.stack(dist[sp.n_adv:], axis=0).min(axis=0)
            for i in range(M):
                if sp.name == "simple_adversary":                       # simple_adversary.py:69-113
                    if i < sp.n_adv:
                        r[:, i] = -np.sum(np.square(self.p_pos[:, i] - goal_pos), axis=-1)
                    else:
                        adv_rew = 0
                        for a in range(sp.n_adv):
                            adv_rew = adv_rew + dist[a]
                        r[:, i] = -good_min + adv_rew
                else:                                                   # simple_push.py:62-78
                    if i < sp.n_adv:
                        neg = np.sqrt(np.sum(np.square(goal_pos - self.p_pos[:, i]), axis=-1))
                        r[:, i] = good_min - neg
                    else:
                        r[:, i] = -dist[i]
        elif sp.name == "simple_world_comm":                            # simple_world_comm.py:135-190
            def dist(pa, pb):
                return np.sqrt(np.sum(np.square(pa - pb), axis=-1))
            food = [L - 4, L - 3]
            for i in range(M):
                if i < sp.n_adv:
                    near = np.stack([dist(self.p_pos[:, a], self.p_pos[:, i]) for a in range(sp.n_adv, M)], 0).min(axis=0)
                    ri = 0.0 - 0.1 * near
                    for ag in range(sp.n_adv, M):                       # += 5 one pair at a time, in this order
                        for adv in range(sp.n_adv):
                            ri = ri + 5.0 * (dist(self.p_pos[:, ag], self.p_pos[:, adv]) < sp.sizes[ag] + sp.sizes[adv])
                    r[:, i] = ri
                else:
                    ri = np.zeros(N)
                    for a in range(sp.n_adv):
                        ri = ri - 5.0 * (dist(self.p_pos[:, a], self.p_pos[:, i]) < sp.sizes[a] + sp.sizes[i])
                    for p in range(2):
                        x = np.abs(self.p_pos[:, i, p])
                        with np.errstate(over="ignore"):
                            b = np.where(x < 0.9, 0.0, np.where(x < 1.0, (x - 0.9) * 10, np.minimum(np.exp(2 * x - 2), 10)))
                        ri = ri - 2 * b
                    fd = [dist(self.lm_pos[:, f], self.p_pos[:, i]) for f in food]
                    for k, f in enumerate(food):
                        ri = ri + 2.0 * (dist(self.p_pos[:, i], self.lm_pos[:, f]) < sp.sizes[i] + sp.food_size)
                    ri = ri + 0.05 * np.minimum(fd[0], fd[1])
                    r[:, i] = ri
        elif sp.name == "simple_crypto":                                # simple_crypto.py:90-125
            goal_color = np.eye(4)[self.goal[:, 0]]
            def sq_err(a):                                              # 0 while the agent has not spoken (:103,:122)
                spoke = (self.comm[:, a] != 0).any(axis=-1)
                return np.where(spoke, np.sum(np.square(self.comm[:, a] - goal_color), axis=-1), 0.0)
            adv_err, good_err = sq_err(0), sq_err(1)
            r[:, 0] = 0.0 - adv_err
            r[:, 1] = adv_err + (0.0 - good_err)
            r[:, 2] = adv_err + (0.0 - good_err)
        else:                                                           # simple_tag.py:73-123
            def hit(a, b):
                d = np.sqrt(np.sum(np.square(self.p_pos[:, a] - self.p_pos[:, b]), axis=-1))
                return d < sp.sizes[a] + sp.sizes[b]
            adv_total = np.zeros(N)
            for ag in range(sp.n_adv, M):
                for adv in range(sp.n_adv):
                    adv_total = adv_total + 10.0 * hit(ag, adv)
            for i in range(M):
                if i < sp.n_adv:
                    r[:, i] = adv_total
                else:
                    ri = np.zeros(N)
                    for a in range(sp.n_adv):
                        ri = ri - 10.0 * hit(a, i)
                    for p in range(2):
                        x = np.abs(self.p_pos[:, i, p])
                        with np.errstate(over="ignore"):
                            b = np.where(x < 0.9, 0.0, np.where(x < 1.0, (x - 0.9) * 10, np.minimum(np.exp(2 * x - 2), 10)))
                        ri = ri - b
                    r[:, i] = ri
        return r
