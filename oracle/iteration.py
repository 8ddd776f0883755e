"""ORACLE (test infrastructure, not product code): one whole MAPPO iteration on the CPU, assembled from the
oracle pieces -- the reference's driver loop restated over the batched oracle env:

  MPERunner.run / warmup / collect / insert     onpolicy/runner/shared/mpe_runner.py:16-139
  Runner.compute / train                        onpolicy/runner/shared/base_runner.py:110-125
  R_MAPPO.train / ppo_update                    onpolicy/algorithms/r_mappo/r_mappo.py:91-219
  SharedReplayBuffer.insert / after_update      onpolicy/utils/shared_buffer.py:79-159

Shared MLP policy (BASELINE configs C1 / C2 / C5).  Env maths: oracle/mpe.py (numpy f64); networks, losses, clipping
and Adam: oracle/policy.py (torch CPU f32 + autograd); returns / advantages / ValueNorm: oracle/buffer.py.
Used by tests/ (trajectory-level parity of the CUDA path) and by bench.py's `cpu_baseline` / `--impl reference`
legs, where it is the thing timed on the host cores ("port" of the reference's CPU path).  Never imported by
the product package.
"""
import numpy as np
import torch

from . import buffer as ob
from . import policy as op
from .mpe import OracleMPE


class OracleMAPPO:
    def __init__(self, args, actor_sd, critic_sd, num_envs=None, seed=None, env_id_offset=0, sample_seed=None):
        """args: the reference's flag namespace.  actor_sd / critic_sd: reference-keyed state dicts (CPU tensors);
        they are cloned and trained in place of nn.Modules."""
        self.a = args
        self.N = int(num_envs if num_envs is not None else args.n_rollout_threads)
        self.M, self.L, self.T = args.num_agents, args.num_landmarks, args.episode_length
        seed = int(args.seed if seed is None else seed)
        self.sample_seed = int(seed if sample_seed is None else sample_seed)
        self.env = OracleMPE(args.scenario_name, self.N, self.M, self.L, self.T, seed=seed,
                             env_id_offset=env_id_offset)
        self.row_offset = env_id_offset * self.M
        self.actor = {k: v.detach().clone().float() for k, v in actor_sd.items()}
        self.critic = {k: v.detach().clone().float() for k, v in critic_sd.items()}
        self.akeys = [k for k in self.actor if "fc_h" not in k]
        self.ckeys = [k for k in self.critic if "fc_h" not in k]
        self.adam = {"actor": {}, "critic": {}}
        self.adam_step = {"actor": 0, "critic": 0}
        self.vn = ob.OracleValueNorm() if args.use_valuenorm else None
        N, M, T = self.N, self.M, self.T
        D = self.env.spec.obs_dims[0]
        self.D, self.S = D, D * M
        f = np.float32
        A = len(op.head_keys(self.actor))
        self.rec = bool(args.use_recurrent_policy or args.use_naive_recurrent_policy)
        self.obs = np.zeros((T + 1, N, M, D), f)
        self.value_preds = np.zeros((T + 1, N, M, 1), f)
        self.returns = np.zeros((T + 1, N, M, 1), f)
        self.masks = np.ones((T + 1, N, M, 1), f)
        self.active_masks = np.ones((T + 1, N, M, 1), f)
        self.actions = np.zeros((T, N, M, A), f)
        self.action_log_probs = np.zeros((T, N, M, A), f)
        self.rewards = np.zeros((T, N, M, 1), f)
        self.rnn_states = np.zeros((T + 1, N, M, args.recurrent_N, args.hidden_size), f)
        self.rnn_states_critic = np.zeros((T + 1, N, M, args.recurrent_N, args.hidden_size), f)
        self.calls = 0

    # ------------------------------------------------------------------ rollout
    def _stack_obs(self, obs_list):
        return np.stack(obs_list, axis=1).astype(np.float32)            # [N, M, D]

    def share_obs_of(self, obs):
        """mpe_runner.py:133-137: concat of all agents' obs, repeated for every agent -> [N, M, S]; the agent's own
        obs with a decentralised critic (ippo)"""
        N, M = obs.shape[0], obs.shape[1]
        if not self.a.use_centralized_V:                                 # mpe_runner.py:136-137: share_obs = obs
            return obs
        return np.repeat(obs.reshape(N, 1, -1), M, axis=1)

    def share_at(self, t):
        """share_obs slot t as [N, M, S]"""
        return self.share_obs_of(self.obs[t])

    def share_rows(self):
        """share_obs[:-1] flattened to [T*N*M, S]"""
        return np.stack([self.share_at(t) for t in range(self.T)]).reshape(self.T * self.N * self.M, -1)

    def warmup(self, state=None):
        self.obs[0] = self._stack_obs(self.env.reset(state))

    @torch.no_grad()
    def collect(self, step, given_actions=None):
        """given_actions [N, M, heads]: evaluate those instead of sampling (replay of a recorded reference rollout)"""
        a = self.a
        N, M = self.N, self.M
        obs = self.obs[step].reshape(N * M, -1)
        sh = self.share_at(step).reshape(N * M, -1)
        RN = self.rnn_states.shape[3]
        z = self.rnn_states[step].reshape(N * M, RN, -1)
        zc = self.rnn_states_critic[step].reshape(N * M, RN, -1)
        mk = self.masks[step].reshape(N * M, 1)
        feats, h = op.actor_features(self.actor, obs, z, mk, a.use_ReLU, a.use_feature_normalization, self.rec)
        logits = op.actor_logits(self.actor, feats, None)
        if given_actions is not None:
            acts = torch.as_tensor(np.asarray(given_actions).reshape(N * M, -1)).long()
            lps, _ = op.evaluate_actions(logits, acts)
        else:
            acts, lps, _ = op.act_forward(logits, False, seed=self.sample_seed, step=self.calls,
                                          row_offset=self.row_offset)
        self.calls += 1
        values, hc = op.critic_values(self.critic, sh, zc, mk, a.use_ReLU, a.use_feature_normalization, self.rec)
        self._h = (h.numpy().reshape(N, M, RN, -1), hc.numpy().reshape(N, M, RN, -1))
        return values.numpy().reshape(N, M, 1), acts.numpy().reshape(N, M, -1), lps.numpy().reshape(N, M, -1)

    def rollout(self, given_actions=None, reset_states=None):
        """given_actions [T, N, M, heads] / reset_states {step: state dict}: replay a recorded rollout"""
        for step in range(self.T):
            values, actions, logp = self.collect(step, None if given_actions is None else given_actions[step])
            obs, rewards, dones, _ = self.env.step(actions, reset_state=None if reset_states is None
                                                   else reset_states.get(step))
            self.obs[step + 1] = self._stack_obs(obs)
            self.actions[step] = actions
            self.action_log_probs[step] = logp
            self.value_preds[step] = values
            self.rewards[step] = rewards
            m = np.ones((self.N, self.M, 1), np.float32)
            m[dones] = 0.0
            self.masks[step + 1] = m
            if self.rec:                                                 # mpe_runner.py:128-129
                h, hc = self._h
                h = h.copy(); hc = hc.copy()
                h[dones] = 0.0; hc[dones] = 0.0
                self.rnn_states[step + 1], self.rnn_states_critic[step + 1] = h, hc

    # ------------------------------------------------------------------ compute + train
    @torch.no_grad()
    def compute(self):
        a = self.a
        N, M = self.N, self.M
        sh = self.share_at(-1).reshape(N * M, -1)
        z = self.rnn_states_critic[-1].reshape(N * M, self.rnn_states_critic.shape[3], -1)
        nv, _ = op.critic_values(self.critic, sh, z, self.masks[-1].reshape(N * M, 1), a.use_ReLU,
                                 a.use_feature_normalization, self.rec)
        vp, ret = ob.compute_returns(self.rewards, self.value_preds, self.masks, nv.numpy().reshape(N, M, 1), a.gamma,
                                     a.gae_lambda, a.use_gae, a.use_proper_time_limits, None, self.vn)
        self.value_preds, self.returns = vp, ret

    def _step_net(self, name, sd, keys, grads, lr):
        a = self.a
        if a.use_max_grad_norm:
            norm, grads = op.clip_grad_norm(grads, a.max_grad_norm)
        else:
            norm = torch.sqrt(sum((g ** 2).sum() for g in grads))
        self.adam_step[name] += 1
        st = self.adam[name]
        for k, g in zip(keys, grads):
            m, v = st.get(k, (torch.zeros_like(g), torch.zeros_like(g)))
            p, m, v = op.adam_step(sd[k].detach(), g, m, v, self.adam_step[name], lr, a.opti_eps,
                                   weight_decay=a.weight_decay)
            sd[k] = p
            st[k] = (m, v)
        return float(norm)

    def ppo_update(self, rows, adv, rnn_rows=None):
        """r_mappo.py:91-164 on the flat rows `rows` (None = all, in storage order).  Recurrent: rows are the
        step-major [L * Nc] rows of a chunked minibatch and rnn_rows [Nc] the rows that seed each chunk."""
        a = self.a
        T, N, M = self.T, self.N, self.M
        B = T * N * M
        take = (lambda x: x) if rows is None else (lambda x: x[rows])
        obs = take(self.obs[:-1].reshape(B, -1))
        sh = take(self.share_rows())
        actions = take(self.actions.reshape(B, -1))
        old_lp = take(self.action_log_probs.reshape(B, -1))
        advb = take(adv.reshape(B, 1))
        vp = take(self.value_preds[:-1].reshape(B, 1))
        ret = take(self.returns[:-1].reshape(B, 1))
        am = take(self.active_masks[:-1].reshape(B, 1))
        nb = obs.shape[0]
        z = zc = np.zeros((nb, 1, a.hidden_size), np.float32)
        ones = np.ones((nb, 1), np.float32)
        if self.rec:
            RN = self.rnn_states.shape[3]
            z = self.rnn_states[:-1].reshape(B, RN, -1)[rnn_rows]
            zc = self.rnn_states_critic[:-1].reshape(B, RN, -1)[rnn_rows]
            ones = take(self.masks[:-1].reshape(B, 1))
        asd = {k: v.requires_grad_("fc_h" not in k) for k, v in self.actor.items()}
        csd = {k: v.requires_grad_("fc_h" not in k) for k, v in self.critic.items()}
        feats, _ = op.actor_features(asd, obs, z, ones, a.use_ReLU, a.use_feature_normalization, self.rec)
        logits = op.actor_logits(asd, feats, None)
        logp, ent = op.evaluate_actions(logits, actions, am if a.use_policy_active_masks else None)
        values, _ = op.critic_values(csd, sh, zc, ones, a.use_ReLU, a.use_feature_normalization, self.rec)
        pl, ratio = op.policy_loss(logp, old_lp, advb, am, a.clip_param, a.use_policy_active_masks)
        agrads = torch.autograd.grad(pl - ent * a.entropy_coef, [asd[k] for k in self.akeys])
        an = self._step_net("actor", self.actor, self.akeys, agrads, a.lr)
        vl = op.value_loss(values, vp, ret, am, a.clip_param, self.vn, a.use_huber_loss, a.huber_delta,
                           a.use_clipped_value_loss, a.use_value_active_masks)
        cgrads = torch.autograd.grad(vl * a.value_loss_coef, [csd[k] for k in self.ckeys])
        cn = self._step_net("critic", self.critic, self.ckeys, cgrads, a.critic_lr)
        self.actor = {k: v.detach() for k, v in self.actor.items()}
        self.critic = {k: v.detach() for k, v in self.critic.items()}
        return dict(value_loss=vl.item(), critic_grad_norm=cn, policy_loss=pl.item(), dist_entropy=ent.item(),
                    actor_grad_norm=an, ratio=ratio.mean().item())

    def recurrent_batches(self, perm, L):
        """shared_buffer.py:385-494: chunks cut from the [N, M, T]-ordered stream."""
        return ob.recurrent_rows(self.T, self.N, self.M, perm, self.a.num_mini_batch, L)

    def train(self, permute=False):
        """r_mappo.py:166-219.  permute=False: num_mini_batch == 1 in storage order (the loss is a mean; the
        reference's randperm only reorders it); permute=True draws torch.randperm like the reference."""
        a = self.a
        _, adv = ob.advantages(self.returns, self.value_preds, self.active_masks, self.vn)
        B = self.T * self.N * self.M
        info = None
        n = 0
        for _ in range(a.ppo_epoch):
            if a.use_recurrent_policy:
                L = a.data_chunk_length
                batches = self.recurrent_batches(torch.randperm(B // L).numpy(), L)
            elif a.use_naive_recurrent_policy:
                batches = ob.naive_recurrent_rows(self.T, self.N * self.M, torch.randperm(self.N * self.M).numpy(),
                                                  a.num_mini_batch)
            elif a.num_mini_batch == 1 and not permute:
                batches = [None]
            else:
                batches = ob.feed_forward_rows(self.T, self.N, self.M, torch.randperm(B).numpy(), a.num_mini_batch)
            for rows in batches:
                rnn_rows = None
                if isinstance(rows, tuple):
                    rows, rnn_rows = rows
                r = self.ppo_update(rows, adv, rnn_rows)
                info = r if info is None else {k: info[k] + r[k] for k in r}
                n += 1
        self.after_update()
        return {k: v / n for k, v in info.items()}

    def after_update(self):
        """shared_buffer.py:149-159"""
        self.obs[0] = self.obs[-1]
        self.masks[0] = self.masks[-1]
        self.active_masks[0] = self.active_masks[-1]
        self.rnn_states[0] = self.rnn_states[-1]
        self.rnn_states_critic[0] = self.rnn_states_critic[-1]

    def iteration(self):
        self.rollout()
        self.compute()
        return self.train()


class _AgentLearner(OracleMAPPO):
    """One agent of the separated setting: its own networks, Adam / ValueNorm state and [T+1, N, 1, ...] storage
    (separated_buffer.py:13-70 with the agent axis kept at size 1 so the shared maths above applies unchanged)."""

    def __init__(self, args, actor_sd, critic_sd, N, D, S, heads, sample_seed, row_offset):
        self.a = args
        self.N, self.M, self.T = N, 1, args.episode_length
        self.sample_seed, self.row_offset = int(sample_seed), int(row_offset)
        self.actor = {k: v.detach().clone().float() for k, v in actor_sd.items()}
        self.critic = {k: v.detach().clone().float() for k, v in critic_sd.items()}
        self.akeys = [k for k in self.actor if "fc_h" not in k]
        self.ckeys = [k for k in self.critic if "fc_h" not in k]
        self.adam = {"actor": {}, "critic": {}}
        self.adam_step = {"actor": 0, "critic": 0}
        self.vn = ob.OracleValueNorm() if args.use_valuenorm else None
        T = self.T
        f = np.float32
        self.D, self.S = D, S
        self.obs = np.zeros((T + 1, N, 1, D), f)
        self.share = np.zeros((T + 1, N, 1, S), f)
        self.value_preds = np.zeros((T + 1, N, 1, 1), f)
        self.returns = np.zeros((T + 1, N, 1, 1), f)
        self.masks = np.ones((T + 1, N, 1, 1), f)
        self.active_masks = np.ones((T + 1, N, 1, 1), f)
        self.actions = np.zeros((T, N, 1, heads), f)
        self.action_log_probs = np.zeros((T, N, 1, heads), f)
        self.rewards = np.zeros((T, N, 1, 1), f)
        self.rec = bool(args.use_recurrent_policy or args.use_naive_recurrent_policy)
        self.rnn_states = np.zeros((T + 1, N, 1, args.recurrent_N, args.hidden_size), f)
        self.rnn_states_critic = np.zeros((T + 1, N, 1, args.recurrent_N, args.hidden_size), f)
        self.calls = 0

    def share_at(self, t):
        return self.share[t]

    def recurrent_batches(self, perm, L):
        """separated_buffer.py:289-393, including the (Nc, L) stack flattened as (L, Nc)."""
        return ob.separated_recurrent_rows(self.T, self.N, perm, self.a.num_mini_batch, L)

    def share_rows(self):
        return self.share[:-1].reshape(self.T * self.N, -1)

    def after_update(self):
        """separated_buffer.py:101-110"""
        OracleMAPPO.after_update(self)
        self.share[0] = self.share[-1]


class OracleSeparatedMAPPO:
    """Separated policies (BASELINE config C3): onpolicy/runner/separated/mpe_runner.py:16-176 and
    runner/separated/base_runner.py:110-130 restated over the batched oracle env, one _AgentLearner per agent."""

    def __init__(self, args, actor_sds, critic_sds, num_envs=None, seed=None, env_id_offset=0, sample_seeds=None):
        self.a = args
        self.N = int(num_envs if num_envs is not None else args.n_rollout_threads)
        self.M, self.T = args.num_agents, args.episode_length
        seed = int(args.seed if seed is None else seed)
        self.env = OracleMPE(args.scenario_name, self.N, self.M, args.num_landmarks, self.T, seed=seed,
                             env_id_offset=env_id_offset)
        dims = self.env.spec.obs_dims
        W = int(sum(dims))
        self.agents = []
        for m in range(self.M):
            S = W if args.use_centralized_V else dims[m]
            heads = sum(1 for k in actor_sds[m] if k.endswith("linear.weight"))
            ss = seed if sample_seeds is None else sample_seeds[m]
            self.agents.append(_AgentLearner(args, actor_sds[m], critic_sds[m], self.N, dims[m], S, heads, ss,
                                             env_id_offset))
        self.heads = max(ag.actions.shape[-1] for ag in self.agents)

    def _store_obs(self, t, obs_list):
        """mpe_runner.py:86-96,160-176: agent m sees its own obs and (centralised) the concatenation of all."""
        joint = np.concatenate([np.asarray(o, np.float32) for o in obs_list], axis=1)
        for m, ag in enumerate(self.agents):
            ag.obs[t, :, 0] = obs_list[m]
            ag.share[t, :, 0] = joint if self.a.use_centralized_V else obs_list[m]

    def warmup(self, state=None):
        self._store_obs(0, self.env.reset(state))

    def rollout(self):
        N, M = self.N, self.M
        for step in range(self.T):
            acts_env = np.zeros((N, M, self.heads), np.int64)
            outs = []
            for m, ag in enumerate(self.agents):
                v, a, lp = ag.collect(step)
                outs.append((v, a, lp))
                acts_env[:, m, :a.shape[-1]] = a[:, 0]
            obs, rewards, dones, _ = self.env.step(acts_env)
            self._store_obs(step + 1, obs)
            for m, ag in enumerate(self.agents):
                v, a, lp = outs[m]
                ag.actions[step], ag.action_log_probs[step], ag.value_preds[step] = a, lp, v
                ag.rewards[step, :, 0] = rewards[:, m]
                mk = np.ones((N, 1), np.float32)
                mk[dones[:, m]] = 0.0
                ag.masks[step + 1, :, 0] = mk
                if ag.rec:                                               # separated/mpe_runner.py:145-146
                    h, hc = ag._h
                    h = h.copy(); hc = hc.copy()
                    h[dones[:, m]] = 0.0; hc[dones[:, m]] = 0.0
                    ag.rnn_states[step + 1], ag.rnn_states_critic[step + 1] = h, hc

    def compute(self):
        for ag in self.agents:
            ag.compute()

    def train(self, permute=False):
        return [ag.train(permute) for ag in self.agents]

    def iteration(self):
        self.rollout()
        self.compute()
        return self.train()
