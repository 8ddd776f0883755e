"""SeparatedReplayBuffer mirror (onpolicy/utils/separated_buffer.py:13-393), device-resident: one buffer per agent,
arrays `[T+1, N, ...]` without the agent axis.

Layout-only differences from the reference:
  * arrays are CUDA tensors;
  * `obs` / `share_obs` may be VIEWS of a joint observation storage shared by all agents' buffers (`joint_obs`
    [T+1, N, W], `obs_offset`): the env kernel writes every agent's observation once, each agent's `obs` is the
    column slice [off, off + D) of it (row stride W) and, with a centralised critic, `share_obs` is the whole row
    (runner/separated/mpe_runner.py:86-96,160-176 builds exactly those arrays by copying);
  * `rnn_states*` are stride-0 zero views unless the policy is recurrent; `available_actions` (all ones for Discrete,
    separated_buffer.py:44-47) is a stride-0 ones view until written.
"""
import numpy as np
import torch

from .. import gather as _gather
from .. import ops
from .shared_buffer import _act_shape, _dev, _shape


class SeparatedReplayBuffer(object):
    def __init__(self, args, obs_space, share_obs_space, act_space, device="cuda", joint_obs=None, obs_offset=0):
        self.episode_length = args.episode_length
        self.n_rollout_threads = args.n_rollout_threads
        self.rnn_hidden_size = args.hidden_size
        self.recurrent_N = args.recurrent_N
        self.gamma = args.gamma
        self.gae_lambda = args.gae_lambda
        self._use_gae = args.use_gae
        self._use_popart = args.use_popart
        self._use_valuenorm = args.use_valuenorm
        self._use_proper_time_limits = args.use_proper_time_limits
        self.device = torch.device(device)
        self.num_agents = 1
        T, N, H = self.episode_length, self.n_rollout_threads, self.rnn_hidden_size
        D, S = _shape(obs_space)[0], _shape(share_obs_space)[0]
        f = dict(dtype=torch.float32, device=self.device)
        self._joint = joint_obs
        self._obs_offset = int(obs_offset)
        if joint_obs is not None:
            assert tuple(joint_obs.shape[:2]) == (T + 1, N) and joint_obs.is_contiguous()
            W = joint_obs.shape[2]
            self.obs = joint_obs[:, :, obs_offset:obs_offset + D]
            if S == W:
                self.share_obs = joint_obs
            elif S == D:
                self.share_obs = self.obs                  # decentralised critic (use_centralized_V = False)
            else:
                raise AssertionError("share_obs width %d is neither the joint width %d nor the obs width %d" % (S, W, D))
        else:
            self.obs = torch.zeros(T + 1, N, D, **f)
            self.share_obs = torch.zeros(T + 1, N, S, **f)
        recurrent = bool(args.use_recurrent_policy or args.use_naive_recurrent_policy)
        if recurrent:
            self.rnn_states = torch.zeros(T + 1, N, self.recurrent_N, H, **f)
            self.rnn_states_critic = torch.zeros_like(self.rnn_states)
        else:
            z = torch.zeros(1, 1, 1, 1, **f)
            self.rnn_states = z.expand(T + 1, N, self.recurrent_N, H)
            self.rnn_states_critic = z.expand(T + 1, N, self.recurrent_N, H)
        self._recurrent = recurrent
        self.value_preds = torch.zeros(T + 1, N, 1, **f)
        self.returns = torch.zeros(T + 1, N, 1, **f)
        if act_space.__class__.__name__ == "Discrete":
            self.available_actions = torch.ones(1, 1, 1, **f).expand(T + 1, N, act_space.n)
            self._avail_all_ones = True
        else:
            self.available_actions = None
            self._avail_all_ones = False
        a = _act_shape(act_space)
        self.actions = torch.zeros(T, N, a, **f)
        self.action_log_probs = torch.zeros(T, N, a, **f)
        self.rewards = torch.zeros(T, N, 1, **f)
        self.masks = torch.ones(T + 1, N, 1, **f)
        self.bad_masks = torch.ones_like(self.masks)
        self.active_masks = torch.ones_like(self.masks)
        self.step = 0
        self.force_permutation = False
        self._adv = None

    # ------------------------------------------------------------------ helpers for the fused path
    def share_obs_unique(self):
        return None

    def _adv_scratch(self):
        if self._adv is None:
            self._adv = torch.empty_like(self.rewards)
        return self._adv

    def flat_views(self):
        """2-D row views (unit inner stride; the row stride may exceed the width) for the update kernels."""
        T, N = self.rewards.shape[0:2]
        B = T * N
        if self._joint is not None:
            J = self._joint[:-1].view(B, self._joint.shape[2])
            D = self.obs.shape[2]
            obs = J[:, self._obs_offset:self._obs_offset + D]
            share = J if self.share_obs is self._joint else obs
        else:
            obs, share = self.obs[:-1].view(B, -1), self.share_obs[:-1].view(B, -1)
        avail = None
        if self.available_actions is not None and not self._avail_all_ones:
            avail = self.available_actions[:-1].reshape(B, -1)
        return dict(B=B, M=1, obs=obs, share=share, share_is_unique=False,
                    actions=self.actions.view(B, -1), logp=self.action_log_probs.view(B, -1),
                    value_preds=self.value_preds[:-1].view(B), returns=self.returns[:-1].view(B),
                    active=self.active_masks[:-1].view(B), avail=avail, masks=self.masks[:-1].view(B),
                    rnn=self.rnn_states[:-1].reshape(B, -1) if self._recurrent else None,
                    rnn_critic=self.rnn_states_critic[:-1].reshape(B, -1) if self._recurrent else None)

    def recurrent_rows(self, num_mini_batch, data_chunk_length):
        """Index form of recurrent_generator (separated_buffer.py:289-393) for the fused kernels: per minibatch
        (rows int64 [Nc * L], Nc, L, seed_rows int64 [Nc]).  The reference stacks a minibatch as (Nc, L) and flattens it
        as if it were (L, Nc) (see recurrent_generator below), so the sample the network sees at step s of slot c is
        flat element p = s * Nc + c of the (Nc, L) stack: chunk p // L, offset p % L -- `rows` is exactly that stack,
        flattened.  The initial state of slot c is still the one stored at the first row of chunk c: `seed_rows`."""
        T, N = self.rewards.shape[0:2]
        L = data_chunk_length
        data_chunks = (N * T) // L
        mini_batch_size = data_chunks // num_mini_batch
        rand = torch.randperm(data_chunks).to(self.device)
        dev = self.device
        ar = torch.arange(L, device=dev)
        for i in range(num_mini_batch):
            idx = rand[i * mini_batch_size:(i + 1) * mini_batch_size]
            pos = idx[:, None] * L + ar[None, :]                                   # [N, T]-ordered stream positions
            rows = ((pos % T) * N + pos // T).reshape(-1).contiguous()              # -> rows of the [T, N] arrays
            seed = ((idx * L) % T) * N + (idx * L) // T
            yield rows, int(idx.numel()), L, seed

    def naive_recurrent_rows(self, num_mini_batch):
        """Index form of naive_recurrent_generator (separated_buffer.py:215-287): whole trajectories, step-major."""
        T, N = self.rewards.shape[0:2]
        per = N // num_mini_batch
        perm = torch.randperm(N).to(self.device)
        tt = torch.arange(T, device=self.device)
        for start in range(0, N, per):
            idx = perm[start:start + per]
            if idx.numel() < per:
                break
            yield (tt[:, None] * N + idx[None, :]).reshape(-1).contiguous(), int(idx.numel()), T, None

    def _put(self, dst, src):
        if src is None:
            return
        if torch.is_tensor(src) and src.is_cuda and src.data_ptr() == dst.data_ptr() and src.stride() == dst.stride():
            return
        dst.copy_(_dev(src, self.device).reshape(dst.shape))

    # ------------------------------------------------------------------ reference API
    def insert(self, share_obs, obs, rnn_states, rnn_states_critic, actions, action_log_probs,
               value_preds, rewards, masks, bad_masks=None, active_masks=None, available_actions=None):
        s = self.step
        if self._joint is None:
            self._put(self.share_obs[s + 1], share_obs)
            self._put(self.obs[s + 1], obs)
        else:
            self._put(self.obs[s + 1], obs)        # a no-op when the env kernel wrote the joint row in place
        if self._recurrent:
            self._put(self.rnn_states[s + 1], rnn_states)
            self._put(self.rnn_states_critic[s + 1], rnn_states_critic)
        self._put(self.actions[s], actions)
        self._put(self.action_log_probs[s], action_log_probs)
        self._put(self.value_preds[s], value_preds)
        self._put(self.rewards[s], rewards)
        self._put(self.masks[s + 1], masks)
        if bad_masks is not None:
            self._put(self.bad_masks[s + 1], bad_masks)
        if active_masks is not None:
            self._put(self.active_masks[s + 1], active_masks)
        if available_actions is not None:
            if self._avail_all_ones:
                self.available_actions = self.available_actions.clone()
                self._avail_all_ones = False
            self._put(self.available_actions[s + 1], available_actions)
        self.step = (self.step + 1) % self.episode_length

    def after_update(self, copy_obs=True):
        """separated_buffer.py:101-110.  copy_obs=False: another agent's buffer already rolled the joint storage."""
        if copy_obs:
            if self._joint is not None:
                self._joint[0].copy_(self._joint[-1])
            else:
                self.share_obs[0].copy_(self.share_obs[-1])
                self.obs[0].copy_(self.obs[-1])
        if self._recurrent:
            self.rnn_states[0].copy_(self.rnn_states[-1])
            self.rnn_states_critic[0].copy_(self.rnn_states_critic[-1])
        self.masks[0].copy_(self.masks[-1])
        self.bad_masks[0].copy_(self.bad_masks[-1])
        self.active_masks[0].copy_(self.active_masks[-1])
        if self.available_actions is not None and not self._avail_all_ones:
            self.available_actions[0].copy_(self.available_actions[-1])

    def compute_returns(self, next_value, value_normalizer=None):
        """separated_buffer.py:118-160 (same four branches as the shared buffer), one scan kernel over N trajectories."""
        nv = _dev(next_value, self.device).reshape(-1).contiguous()
        vn = None
        if (self._use_popart or self._use_valuenorm) and value_normalizer is not None:
            vn = value_normalizer.stats
        ops.gae_returns(self.rewards, self.value_preds, self.returns, self.masks, nv, self.gamma, self.gae_lambda,
                        use_gae=self._use_gae, use_proper_time_limits=self._use_proper_time_limits,
                        bad_masks=self.bad_masks, active_masks=self.active_masks, vn_stats=vn,
                        advantages_raw=self._adv_scratch())

    # ------------------------------------------------------------------ generators
    ORDER = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "value_preds", "returns", "masks",
             "active_masks", "old_action_log_probs", "adv_targ", "available_actions"]

    def _fields(self, advantages):
        c = lambda x: x if x.is_contiguous() or any(s == 0 for s in x.stride()) else x.contiguous()
        adv = _dev(advantages, self.device)
        return dict(share_obs=c(self.share_obs[:-1]), obs=c(self.obs[:-1]), rnn_states=self.rnn_states[:-1],
                    rnn_states_critic=self.rnn_states_critic[:-1], actions=self.actions,
                    value_preds=self.value_preds[:-1], returns=self.returns[:-1], masks=self.masks[:-1],
                    active_masks=self.active_masks[:-1], old_action_log_probs=self.action_log_probs, adv_targ=adv,
                    available_actions=None if self.available_actions is None else self.available_actions[:-1])

    def _yield(self, fields, rows, rnn_rows):
        out = []
        for name in self.ORDER:
            src = fields[name]
            if src is None:
                out.append(None)
                continue
            tail = src.shape[2:]
            r = rnn_rows if (name.startswith("rnn_states") and rnn_rows is not None) else rows
            out.append(_gather.gather_rows(src, 2, r).view(r.numel(), *tail))
        return tuple(out)

    def feed_forward_generator(self, advantages, num_mini_batch=None, mini_batch_size=None):
        """separated_buffer.py:162-213"""
        T, N = self.rewards.shape[0:2]
        batch_size = N * T
        if mini_batch_size is None:
            assert batch_size >= num_mini_batch, (
                "PPO requires the number of processes ({}) * number of steps ({}) = {} "
                "to be greater than or equal to the number of PPO mini batches ({}).".format(N, T, N * T, num_mini_batch))
            mini_batch_size = batch_size // num_mini_batch
        rand = torch.randperm(batch_size).to(self.device)
        fields = self._fields(advantages)
        for i in range(num_mini_batch):
            yield self._yield(fields, rand[i * mini_batch_size:(i + 1) * mini_batch_size], None)

    def naive_recurrent_generator(self, advantages, num_mini_batch):
        """separated_buffer.py:215-287"""
        T, N = self.rewards.shape[0:2]
        assert N >= num_mini_batch, (
            "PPO requires the number of processes ({}) to be greater than or equal to the number of "
            "PPO mini batches ({}).".format(N, num_mini_batch))
        per = N // num_mini_batch
        perm = torch.randperm(N).to(self.device)
        fields = self._fields(advantages)
        tt = torch.arange(T, device=self.device)
        for start in range(0, N, per):
            idx = perm[start:start + per]
            if idx.numel() < per:
                break
            yield self._yield(fields, (tt[:, None] * N + idx[None, :]).reshape(-1), idx)

    def recurrent_generator(self, advantages, num_mini_batch, data_chunk_length):
        """separated_buffer.py:289-393: data reordered to [N, T], cut into chunks of L, a minibatch stacked as (Nc, L)
        and then flattened AS IF it were (L, Nc) -- reproduced as is."""
        T, N = self.rewards.shape[0:2]
        L = data_chunk_length
        data_chunks = (N * T) // L
        mini_batch_size = data_chunks // num_mini_batch
        rand = torch.randperm(data_chunks).to(self.device)
        dev = self.device
        n, t = torch.meshgrid(torch.arange(N, device=dev), torch.arange(T, device=dev), indexing="ij")
        stream = (t * N + n).reshape(-1)
        fields = self._fields(advantages)
        for i in range(num_mini_batch):
            idx = rand[i * mini_batch_size:(i + 1) * mini_batch_size]
            pos = idx[:, None] * L + torch.arange(L, device=dev)[None, :]
            yield self._yield(fields, stream[pos].reshape(-1), stream[idx * L])
