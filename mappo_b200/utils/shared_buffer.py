"""SharedReplayBuffer mirror (onpolicy/utils/shared_buffer.py:14-494), device-resident.

Same constructor, public array attributes (shapes and dtypes of shared_buffer.py:45-75) and methods.  Differences
that are layout only:
  * arrays are CUDA tensors; the env and policy kernels write straight into their slots;
  * `share_obs` [T+1,N,M,S]: for homogeneous agents with a centralised critic it is the reference's
    M-times replicated concatenation of all agents' obs (runner/shared/mpe_runner.py:133-137), so it is kept
    as a stride-0 VIEW of `obs` (one copy per env) -- same values at every index;
  * `rnn_states{,_critic}` are stride-0 zero views unless the policy is recurrent;
  * `available_actions` (all ones for Discrete, shared_buffer.py:58-60) is a stride-0 ones view until written.
"""
import numpy as np
import torch

from .. import ops
from .. import gather as _gather


def _shape(space):
    if space.__class__.__name__ == "Box":
        return tuple(space.shape)
    if space.__class__.__name__ == "list":
        return tuple(space)
    raise NotImplementedError


def _act_shape(space):
    name = space.__class__.__name__
    if name == "Discrete":
        return 1
    if name == "MultiDiscrete":
        return space.shape
    raise NotImplementedError(name)


def _dev(x, device):
    """Host array / tensor -> float32 device tensor.  An array the library itself produced from a device tensor (the
    vec env's host outputs, host_mirror.t2n) is served by its device twin instead of a PCIe copy."""
    from .. import host_mirror
    twin = host_mirror.lookup(x)
    if twin is not None and twin.device == torch.device(device):
        return twin if twin.dtype == torch.float32 else twin.to(torch.float32)
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    return x.to(device=device, dtype=torch.float32, non_blocking=True)


class SharedReplayBuffer(object):
    def __init__(self, args, num_agents, obs_space, cent_obs_space, act_space, device="cuda"):
        self.episode_length = args.episode_length
        self.n_rollout_threads = args.n_rollout_threads
        self.hidden_size = args.hidden_size
        self.recurrent_N = args.recurrent_N
        self.gamma = args.gamma
        self.gae_lambda = args.gae_lambda
        self._use_gae = args.use_gae
        self._use_popart = args.use_popart
        self._use_valuenorm = args.use_valuenorm
        self._use_proper_time_limits = args.use_proper_time_limits
        self.num_agents = num_agents
        self.device = torch.device(device)
        T, N, M, H = self.episode_length, self.n_rollout_threads, num_agents, self.hidden_size
        D, S = _shape(obs_space)[0], _shape(cent_obs_space)[0]
        dev = self.device
        f = dict(dtype=torch.float32, device=dev)
        self.obs = torch.zeros(T + 1, N, M, D, **f)
        self._share_is_view = (S == M * D) and bool(getattr(args, "use_centralized_V", True))
        if self._share_is_view:
            self._share_unique = self.obs.view(T + 1, N, S)
            self.share_obs = self._share_unique.unsqueeze(2).expand(T + 1, N, M, S)
        elif S == D and not bool(getattr(args, "use_centralized_V", True)):
            # decentralised critic (ippo): share_obs IS obs (mpe_runner.py:136-137), kept as the same storage
            self._share_unique = None
            self.share_obs = self.obs
        else:
            self._share_unique = None
            self.share_obs = torch.zeros(T + 1, N, M, S, **f)
        recurrent = bool(args.use_recurrent_policy or args.use_naive_recurrent_policy)
        if recurrent:
            self.rnn_states = torch.zeros(T + 1, N, M, self.recurrent_N, H, **f)
            self.rnn_states_critic = torch.zeros_like(self.rnn_states)
        else:
            z = torch.zeros(1, 1, 1, 1, 1, **f)
            self.rnn_states = z.expand(T + 1, N, M, self.recurrent_N, H)
            self.rnn_states_critic = z.expand(T + 1, N, M, self.recurrent_N, H)
        self._recurrent = recurrent
        self.value_preds = torch.zeros(T + 1, N, M, 1, **f)
        self.returns = torch.zeros_like(self.value_preds)
        if act_space.__class__.__name__ == "Discrete":
            self.available_actions = torch.ones(1, 1, 1, 1, **f).expand(T + 1, N, M, act_space.n)
            self._avail_all_ones = True
        else:
            self.available_actions = None
            self._avail_all_ones = False
        a = _act_shape(act_space)
        self.actions = torch.zeros(T, N, M, a, **f)
        self.action_log_probs = torch.zeros(T, N, M, a, **f)
        self.rewards = torch.zeros(T, N, M, 1, **f)
        self.masks = torch.ones(T + 1, N, M, 1, **f)
        self.bad_masks = torch.ones_like(self.masks)
        self.active_masks = torch.ones_like(self.masks)
        self.step = 0
        self.force_permutation = False     # True: draw torch.randperm even when num_mini_batch == 1
        self._adv = None
        self.all_reduce = None

    # ------------------------------------------------------------------ helpers for the fused path
    def share_obs_unique(self):
        return self._share_unique

    def _adv_scratch(self):
        if self._adv is None:
            self._adv = torch.empty_like(self.rewards)
        return self._adv

    def flat_views(self):
        """2-D row views for the update kernels.  share_is_unique: `share` holds one row per (t, env) serving all M
        agents of that env (the stride-0 view above), else one row per sample."""
        T, N, M = self.rewards.shape[0:3]
        B = T * N * M
        avail = None
        if self.available_actions is not None and not self._avail_all_ones:
            avail = self.available_actions[:-1].reshape(B, -1)
        u = self._share_unique
        share = u[:-1].view(T * N, -1) if u is not None else self.share_obs[:-1].view(B, -1)
        return dict(B=B, M=M, obs=self.obs[:-1].view(B, -1), share=share, share_is_unique=u is not None,
                    actions=self.actions.view(B, -1), logp=self.action_log_probs.view(B, -1),
                    value_preds=self.value_preds[:-1].view(B), returns=self.returns[:-1].view(B),
                    active=self.active_masks[:-1].view(B), avail=avail, masks=self.masks[:-1].view(B),
                    rnn=self.rnn_states[:-1].reshape(B, -1) if self._recurrent else None,
                    rnn_critic=self.rnn_states_critic[:-1].reshape(B, -1) if self._recurrent else None)

    def recurrent_rows(self, num_mini_batch, data_chunk_length):
        """Index form of recurrent_generator (shared_buffer.py:385-494): per minibatch (rows int64 [L * Nc] step-major
        into the flat [T*N*M] sample arrays, Nc, L)."""
        T, N, M = self.rewards.shape[0:3]
        L = data_chunk_length
        data_chunks = (N * T * M) // L
        mini_batch_size = data_chunks // num_mini_batch
        dev = self.device
        if num_mini_batch == 1 and not self.force_permutation:
            # one minibatch = every chunk: the losses are means over it and each chunk carries its own initial state, so
            # the permutation (shared_buffer.py:397) only reorders the sums -- chunks are taken in storage order, the
            # index is built once, and nothing on the path depends on the host any more (CUDA-graph capture)
            key = (L, mini_batch_size)
            if getattr(self, "_chunk_rows_key", None) != key:
                idx = torch.arange(mini_batch_size, device=dev)
                pos = idx[None, :] * L + torch.arange(L, device=dev)[:, None]
                self._chunk_rows = ((pos % T) * (N * M) + pos // T).reshape(-1).contiguous()
                self._chunk_rows_key = key
            yield self._chunk_rows, mini_batch_size, L
            return
        rand = torch.randperm(data_chunks).to(self.device)
        for i in range(num_mini_batch):
            idx = rand[i * mini_batch_size:(i + 1) * mini_batch_size]
            pos = idx[None, :] * L + torch.arange(L, device=dev)[:, None]            # stream positions, [N, M, T] order
            nm, t = pos // T, pos % T
            yield (t * (N * M) + nm).reshape(-1).contiguous(), int(idx.numel()), L

    def naive_recurrent_rows(self, num_mini_batch):
        """Index form of naive_recurrent_generator (shared_buffer.py:288-383): whole trajectories."""
        T, N, M = self.rewards.shape[0:3]
        batch_size = N * M
        per = batch_size // num_mini_batch
        perm = torch.randperm(batch_size).to(self.device)
        tt = torch.arange(T, device=self.device)
        for start in range(0, batch_size, per):
            idx = perm[start:start + per]
            if idx.numel() < per:
                break
            yield (tt[:, None] * batch_size + idx[None, :]).reshape(-1).contiguous(), int(idx.numel()), T

    def _put(self, dst, src):
        """Copy unless `src` already IS the destination slot (the kernels wrote in place)."""
        if src is None:
            return
        if torch.is_tensor(src) and src.is_cuda and src.data_ptr() == dst.data_ptr() and src.numel() == dst.numel():
            return
        d = _dev(src, self.device)
        pend = getattr(self, "_pending", None)
        if pend is not None and d.dtype == dst.dtype == torch.float32 and d.is_contiguous() and dst.is_contiguous() \
                and d.numel() == dst.numel():
            pend.append((dst, d))               # inside insert(): batched into one launch (mappo_copy_segments)
        else:
            dst.copy_(d.reshape(dst.shape))

    def _materialise_avail(self):
        if self._avail_all_ones:
            self.available_actions = self.available_actions.clone()
            self._avail_all_ones = False

    # ------------------------------------------------------------------ reference API
    def insert(self, share_obs, obs, rnn_states_actor, rnn_states_critic, actions, action_log_probs,
               value_preds, rewards, masks, bad_masks=None, active_masks=None, available_actions=None):
        s = self.step
        self._pending = []                      # (dst, src) device copies of this insert: ONE launch at the end
        self._put(self.obs[s + 1], obs)
        if not self._share_is_view:
            self._put(self.share_obs[s + 1], share_obs)
        if self._recurrent:
            self._put(self.rnn_states[s + 1], rnn_states_actor)
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
            self._materialise_avail()
            self._put(self.available_actions[s + 1], available_actions)
        pending, self._pending = self._pending, None
        if pending:
            _gather.copy_segments(pending)
        self.step = (self.step + 1) % self.episode_length

    def after_update(self):
        """shared_buffer.py:149-159"""
        self.obs[0].copy_(self.obs[-1])
        if not self._share_is_view:
            self.share_obs[0].copy_(self.share_obs[-1])
        if self._recurrent:
            self.rnn_states[0].copy_(self.rnn_states[-1])
            self.rnn_states_critic[0].copy_(self.rnn_states_critic[-1])
        self.masks[0].copy_(self.masks[-1])
        self.bad_masks[0].copy_(self.bad_masks[-1])
        self.active_masks[0].copy_(self.active_masks[-1])
        if self.available_actions is not None and not self._avail_all_ones:
            self.available_actions[0].copy_(self.available_actions[-1])

    def compute_returns(self, next_value, value_normalizer=None):
        """shared_buffer.py:168-224: one scan kernel, one thread per (env, agent) trajectory."""
        nv = _dev(next_value, self.device).reshape(-1).contiguous()
        vn = None
        if (self._use_popart or self._use_valuenorm) and value_normalizer is not None:
            vn = value_normalizer.stats
        ops.gae_returns(self.rewards, self.value_preds, self.returns, self.masks, nv, self.gamma, self.gae_lambda,
                        use_gae=self._use_gae, use_proper_time_limits=self._use_proper_time_limits,
                        bad_masks=self.bad_masks, active_masks=self.active_masks, vn_stats=vn,
                        advantages_raw=self._adv_scratch())

    # ------------------------------------------------------------------ generators (API parity; the trainer fuses
    # the gather into its kernels instead of materialising minibatches)
    def _fields(self, advantages):
        T, N, M = self.rewards.shape[0:3]
        adv = _dev(advantages, self.device)
        return dict(share_obs=self.share_obs[:-1], obs=self.obs[:-1], rnn_states=self.rnn_states[:-1],
                    rnn_states_critic=self.rnn_states_critic[:-1], actions=self.actions,
                    value_preds=self.value_preds[:-1], returns=self.returns[:-1], masks=self.masks[:-1],
                    active_masks=self.active_masks[:-1], old_action_log_probs=self.action_log_probs, adv_targ=adv,
                    available_actions=None if self.available_actions is None else self.available_actions[:-1])

    ORDER = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "value_preds", "returns", "masks",
             "active_masks", "old_action_log_probs", "adv_targ", "available_actions"]

    def _yield(self, fields, rows, rnn_rows):
        out = []
        for name in self.ORDER:
            src = fields[name]
            if src is None:
                out.append(None)
                continue
            tail = src.shape[3:]
            r = rnn_rows if (name.startswith("rnn_states") and rnn_rows is not None) else rows
            out.append(_gather.gather_rows(src, 3, r).view(r.numel(), *tail))
        return tuple(out)

    def feed_forward_generator(self, advantages, num_mini_batch=None, mini_batch_size=None):
        """shared_buffer.py:226-286"""
        T, N, M = self.rewards.shape[0:3]
        batch_size = N * T * M
        if mini_batch_size is None:
            assert batch_size >= num_mini_batch, (
                "PPO requires the number of processes ({}) * number of steps ({}) * number of agents ({}) = {} "
                "to be greater than or equal to the number of PPO mini batches ({}).".format(
                    N, T, M, batch_size, num_mini_batch))
            mini_batch_size = batch_size // num_mini_batch
        rand = torch.randperm(batch_size).to(self.device)
        fields = self._fields(advantages)
        for i in range(num_mini_batch):
            yield self._yield(fields, rand[i * mini_batch_size:(i + 1) * mini_batch_size], None)

    def recurrent_generator(self, advantages, num_mini_batch, data_chunk_length):
        """shared_buffer.py:385-494 (chunks cut the [N,M,T]-ordered stream; they straddle trajectories when
        data_chunk_length does not divide episode_length -- reproduced, not fixed)."""
        T, N, M = self.rewards.shape[0:3]
        batch_size = N * T * M
        data_chunks = batch_size // data_chunk_length
        mini_batch_size = data_chunks // num_mini_batch
        rand = torch.randperm(data_chunks).to(self.device)
        L = data_chunk_length
        dev = self.device
        n, m, t = torch.meshgrid(torch.arange(N, device=dev), torch.arange(M, device=dev),
                                 torch.arange(T, device=dev), indexing="ij")
        stream = ((t * N + n) * M + m).reshape(-1)
        fields = self._fields(advantages)
        for i in range(num_mini_batch):
            idx = rand[i * mini_batch_size:(i + 1) * mini_batch_size]
            pos = idx[None, :] * L + torch.arange(L, device=dev)[:, None]
            yield self._yield(fields, stream[pos].reshape(-1), stream[idx * L])

    def naive_recurrent_generator(self, advantages, num_mini_batch):
        """shared_buffer.py:288-383"""
        T, N, M = self.rewards.shape[0:3]
        batch_size = N * M
        assert N * M >= num_mini_batch, (
            "PPO requires the number of processes ({})* number of agents ({}) to be greater than or equal to the "
            "number of PPO mini batches ({}).".format(N, M, num_mini_batch))
        per = batch_size // num_mini_batch
        perm = torch.randperm(batch_size).to(self.device)
        fields = self._fields(advantages)
        tt = torch.arange(T, device=self.device)
        for start in range(0, batch_size, per):
            idx = perm[start:start + per]
            if idx.numel() < per:
                break
            rows = (tt[:, None] * batch_size + idx[None, :]).reshape(-1)
            yield self._yield(fields, rows, idx)
