"""R_Actor / R_Critic mirrors (onpolicy/algorithms/r_mappo/algorithm/r_actor_critic.py:12-165).

Same constructor arguments, method names, return order and `state_dict()` keys as the reference, so
`actor.pt` / `critic.pt` checkpoints are interchangeable.  All trainable parameters of a network are
views into ONE flat float32 CUDA buffer (`.flat`, layout in include/mappo_b200.h section 3) which the
fused kernels read; there is no PyTorch forward path."""
import copy

import numpy as np
import torch
import torch.nn as nn

from .... import ops
from ....envs.spaces import head_sizes


def check(x, device):
    """algorithms/utils/util.py:15-17 plus the move to the device."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    return x.to(device=device, dtype=torch.float32, non_blocking=True)


class _Node(nn.Module):
    """Anonymous container used to reproduce the reference's module tree (and therefore its key names)."""


def _attach(root, dotted, param):
    parts = dotted.split(".")
    node = root
    for p in parts[:-1]:
        if p not in node._modules:
            node.add_module(p, _Node())
        node = node._modules[p]
    node.register_parameter(parts[-1], param)


def _reference_init(args, in_dim, head_dims, is_critic):
    """Initial values, drawn exactly as the reference draws them (same torch CPU RNG consumption order):
    MLPLayer.__init__ mlp.py:6-22 (fc1, then the fc_h template, fc2 = deep copies of fc_h),
    RNNLayer.__init__ rnn.py:7-22, Categorical.__init__ distributions.py:56-62 / v_out r_actor_critic.py:134-140.
    Returns an ordered {reference key: CPU tensor}."""
    H = args.hidden_size
    use_relu, orth = bool(args.use_ReLU), bool(args.use_orthogonal)
    w_init = nn.init.orthogonal_ if orth else nn.init.xavier_uniform_
    gain = nn.init.calculate_gain("relu" if use_relu else "tanh")
    out = {}

    def linear(i, o, g):
        lin = nn.Linear(i, o)
        w_init(lin.weight.data, gain=g)
        nn.init.constant_(lin.bias.data, 0)
        return lin

    if args.use_feature_normalization:
        out["base.feature_norm.weight"] = torch.ones(in_dim)
        out["base.feature_norm.bias"] = torch.zeros(in_dim)
    fc1 = linear(in_dim, H, gain)
    fc_h = linear(H, H, gain)
    out["base.mlp.fc1.0.weight"], out["base.mlp.fc1.0.bias"] = fc1.weight.data, fc1.bias.data
    out["base.mlp.fc1.2.weight"], out["base.mlp.fc1.2.bias"] = torch.ones(H), torch.zeros(H)
    out["base.mlp.fc_h.0.weight"], out["base.mlp.fc_h.0.bias"] = fc_h.weight.data, fc_h.bias.data
    out["base.mlp.fc_h.2.weight"], out["base.mlp.fc_h.2.bias"] = torch.ones(H), torch.zeros(H)
    for l in range(args.layer_N):
        out["base.mlp.fc2.%d.0.weight" % l] = fc_h.weight.data.clone()
        out["base.mlp.fc2.%d.0.bias" % l] = fc_h.bias.data.clone()
        out["base.mlp.fc2.%d.2.weight" % l], out["base.mlp.fc2.%d.2.bias" % l] = torch.ones(H), torch.zeros(H)
    if args.use_recurrent_policy or args.use_naive_recurrent_policy:
        gru = nn.GRU(H, H, num_layers=args.recurrent_N)
        for name, prm in gru.named_parameters():
            if "bias" in name:
                nn.init.constant_(prm, 0)
            elif orth:
                nn.init.orthogonal_(prm)
            else:
                nn.init.xavier_uniform_(prm)
        for name, prm in gru.named_parameters():
            out["rnn.rnn." + name] = prm.data
        out["rnn.norm.weight"], out["rnn.norm.bias"] = torch.ones(H), torch.zeros(H)
    if is_critic:
        v = linear(H, 1, 1)
        out["v_out.weight"], out["v_out.bias"] = v.weight.data, v.bias.data
    elif len(head_dims) == 1:
        a = linear(H, head_dims[0], args.gain)
        out["act.action_out.linear.weight"], out["act.action_out.linear.bias"] = a.weight.data, a.bias.data
    else:
        for i, d in enumerate(head_dims):
            a = linear(H, d, args.gain)
            out["act.action_outs.%d.linear.weight" % i] = a.weight.data
            out["act.action_outs.%d.linear.bias" % i] = a.bias.data
    return out


class _FlatNet(nn.Module):
    def _build(self, args, in_dim, head_dims, is_critic, device):
        self.hidden_size = args.hidden_size
        self._recurrent = bool(args.use_recurrent_policy or args.use_naive_recurrent_policy)
        self._recurrent_N = args.recurrent_N
        self.device_ = torch.device(device)
        # --use_popart (r_actor_critic.py:139-140): the value head is a PopArt whose tensors, on the reference's default
        # cuda device, are not registered parameters (algorithms/utils/popart.py here): kept in `.flat` for the kernels,
        # left out of named_parameters() / state_dict() and of the optimiser
        self._head_outside_optimizer = bool(args.use_popart) and is_critic
        if self._recurrent and not 1 <= args.recurrent_N <= 2:
            raise NotImplementedError("recurrent_N = %d (this build: 1 or 2 GRU layers)" % args.recurrent_N)
        self.cfg = ops.net_config(in_dim, args.hidden_size, args.layer_N, args.use_ReLU,
                                  args.use_feature_normalization, args.recurrent_N if self._recurrent else 0,
                                  head_dims, is_critic)
        init = _reference_init(args, in_dim, head_dims, is_critic)
        live = [k for k in init if "fc_h" not in k]
        total = sum(init[k].numel() for k in live)
        assert total == ops.net_param_count(self.cfg), (total, ops.net_param_count(self.cfg))
        flat = torch.empty(total, dtype=torch.float32, device=self.device_)
        self.register_buffer("flat", flat, persistent=False)
        off = 0
        self._slices, self._shapes = {}, {}
        for k, v in init.items():
            if "fc_h" in k:      # registered-but-never-executed template (mlp.py:20-27): kept for checkpoints
                p = nn.Parameter(v.to(self.device_), requires_grad=False)
            else:
                n = v.numel()
                flat[off:off + n].copy_(v.reshape(-1))
                p = nn.Parameter(flat[off:off + n].view(v.shape), requires_grad=False)
                self._slices[k] = (off, off + n)
                self._shapes[k] = tuple(v.shape)
                off += n
            if self._head_outside_optimizer and k.startswith("v_out."):
                continue         # PopArt's weight / bias: plain tensors on cuda, not parameters
            _attach(self, k, p)
        assert off == total
        if self._head_outside_optimizer:
            from ...utils.popart import PopArt
            self.v_out = PopArt(args.hidden_size, 1, device=self.device_,
                                weight=self.flat_view("v_out.weight").view(self._shapes["v_out.weight"]),
                                bias=self.flat_view("v_out.bias"))

    def flat_view(self, key):
        a, b = self._slices[key]
        return self.flat[a:b]

    def unflatten(self, vec):
        """{reference key: view} of a flat vector laid out like `.flat` (gradients, Adam moments)."""
        return {k: vec[a:b].view(self._shapes[k]) for k, (a, b) in self._slices.items()}

    def load_state_dict(self, state_dict, strict=True):
        own = dict(self.named_parameters())
        missing = [k for k in own if k not in state_dict]
        unexpected = [k for k in state_dict if k not in own]
        if strict and (missing or unexpected):
            raise RuntimeError("load_state_dict: missing %s unexpected %s" % (missing, unexpected))
        with torch.no_grad():
            for k, v in state_dict.items():
                if k in own:
                    own[k].copy_(torch.as_tensor(v).to(own[k]))


class R_Actor(_FlatNet):
    def __init__(self, args, obs_space, action_space, device=torch.device("cuda")):
        super().__init__()
        self._use_policy_active_masks = args.use_policy_active_masks
        obs_shape = obs_space.shape
        if len(obs_shape) == 3:
            raise NotImplementedError("CNN base: MPE observations are vectors (r_actor_critic.py:33)")
        self.head_dims = head_sizes(action_space)
        self.discrete = action_space.__class__.__name__ == "Discrete"
        self._build(args, obs_shape[0], self.head_dims, False, device)
        self.sample_seed = int(getattr(args, "seed", 1))
        self._calls = 0
        self.step_dev = None            # int64 device scalar: the sampling counter when a CUDA graph replays the rollout

    def use_device_step(self, on=True):
        """Keep the sampling counter (Philox counter word 1) in HBM, advanced on the stream after every forward, so
        a captured rollout draws fresh numbers on every replay.  The stream of draws is the same as with the host
        counter."""
        self.step_dev = torch.tensor(self._calls, dtype=torch.int64, device=self.device_) if on else None

    def forward(self, obs, rnn_states, masks, available_actions=None, deterministic=False):
        """-> (actions int64 [B, heads], action_log_probs [B, heads], rnn_states)   r_actor_critic.py:43-70"""
        obs = check(obs, self.device_)
        B = obs.shape[0]
        nh = len(self.head_dims)
        actions = torch.empty(B, nh, dtype=torch.float32, device=self.device_)
        logp = torch.empty(B, nh, dtype=torch.float32, device=self.device_)
        rnn_states = check(rnn_states, self.device_)
        if self._recurrent:
            rnn_in = rnn_states.to(torch.float32).contiguous()
            rnn_out = torch.empty_like(rnn_in)
            self.forward_into(obs, actions, logp, available_actions=check(available_actions, self.device_),
                              deterministic=deterministic, rnn_in=rnn_in,
                              masks=check(masks, self.device_).to(torch.float32).contiguous(), rnn_out=rnn_out)
            return actions.long(), logp, rnn_out
        self.forward_into(obs, actions, logp, available_actions=check(available_actions, self.device_),
                          deterministic=deterministic)
        return actions.long(), logp, rnn_states

    def forward_into(self, obs, actions_out, logp_out, available_actions=None, deterministic=False, step=None,
                     row_offset=0, rnn_in=None, masks=None, rnn_out=None, given_actions=None):
        """Hot-path entry: writes float32 action ids and log-probs (and, recurrent, the next rnn state) straight into
        buffer slots.  given_actions (float32 ids, same shape as actions_out): evaluate those instead of sampling
        (replay of a recorded rollout); they are copied into actions_out."""
        if given_actions is not None:
            if actions_out.data_ptr() != given_actions.data_ptr():
                actions_out.copy_(given_actions.reshape(actions_out.shape))
            self._calls += 1
            ga = actions_out.reshape(obs.shape[0], -1)
            if self._recurrent:
                ops.actor_forward_rnn(self.cfg, self.flat, obs, rnn_in, masks, rnn_out, None, logp_out, given_actions=ga)
            else:
                ops.actor_forward(self.cfg, self.flat, obs, None, logp_out, given_actions=ga)
            return
        dev_step = self.step_dev if step is None else None
        if step is None:
            step = self._calls
        self._calls += 1
        if not self.discrete:
            available_actions = None
        if dev_step is not None and self._recurrent:
            ops.actor_forward_rnn(self.cfg, self.flat, obs, rnn_in, masks, rnn_out, actions_out, logp_out,
                                  available_actions=None if available_actions is None else available_actions.contiguous(),
                                  deterministic=deterministic, seed=self.sample_seed, step=0, row_offset=row_offset,
                                  step_offset=dev_step)
            dev_step.add_(1)
            return
        if dev_step is not None:
            ops.actor_forward(self.cfg, self.flat, obs, actions_out, logp_out,
                              available_actions=None if available_actions is None else available_actions.contiguous(),
                              deterministic=deterministic, seed=self.sample_seed, step=0, row_offset=row_offset,
                              step_offset=dev_step)
            dev_step.add_(1)
            return
        if self._recurrent:
            ops.actor_forward_rnn(self.cfg, self.flat, obs, rnn_in, masks, rnn_out, actions_out, logp_out,
                                  available_actions=None if available_actions is None else available_actions.contiguous(),
                                  deterministic=deterministic, seed=self.sample_seed, step=step, row_offset=row_offset)
            return
        ops.actor_forward(self.cfg, self.flat, obs, actions_out, logp_out,
                          available_actions=None if available_actions is None else available_actions.contiguous(),
                          deterministic=deterministic, seed=self.sample_seed, step=step, row_offset=row_offset)

    def evaluate_actions(self, obs, rnn_states, action, masks, available_actions=None, active_masks=None):
        """-> (action_log_probs [B, heads], dist_entropy scalar)   r_actor_critic.py:72-107, act.py:107-162"""
        obs = check(obs, self.device_)
        B, nh = obs.shape[0], len(self.head_dims)
        if self._recurrent:
            return self._evaluate_actions_rnn(obs, rnn_states, action, masks, available_actions, active_masks)
        action = check(action, self.device_).reshape(B, nh).contiguous()
        logp = torch.empty(B, nh, dtype=torch.float32, device=self.device_)
        ent = torch.empty(B, dtype=torch.float32, device=self.device_)
        avail = check(available_actions, self.device_) if self.discrete else None
        ops.actor_forward(self.cfg, self.flat, obs, None, logp, available_actions=None if avail is None else avail.contiguous(),
                          given_actions=action, entropy_out=ent)
        if active_masks is not None and self._use_policy_active_masks:
            am = check(active_masks, self.device_).reshape(B)
            dist_entropy = (ent * am).sum() / am.sum()
        else:
            dist_entropy = ent.mean()
        return logp, dist_entropy


    def _evaluate_actions_rnn(self, obs, rnn_states, action, masks, available_actions, active_masks):
        """rnn.py:30-77 (training branch): obs [L*N, D] step-major, rnn_states [N, 1, H] -- L sequential single-step
        evaluations with the hidden state carried on the device."""
        B, nh = obs.shape[0], len(self.head_dims)
        h = check(rnn_states, self.device_).to(torch.float32)
        h = h.reshape(-1, self._recurrent_N * self.hidden_size).contiguous()       # [N, recurrent_N * H]
        N = h.shape[0]
        Lc = B // N
        action = check(action, self.device_).reshape(B, nh).to(torch.float32).contiguous()
        masks = check(masks, self.device_).to(torch.float32).reshape(B).contiguous()
        logp = torch.empty(B, nh, dtype=torch.float32, device=self.device_)
        ent = torch.empty(B, dtype=torch.float32, device=self.device_)
        avail = check(available_actions, self.device_) if self.discrete else None
        for s in range(Lc):
            sl = slice(s * N, (s + 1) * N)
            h_next = torch.empty_like(h)
            ops.actor_forward_rnn(self.cfg, self.flat, obs[sl], h, masks[sl], h_next, None, logp[sl],
                                  available_actions=None if avail is None else avail[sl].contiguous(),
                                  given_actions=action[sl], entropy_out=ent[sl])
            h = h_next
        if active_masks is not None and self._use_policy_active_masks:
            am = check(active_masks, self.device_).reshape(B)
            return logp, (ent * am).sum() / am.sum()
        return logp, ent.mean()


class R_Critic(_FlatNet):
    def __init__(self, args, cent_obs_space, device=torch.device("cuda")):
        super().__init__()
        shape = cent_obs_space.shape
        if len(shape) == 3:
            raise NotImplementedError("CNN base: MPE observations are vectors")
        self._build(args, shape[0], [1], True, device)

    def forward(self, cent_obs, rnn_states, masks):
        """-> (values [B, 1], rnn_states)   r_actor_critic.py:146-165"""
        cent_obs = check(cent_obs, self.device_)
        if (not self._recurrent and torch.is_tensor(cent_obs) and cent_obs.dim() == 3 and cent_obs.stride(1) == 0
                and cent_obs.stride(2) == 1):
            # [N, M, S] with the M agent rows of an env being ONE stored row (the buffer's share_obs view,
            # utils/shared_buffer.py): evaluate each env once and write its value M times -- the same numbers as the
            # reference's M replicated rows (shared/mpe_runner.py:133-137), a third of the work
            N, M = cent_obs.shape[0], cent_obs.shape[1]
            values = torch.empty(N * M, 1, dtype=torch.float32, device=self.device_)
            self.forward_into(cent_obs[:, 0, :], values, group=M)
            return values, check(rnn_states, self.device_)
        if cent_obs.dim() == 3:
            cent_obs = cent_obs.reshape(-1, cent_obs.shape[-1])
        values = torch.empty(cent_obs.shape[0], 1, dtype=torch.float32, device=self.device_)
        if self._recurrent:
            rnn_in = check(rnn_states, self.device_).to(torch.float32).contiguous()
            rnn_out = torch.empty_like(rnn_in)
            self.forward_into(cent_obs, values, rnn_in=rnn_in,
                              masks=check(masks, self.device_).to(torch.float32).contiguous(), rnn_out=rnn_out)
            return values, rnn_out
        self.forward_into(cent_obs, values)
        return values, check(rnn_states, self.device_)

    def forward_into(self, cent_obs, values_out, group=1, rnn_in=None, masks=None, rnn_out=None, share_rows_div=1):
        """Recurrent: one value per (env, agent) sample; sample i reads share-obs row i // share_rows_div."""
        if self._recurrent:
            ops.critic_forward_rnn(self.cfg, self.flat, cent_obs, rnn_in, masks, rnn_out, values_out,
                                   share_rows_div=share_rows_div)
            return
        ops.critic_forward(self.cfg, self.flat, cent_obs, values_out, group=group)
