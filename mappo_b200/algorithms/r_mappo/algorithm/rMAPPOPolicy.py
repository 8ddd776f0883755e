"""R_MAPPOPolicy mirror (onpolicy/algorithms/r_mappo/algorithm/rMAPPOPolicy.py:7-127): owns the actor, the critic
and their two Adam optimisers; same method names, argument order and return order."""
import math

import torch

from .... import ops
from .r_actor_critic import R_Actor, R_Critic, check


class FusedAdam:
    """State of torch.optim.Adam(lr, eps=opti_eps, weight_decay) for one flat parameter buffer
    (rMAPPOPolicy.py:31-37).  The update itself runs inside mappo_adam_step together with the gradient
    clipping; `param_groups[0]['lr']` is honoured so `update_linear_schedule` keeps working."""

    def __init__(self, net, lr, eps, weight_decay, betas=(0.9, 0.999)):
        self.net = net
        self.frozen_head = bool(getattr(net, "_head_outside_optimizer", False))   # PopArt head (see utils/popart.py)
        self.param_groups = [dict(lr=lr, eps=eps, weight_decay=weight_decay, betas=betas)]
        self.exp_avg = torch.zeros_like(net.flat)
        self.exp_avg_sq = torch.zeros_like(net.flat)
        self.step_count = 0
        self.step_dev = None            # int64 device scalar mirroring step_count (CUDA-graph replays)

    def use_device_step(self, on=True):
        """Keep the step count in HBM: mappo_adam_step then derives the bias corrections from it and increments it."""
        self.step_dev = torch.tensor(self.step_count, dtype=torch.int64, device=self.exp_avg.device) if on else None

    def zero_grad(self):
        pass

    def config(self, max_grad_norm, use_max_grad_norm):
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        t = self.step_count + 1
        return ops.AdamConfig(g["lr"], b1, b2, g["eps"], g["weight_decay"], float(max_grad_norm),
                              int(bool(use_max_grad_norm)), 1.0 - b1 ** t, math.sqrt(1.0 - b2 ** t), 1.0 - b1, 1.0 - b2,
                              float(b1), float(b2), int(self.frozen_head), 0)

    def state_dict(self):
        return dict(exp_avg=self.exp_avg.clone(), exp_avg_sq=self.exp_avg_sq.clone(), step=self.step_count,
                    param_groups=[dict(g) for g in self.param_groups])

    def load_state_dict(self, sd):
        self.exp_avg.copy_(sd["exp_avg"]); self.exp_avg_sq.copy_(sd["exp_avg_sq"]); self.step_count = int(sd["step"])
        if self.step_dev is not None:
            self.step_dev.fill_(self.step_count)
        self.param_groups = [dict(g) for g in sd["param_groups"]]


class R_MAPPOPolicy:
    def __init__(self, args, obs_space, cent_obs_space, act_space, device=torch.device("cuda")):
        self.device = torch.device(device)
        self.lr = args.lr
        self.critic_lr = args.critic_lr
        self.opti_eps = args.opti_eps
        self.weight_decay = args.weight_decay
        self.obs_space = obs_space
        self.share_obs_space = cent_obs_space
        self.act_space = act_space
        self.actor = R_Actor(args, self.obs_space, self.act_space, self.device)
        self.critic = R_Critic(args, self.share_obs_space, self.device)
        self.actor_optimizer = FusedAdam(self.actor, self.lr, self.opti_eps, self.weight_decay)
        self.critic_optimizer = FusedAdam(self.critic, self.critic_lr, self.opti_eps, self.weight_decay)

    def lr_decay(self, episode, episodes):
        """utils/util.py:17-21 (update_linear_schedule) for both optimisers."""
        for opt, lr0 in ((self.actor_optimizer, self.lr), (self.critic_optimizer, self.critic_lr)):
            lr = lr0 - (lr0 * (episode / float(episodes)))
            for g in opt.param_groups:
                g["lr"] = lr

    def get_actions(self, cent_obs, obs, rnn_states_actor, rnn_states_critic, masks, available_actions=None,
                    deterministic=False):
        actions, action_log_probs, rnn_states_actor = self.actor(obs, rnn_states_actor, masks, available_actions,
                                                                 deterministic)
        values, rnn_states_critic = self.critic(cent_obs, rnn_states_critic, masks)
        return values, actions, action_log_probs, rnn_states_actor, rnn_states_critic

    def get_values(self, cent_obs, rnn_states_critic, masks):
        values, _ = self.critic(cent_obs, rnn_states_critic, masks)
        return values

    def evaluate_actions(self, cent_obs, obs, rnn_states_actor, rnn_states_critic, action, masks,
                         available_actions=None, active_masks=None):
        action_log_probs, dist_entropy = self.actor.evaluate_actions(obs, rnn_states_actor, action, masks,
                                                                     available_actions, active_masks)
        values, _ = self.critic(cent_obs, rnn_states_critic, masks)
        return values, action_log_probs, dist_entropy

    def act(self, obs, rnn_states_actor, masks, available_actions=None, deterministic=False):
        actions, _, rnn_states_actor = self.actor(obs, rnn_states_actor, masks, available_actions, deterministic)
        return actions, rnn_states_actor
