"""Shared by the CPU oracle test and the GPU parity tests: run one PPO update on a fixture's sample with the
oracle (oracle/policy.py + oracle/buffer.py) and return everything the reference returned."""
import numpy as np
import torch

from oracle import policy as op
from oracle.buffer import OracleValueNorm


def fx_args(fx):
    return {k[5:]: fx[k].item() for k in fx.files if k.startswith("args_")}


def sd_from(fx, prefix):
    return {k[len(prefix):]: torch.tensor(fx[k]) for k in fx.files if k.startswith(prefix)}


def trainable(sd):
    """Parameters the reference's optimiser actually moves: everything but the dead fc_h template."""
    return [k for k in sd if "fc_h" not in k]


def oracle_update(fx, actor_sd, critic_sd, vn_stats, adam_state, step):
    """One ppo_update (r_mappo.py:91-164) restated with the oracle.  Mutates nothing; returns a dict."""
    a = fx_args(fx)
    rec = bool(fx["recurrent"])
    s = {k[7:]: fx[k] for k in fx.files if k.startswith("sample_")}
    B = int(fx["B"])
    zeros = np.zeros((B, 1, int(a["hidden_size"])), np.float32)
    actor_sd = {k: v.clone().requires_grad_("fc_h" not in k) for k, v in actor_sd.items()}
    critic_sd = {k: v.clone().requires_grad_("fc_h" not in k) for k, v in critic_sd.items()}
    feats, _ = op.actor_features(actor_sd, s["obs"], s.get("rnn_states", zeros), s["masks"], a["use_ReLU"],
                                 a["use_feature_normalization"], rec)
    logits = op.actor_logits(actor_sd, feats, s.get("available_actions"))
    logp, ent = op.evaluate_actions(logits, s["actions"], s["active_masks"] if a["use_policy_active_masks"] else None)
    values, _ = op.critic_values(critic_sd, s["share_obs"], s.get("rnn_states_critic", zeros), s["masks"],
                                 a["use_ReLU"], a["use_feature_normalization"], rec)
    pl, ratio = op.policy_loss(logp, s["old_action_log_probs"], s["adv_targ"], s["active_masks"], a["clip_param"],
                               a["use_policy_active_masks"])
    akeys, ckeys = trainable(actor_sd), trainable(critic_sd)
    agrads = torch.autograd.grad(pl - ent * a["entropy_coef"], [actor_sd[k] for k in akeys])
    vn = None
    if a["use_valuenorm"]:
        vn = OracleValueNorm(); vn.load(vn_stats)
    vl = op.value_loss(values, s["value_preds"], s["returns"], s["active_masks"], a["clip_param"], vn,
                       a["use_huber_loss"], a["huber_delta"], a["use_clipped_value_loss"], a["use_value_active_masks"])
    cgrads = torch.autograd.grad(vl * a["value_loss_coef"], [critic_sd[k] for k in ckeys])
    out = dict(policy_loss=pl.item(), dist_entropy=ent.item(), value_loss=vl.item(), ratio_mean=ratio.mean().item(),
               actor_grads=dict(zip(akeys, agrads)), critic_grads=dict(zip(ckeys, cgrads)),
               vn=None if vn is None else vn.stats())
    for name, keys, grads, sd, lr in (("actor", akeys, agrads, actor_sd, a["lr"]), ("critic", ckeys, cgrads, critic_sd, a["critic_lr"])):
        if a["use_max_grad_norm"]:
            norm, clipped = op.clip_grad_norm(grads, a["max_grad_norm"])
        else:
            norm, clipped = torch.sqrt(sum((g ** 2).sum() for g in grads)), list(grads)
        out[name + "_grad_norm"] = float(norm)
        new_sd = {k: v.detach().clone() for k, v in sd.items()}
        new_state = {}
        for k, g in zip(keys, clipped):
            m, v = adam_state[name].get(k, (torch.zeros_like(g), torch.zeros_like(g)))
            p, m, v = op.adam_step(sd[k].detach(), g, m, v, step, lr, a["opti_eps"], weight_decay=a["weight_decay"])
            new_sd[k] = p
            new_state[k] = (m, v)
        out[name + "_sd"] = new_sd
        out[name + "_state"] = new_state
    return out
