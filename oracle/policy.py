"""ORACLE (test infrastructure, not product code): CPU float32 restatement of the reference's actor/critic
forward, the PPO losses, gradient clipping and Adam, as plain functions of a `state_dict`.

  MLPBase / MLPLayer             onpolicy/algorithms/utils/mlp.py:6-59   (act BEFORE LayerNorm; fc_h unused)
  RNNLayer (GRU + LayerNorm)     onpolicy/algorithms/utils/rnn.py:7-80
  ACTLayer / FixedCategorical    onpolicy/algorithms/utils/act.py:41-162, distributions.py:14-68
  R_Actor / R_Critic             onpolicy/algorithms/r_mappo/algorithm/r_actor_critic.py:20-165
  ppo_update / cal_value_loss    onpolicy/algorithms/r_mappo/r_mappo.py:52-164
  huber_loss / mse_loss          onpolicy/utils/util.py:23-29
  clip_grad_norm_ / Adam         torch 2.11 semantics as used at r_mappo.py:143-162, rMAPPOPolicy.py:31-37

torch (CPU) supplies the dense arithmetic (F.linear, F.layer_norm, tanh) and autograd; the network
structure, loss, clipping and optimiser are restated here.  Pinned against fixtures recorded from the
reference's own R_MAPPO.ppo_update (tests/golden/ppo_*.npz) by tests/test_oracle_policy.py.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import philox

STREAM_SAMPLE = 0x53414D50  # 'SAMP'


def _t(x):
    return x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x), dtype=torch.float32)


def mlp_base(sd, x, use_relu, use_feature_norm=True, layer_N=None, prefix="base."):
    act = torch.relu if use_relu else torch.tanh
    if layer_N is None:                                     # mlp.py:20-28: as many fc2 blocks as the state dict holds
        layer_N = 0
        while prefix + "mlp.fc2.%d.0.weight" % layer_N in sd:
            layer_N += 1
    if use_feature_norm:
        x = F.layer_norm(x, x.shape[-1:], sd[prefix + "feature_norm.weight"], sd[prefix + "feature_norm.bias"])
    x = act(F.linear(x, sd[prefix + "mlp.fc1.0.weight"], sd[prefix + "mlp.fc1.0.bias"]))
    x = F.layer_norm(x, x.shape[-1:], sd[prefix + "mlp.fc1.2.weight"], sd[prefix + "mlp.fc1.2.bias"])
    for i in range(layer_N):
        p = prefix + "mlp.fc2.%d." % i
        x = act(F.linear(x, sd[p + "0.weight"], sd[p + "0.bias"]))
        x = F.layer_norm(x, x.shape[-1:], sd[p + "2.weight"], sd[p + "2.bias"])
    return x


def gru_cell(sd, x, h, prefix="rnn.rnn.", layer=0):
    """One torch.nn.GRU step of layer `layer`, gate order (r, z, n); n = tanh(W_in x + b_in + r*(W_hn h + b_hn))."""
    gi = F.linear(x, sd[prefix + "weight_ih_l%d" % layer], sd[prefix + "bias_ih_l%d" % layer])
    gh = F.linear(h, sd[prefix + "weight_hh_l%d" % layer], sd[prefix + "bias_hh_l%d" % layer])
    i_r, i_z, i_n = gi.chunk(3, -1)
    h_r, h_z, h_n = gh.chunk(3, -1)
    r = torch.sigmoid(i_r + h_r)
    z = torch.sigmoid(i_z + h_z)
    n = torch.tanh(i_n + r * h_n)
    return (1 - z) * n + z * h


def rnn_layer(sd, x, hxs, masks):
    """rnn.py:24-80: h is multiplied by mask[t] before every step; the feature passed
    on is LayerNorm(h_new), the state returned is the raw h_new.  x: [T*N, H] time-major, hxs: [N, 1, H]."""
    N = hxs.shape[0]
    T = x.shape[0] // N
    x = x.view(T, N, -1)
    m = masks.view(T, N, 1)
    RN = hxs.shape[1]                                   # recurrent_N stacked layers (nn.GRU(num_layers=recurrent_N), rnn.py:13)
    h = [hxs[:, l] for l in range(RN)]
    outs = []
    for t in range(T):
        inp = x[t]
        for l in range(RN):                             # every layer's state is masked before the step (rnn.py:27,62-66)
            h[l] = gru_cell(sd, inp, h[l] * m[t], layer=l)
            inp = h[l]
        outs.append(inp)
    y = torch.stack(outs, 0).reshape(T * N, -1)
    y = F.layer_norm(y, y.shape[-1:], sd["rnn.norm.weight"], sd["rnn.norm.bias"])
    return y, torch.stack(h, 1)


def head_keys(sd):
    if "act.action_out.linear.weight" in sd:
        return ["act.action_out.linear."]
    ks, i = [], 0
    while "act.action_outs.%d.linear.weight" % i in sd:
        ks.append("act.action_outs.%d.linear." % i)
        i += 1
    return ks


def actor_features(sd, obs, rnn_states, masks, use_relu, use_feature_norm, recurrent):
    x = mlp_base(sd, _t(obs), use_relu, use_feature_norm)
    h = _t(rnn_states)
    if recurrent:
        x, h = rnn_layer(sd, x, h, _t(masks))
    return x, h


def actor_logits(sd, feats, available_actions=None):
    outs = []
    for i, k in enumerate(head_keys(sd)):
        l = F.linear(feats, sd[k + "weight"], sd[k + "bias"])
        if available_actions is not None and i == 0 and len(head_keys(sd)) == 1:
            l = torch.where(_t(available_actions) == 0, torch.full_like(l, -1e10), l)   # distributions.py:66-67
        outs.append(l)
    return outs


def critic_values(sd, share_obs, rnn_states, masks, use_relu, use_feature_norm, recurrent):
    x = mlp_base(sd, _t(share_obs), use_relu, use_feature_norm)
    h = _t(rnn_states)
    if recurrent:
        x, h = rnn_layer(sd, x, h, _t(masks))
    return F.linear(x, sd["v_out.weight"], sd["v_out.bias"]), h


def sample_uniforms(rows, step, head, seed, row_offset=0):
    """u in (0,1) float32 for (row, step, head): Philox(counter=(row_offset+row, step, head, 0), key=(seed,'SAMP')).x,
    24-bit: ((x >> 8) + 0.5) * 2^-24."""
    r = philox.philox4x32(np.arange(rows, dtype=np.uint64) + np.uint64(row_offset), np.uint64(step), np.uint64(head),
                          np.uint64(0), seed, STREAM_SAMPLE)[0]
    return ((r >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)


def act_forward(logits_list, deterministic, seed=0, step=0, row_offset=0):
    """ACTLayer.forward (act.py:41-81): per-head action and log-prob, concatenated.  Sampling is inverse-CDF
    on exp(log_softmax) accumulated in float32 in index order (first a with u < cumsum)."""
    acts, lps, margins = [], [], []
    for h, l in enumerate(logits_list):
        logp = l - torch.logsumexp(l, -1, keepdim=True)
        if deterministic:
            a = logp.argmax(-1)
            margin = torch.full((l.shape[0],), 1.0)
        else:
            u = torch.as_tensor(sample_uniforms(l.shape[0], step, h, seed, row_offset))
            p = torch.exp(logp)
            cum = torch.zeros(l.shape[0])
            a = torch.full((l.shape[0],), l.shape[1] - 1, dtype=torch.long)
            done = torch.zeros(l.shape[0], dtype=torch.bool)
            margin = torch.full((l.shape[0],), 1.0)
            for q in range(l.shape[1]):
                cum = cum + p[:, q]
                margin = torch.minimum(margin, (u - cum).abs())
                hit = (u < cum) & ~done
                a[hit] = q
                done |= hit
        acts.append(a)
        lps.append(logp.gather(-1, a.unsqueeze(-1)).squeeze(-1))
        margins.append(margin)
    return torch.stack(acts, -1), torch.stack(lps, -1), torch.stack(margins, -1)


def evaluate_actions(logits_list, actions, active_masks=None):
    """ACTLayer.evaluate_actions (act.py:107-162): per-head log-probs [B, heads]; entropy = mean over heads of
    the (active-masked) mean entropy."""
    actions = _t(actions).long()
    lps, ents = [], []
    for h, l in enumerate(logits_list):
        logp = l - torch.logsumexp(l, -1, keepdim=True)
        lps.append(logp.gather(-1, actions[:, h:h + 1]).squeeze(-1))
        ent = -(torch.exp(logp) * logp).sum(-1)
        if active_masks is not None:
            am = _t(active_masks).squeeze(-1)
            ents.append((ent * am).sum() / am.sum())
        else:
            ents.append(ent.mean())
    return torch.stack(lps, -1), sum(ents) / len(ents)


def huber_loss(e, d):
    a = (e.abs() <= d).float()
    b = (e.abs() > d).float()
    return a * e ** 2 / 2 + b * d * (e.abs() - d / 2)


def policy_loss(action_log_probs, old_log_probs, adv, active_masks, clip_param, use_policy_active_masks=True):
    """r_mappo.py:124-136."""
    ratio = torch.exp(action_log_probs - _t(old_log_probs))
    adv = _t(adv)
    surr1 = ratio * adv
    surr2 = torch.clamp(ratio, 1.0 - clip_param, 1.0 + clip_param) * adv
    term = -torch.sum(torch.min(surr1, surr2), dim=-1, keepdim=True)
    if use_policy_active_masks:
        am = _t(active_masks)
        return (term * am).sum() / am.sum(), ratio
    return term.mean(), ratio


def value_loss(values, value_preds, returns, active_masks, clip_param, normalizer=None, use_huber=True,
               huber_delta=10.0, use_clipped=True, use_value_active_masks=True):
    """r_mappo.py:52-89.  `normalizer`: oracle.buffer.OracleValueNorm, updated here BEFORE normalising (:65-67)."""
    value_preds, returns = _t(value_preds), _t(returns)
    clipped = value_preds + (values - value_preds).clamp(-clip_param, clip_param)
    if normalizer is not None:
        normalizer.update(returns.numpy())
        target = torch.as_tensor(normalizer.normalize(returns.numpy()))
    else:
        target = returns
    e_c, e_o = target - clipped, target - values
    if use_huber:
        l_c, l_o = huber_loss(e_c, huber_delta), huber_loss(e_o, huber_delta)
    else:
        l_c, l_o = e_c ** 2 / 2, e_o ** 2 / 2
    l = torch.max(l_o, l_c) if use_clipped else l_o
    if use_value_active_masks:
        am = _t(active_masks)
        return (l * am).sum() / am.sum()
    return l.mean()


def clip_grad_norm(grads, max_norm):
    """nn.utils.clip_grad_norm_: total L2 norm over all grads; scale by clamp(max_norm/(norm+1e-6), max=1)."""
    total = torch.sqrt(sum((g.double() ** 2).sum() for g in grads)).float()
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    return total, [g * coef for g in grads]


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, eps=1e-5, beta1=0.9, beta2=0.999, weight_decay=0.0):
    """torch.optim.Adam single-tensor update (no amsgrad), `step` counted from 1."""
    if weight_decay != 0:
        grad = grad + weight_decay * param
    exp_avg = exp_avg + (grad - exp_avg) * (1 - beta1)
    exp_avg_sq = exp_avg_sq * beta2 + (1 - beta2) * grad * grad
    bc1 = 1 - beta1 ** step
    bc2_sqrt = math.sqrt(1 - beta2 ** step)
    denom = exp_avg_sq.sqrt() / bc2_sqrt + eps
    return param - (lr / bc1) * (exp_avg / denom), exp_avg, exp_avg_sq
