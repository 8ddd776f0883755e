"""GPU: gradient parity at batch sizes of the order the benchmark runs (north_star: losses and gradients within 1e-4).

The reference-recorded fixtures (tests/golden/ppo_*.npz) have 257-700 rows: 3-6 tiles on a handful of CTAs.  The
benchmark feeds 4.9 M rows: ~260 tiles per CTA accumulate into the same tensor-memory weight-gradient blocks
(csrc/mlp_tc.cu), the fp16-pair split meets the whole range of activations, and 296 partial rows are summed
(csrc/optim.cu).  Here every update kernel runs >= 200 k rows (>= 10 tiles on each of the 148 CTAs) and its loss and
PRE-clip gradients are compared with the oracle (oracle/policy.py: the reference's network / loss restated; pinned by
tests/test_oracle_policy.py) evaluated in FLOAT64 with torch autograd on the same inputs -- so the comparison measures
the kernels' error, not the fp32 oracle's own summation error at this size:
  * MODE 2  mappo_actor_update (D = 18, 5 logits, Tanh: the C2 actor), 300,087 rows (ragged last tile)
  * MODE 3  mappo_critic_update with group = 3 (C2's centralised critic: one share_obs row serves the 3 agents)
  * the packed wide critic at S = 460, group = 10 (C5), packed and unpacked first layer
  * the recurrent chain (simple_reference GRU actor + critic, MultiDiscrete heads), 20,480 chunks of 10 steps
Tolerance per parameter tensor: rtol 1e-4, atol 1e-4 * max|g| (the form tests/test_gpu_policy.py uses on the fixtures).

ReLU networks (the reference's C3 / C4 scripts; C1 / C2 / C5 pass --use_ReLU, a store_false flag = Tanh): a sample whose
pre-activation lies within rounding of 0 has derivative 0 or 1 depending on the last bit, so ANY float32 implementation
differs from the float64 gradient by whole per-sample contributions there (torch's own fp32 autograd is 40x farther
from fp64 with ReLU than with Tanh at 300 k rows: 3.8e-5 vs 1e-7 of max|g|, probed).  The comparison is only
well-posed away from the kink, so the ReLU cases switch off (active_masks = 0) the ~0.1 % of samples that have a
float64 pre-activation within 1e-5 of zero in either base layer; every other sample is compared as strictly as Tanh."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _policy(flags, D, S, act, seed=0, perturb=0.05):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.spaces import Box
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    from mappo_b200.algorithms.r_mappo.r_mappo import R_MAPPO
    a = parse_args(("--scenario_name simple_spread --ppo_epoch 1 --num_mini_batch 1 " + flags).split())
    torch.manual_seed(seed)
    box = lambda n: Box(-np.inf, np.inf, (n,), np.float32)
    pol = R_MAPPOPolicy(a, box(D), box(S), act, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(seed + 1)
    for net in (pol.actor, pol.critic):                     # LayerNorm affines and biases away from (1, 0)
        net.flat.add_(perturb * torch.randn(net.flat.shape, device="cuda", generator=g))
    return a, pol, R_MAPPO(a, pol, device="cuda"), g


def _sd64(net):
    return {k: v.detach().cpu().double().clone().requires_grad_("fc_h" not in k) for k, v in net.state_dict().items()}


def _d(x):
    return None if x is None else x.detach().cpu().double()


def _threads():
    from conftest import DEFAULT_TORCH_THREADS
    return max(1, DEFAULT_TORCH_THREADS)


def _check(net, got_flat, want, what):
    worst = 0.0
    for k, gv in net.unflatten(got_flat).items():
        w = want[k].numpy()
        g = gv.cpu().numpy().astype(np.float64)
        scale = np.abs(w).max()
        np.testing.assert_allclose(g, w, rtol=1e-4, atol=1e-4 * scale + 1e-12, err_msg="%s: %s" % (what, k))
        worst = max(worst, float(np.abs(g - w).max() / (scale + 1e-30)))
    print("%s: worst |dg| / max|g| over the parameter tensors = %.2e" % (what, worst))


def _actor_oracle64(a, pol, obs, actions, old_logp, adv, active, rnn=None, masks=None):
    from oracle import policy as op
    sd = _sd64(pol.actor)
    rec = rnn is not None
    B = obs.shape[0]
    z = torch.zeros(B, 1, 64, dtype=torch.float64) if not rec else _d(rnn)
    m = torch.ones(B, 1, dtype=torch.float64) if not rec else _d(masks)
    feats, _ = op.actor_features(sd, _d(obs), z, m, a.use_ReLU, a.use_feature_normalization, rec)
    logits = op.actor_logits(sd, feats, None)
    logp, ent = op.evaluate_actions(logits, _d(actions), _d(active) if a.use_policy_active_masks else None)
    pl, ratio = op.policy_loss(logp, _d(old_logp), _d(adv), _d(active), a.clip_param, a.use_policy_active_masks)
    keys = [k for k in sd if "fc_h" not in k]
    grads = torch.autograd.grad(pl - ent * a.entropy_coef, [sd[k] for k in keys])
    return pl.item(), ent.item(), ratio.mean().item(), dict(zip(keys, grads))


def _critic_oracle64(a, pol, tr, share, group, value_preds, returns, active, vn_before, rnn=None, masks=None):
    """share [rows_c, S]: row i serves samples i*group .. i*group + group - 1 (the buffer's stride-0 share_obs view)."""
    from oracle import policy as op
    from oracle.buffer import OracleValueNorm
    sd = _sd64(pol.critic)
    rec = rnn is not None
    rows = share.shape[0]
    z = torch.zeros(rows, 1, 64, dtype=torch.float64) if not rec else _d(rnn)
    m = torch.ones(rows, 1, dtype=torch.float64) if not rec else _d(masks)
    values, _ = op.critic_values(sd, _d(share), z, m, a.use_ReLU, a.use_feature_normalization, rec)
    if group > 1:
        values = values.repeat_interleave(group, dim=0)
    vn = None
    if vn_before is not None:
        vn = OracleValueNorm(); vn.load(vn_before)
    vl = op.value_loss(values, _d(value_preds), returns.detach().cpu().float(), _d(active), a.clip_param, vn,
                       a.use_huber_loss, a.huber_delta, a.use_clipped_value_loss, a.use_value_active_masks)
    keys = [k for k in sd if "fc_h" not in k]
    grads = torch.autograd.grad(vl * a.value_loss_coef, [sd[k] for k in keys])
    return vl.item(), dict(zip(keys, grads))


def _near_kink(net, x, delta=1e-5):
    """bool [rows]: some pre-activation of the MLP base (mlp.py:44-59) is within `delta` of 0 in float64"""
    import torch.nn.functional as F
    sd = {k: v.detach().cpu().double() for k, v in net.state_dict().items()}
    x = _d(x)
    if "base.feature_norm.weight" in sd:
        x = F.layer_norm(x, x.shape[-1:], sd["base.feature_norm.weight"], sd["base.feature_norm.bias"])
    z1 = F.linear(x, sd["base.mlp.fc1.0.weight"], sd["base.mlp.fc1.0.bias"])
    n1 = F.layer_norm(torch.relu(z1), (64,), sd["base.mlp.fc1.2.weight"], sd["base.mlp.fc1.2.bias"])
    z2 = F.linear(n1, sd["base.mlp.fc2.0.0.weight"], sd["base.mlp.fc2.0.0.bias"])
    return ((z1.abs() < delta).any(1) | (z2.abs() < delta).any(1)).cuda()


def _mask_kinks(a, pol, obs, share, group, active):
    """ReLU only: active = 0 for samples whose actor row or critic row sits on a kink (see the module docstring)."""
    if not a.use_ReLU:
        return active
    bad = _near_kink(pol.actor, obs) | _near_kink(pol.critic, share).repeat_interleave(group)
    frac = bad.float().mean().item()
    assert frac < 0.02, frac
    print("ReLU: %.3f %% of the samples masked as near-kink" % (100 * frac))
    return active * (~bad).float().view(active.shape)


def _mlp_batch(pol, g, B, D, A, rows_c, S):
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    obs = r(B, D) * (0.4 + r(1, D).abs()) + 0.3 * r(1, D)            # MPE-like: O(1) columns of different scale
    share = r(rows_c, S) * (0.4 + r(1, S).abs()) + 0.3 * r(1, S)
    acts = torch.empty(B, 1, device="cuda"); logp = torch.empty(B, 1, device="cuda")
    pol.actor.forward_into(obs, acts, logp, step=3)                  # actions the policy itself samples
    old_logp = logp + 0.15 * r(B, 1)                                 # ratios on both sides of the clip range
    adv = r(B)
    active = (torch.rand(B, device="cuda", generator=g) > 0.1).float()
    vp = 0.5 * r(B)
    ret = -5.0 + 3.0 * r(B)
    return obs, share, acts, old_logp, adv, active, vp, ret


@pytest.mark.parametrize("flags,group", [("--use_ReLU", 3), ("", 1)])
def test_actor_and_critic_update_gradients_at_300k_rows(flags, group):
    """C2's networks: actor D = 18 -> 5 logits, centralised critic S = 54 with `group` agents per share_obs row.
    `--use_ReLU` is store_false in the reference's parser: that case is Tanh (the benchmark's setting)."""
    from mappo_b200.envs.spaces import Discrete
    D, S, A = 18, 54, 5
    rows_c = 100_029 if group == 3 else 300_077
    B = rows_c * group
    a, pol, tr, g = _policy(flags, D, S, Discrete(A))
    obs, share, acts, old_logp, adv, active, vp, ret = _mlp_batch(pol, g, B, D, A, rows_c, S)
    active = _mask_kinks(a, pol, obs, share, group, active)
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    vn_before = tr.value_normalizer.stats.cpu().numpy().copy()
    torch.set_num_threads(_threads())
    try:
        pl, ent, ratio, ga = _actor_oracle64(a, pol, obs, acts, old_logp.view(B, 1), adv.view(B, 1), active.view(B, 1))
        vl, gc = _critic_oracle64(a, pol, tr, share, group, vp.view(B, 1), ret.view(B, 1), active.view(B, 1), vn_before)
    finally:
        torch.set_num_threads(1)
    ws = tr._update(obs, share, group, acts, old_logp, adv, vp, ret, active, None, B, update_actor=True)
    info = tr._info.cpu().numpy()
    np.testing.assert_allclose(info[0], pl, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(info[2], ent, rtol=1e-5)
    np.testing.assert_allclose(info[3], ratio, rtol=1e-5)
    np.testing.assert_allclose(info[4], vl, rtol=1e-4, atol=1e-6)
    Pa = pol.actor.flat.numel()
    _check(pol.actor, ws["grad"][:Pa], ga, "actor update, %d rows" % B)
    _check(pol.critic, ws["grad"][Pa:], gc, "critic update, %d share rows x group %d" % (rows_c, group))


@pytest.mark.parametrize("packed,relu", [(True, False), (False, False), (True, True)])
def test_wide_critic_gradients_at_200k_rows(packed, relu):
    """C5's centralised critic: S = 10 x 46 = 460 columns, one row per env serving 10 agents; the first layer as the
    packed (fp16-pair tiles written once, bulk-copy-fed streaming GEMMs) and the unpacked streaming kernels."""
    from mappo_b200 import ops
    from mappo_b200.envs.spaces import Discrete
    D, S, A, M = 46, 460, 5, 10
    rows_c = 200_064 + 5
    B = rows_c * M
    a, pol, tr, g = _policy("--num_agents 10" + ("" if relu else " --use_ReLU"), D, S, Discrete(A))
    assert bool(a.use_ReLU) == relu
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    share = r(rows_c, S) * (0.4 + r(1, S).abs()) + 0.3 * r(1, S)
    obs = r(B, D)
    acts = torch.randint(0, A, (B, 1), device="cuda", generator=g).float()
    old_logp = -1.6 + 0.1 * r(B, 1)
    adv, vp, ret = r(B), 0.5 * r(B), -5.0 + 3.0 * r(B)
    active = (torch.rand(B, device="cuda", generator=g) > 0.1).float()
    if relu:
        active = active * (~_near_kink(pol.critic, share).repeat_interleave(M)).float()
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    vn_before = tr.value_normalizer.stats.cpu().numpy().copy()
    torch.set_num_threads(_threads())
    try:
        vl, gc = _critic_oracle64(a, pol, tr, share, M, vp.view(B, 1), ret.view(B, 1), active.view(B, 1), vn_before)
    finally:
        torch.set_num_threads(1)
    pk = ops.critic_pack_inputs(pol.critic.cfg, share) if packed else None
    assert (pk is not None) == packed
    ws = tr._update(obs, share, M, acts, old_logp, adv, vp, ret, active, None, B, packed=pk)
    np.testing.assert_allclose(tr._info[4].item(), vl, rtol=1e-4, atol=1e-6)
    _check(pol.critic, ws["grad"][pol.actor.flat.numel():], gc,
           "wide critic (%s), %d share rows x group %d" % ("packed" if packed else "unpacked", rows_c, M))


@pytest.mark.parametrize("relu", [False, True])
def test_recurrent_update_gradients_at_20k_chunks(relu):
    """C4's networks (simple_reference: D = 21, S = 42, MultiDiscrete heads 5 + 10, GRU + LayerNorm) on 20,480 chunks of
    10 steps with random masks inside the chunks and random initial states: back-propagation through time on the
    tensor cores against float64 autograd through the oracle's GRU."""
    from mappo_b200.envs.spaces import MultiDiscrete
    D, S, L, Nc = 21, 42, 10, 20_480
    B = L * Nc
    # train_mpe_reference.sh does not pass --use_ReLU: C4 itself is the ReLU case
    a, pol, tr, g = _policy("--algorithm_name rmappo --data_chunk_length 10" + ("" if relu else " --use_ReLU"), D, S,
                            MultiDiscrete([[0, 4], [0, 9]]))
    assert bool(a.use_ReLU) == relu
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    obs = r(B, D) * (0.4 + r(1, D).abs())
    share = r(B, S) * (0.4 + r(1, S).abs())
    rnn_a, rnn_c = 0.5 * r(Nc, 1, 64), 0.5 * r(Nc, 1, 64)
    masks = (torch.rand(B, 1, device="cuda", generator=g) > 0.08).float()
    acts = torch.stack([torch.randint(0, 5, (B,), device="cuda", generator=g),
                        torch.randint(0, 10, (B,), device="cuda", generator=g)], -1).float()
    old_logp = torch.stack([-1.6 + 0.1 * r(B), -2.3 + 0.1 * r(B)], -1)
    adv, vp, ret = r(B, 1), 0.5 * r(B, 1), -5.0 + 3.0 * r(B, 1)
    active = (torch.rand(B, 1, device="cuda", generator=g) > 0.1).float()
    active = _mask_kinks(a, pol, obs, share, 1, active)
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    vn_before = tr.value_normalizer.stats.cpu().numpy().copy()
    torch.set_num_threads(_threads())
    try:
        pl, ent, ratio, ga = _actor_oracle64(a, pol, obs, acts, old_logp, adv, active, rnn=rnn_a, masks=masks)
        vl, gc = _critic_oracle64(a, pol, tr, share, 1, vp, ret, active, vn_before, rnn=rnn_c, masks=masks)
    finally:
        torch.set_num_threads(1)
    sample = (share, obs, rnn_a, rnn_c, acts, vp, ret, masks, active, old_logp, adv, None)
    got = tr.ppo_update(sample)
    np.testing.assert_allclose(got[2].item(), pl, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(got[3].item(), ent, rtol=1e-5)
    np.testing.assert_allclose(got[0].item(), vl, rtol=1e-4, atol=1e-6)
    ws = next(iter(tr._ws.values()))
    Pa = pol.actor.flat.numel()
    _check(pol.actor, ws["grad"][:Pa], ga, "recurrent actor update, %d chunks x %d steps" % (Nc, L))
    _check(pol.critic, ws["grad"][Pa:], gc, "recurrent critic update, %d chunks x %d steps" % (Nc, L))
