"""GPU: fused actor/critic kernels, PPO update, clipping and Adam against (a) fixtures recorded from the
reference's own R_MAPPO.ppo_update and (b) the oracle.  Tolerances are north_star's: losses and gradients 1e-4."""
import os
import types

import numpy as np
import pytest
import torch

from ppo_common import fx_args, sd_from

pytestmark = pytest.mark.gpu
CASES = ["mlp_tanh", "mlp_relu_opts", "mlp_listener", "mlp_layer2", "mlp_layer3_relu"]   # the last two: layer_N = 2, 3


def make_space(n):
    from mappo_b200.envs.spaces import Box
    return Box(-np.inf, np.inf, (n,), np.float32)


def build(fx, seed=1):
    from mappo_b200.envs.spaces import Discrete, MultiDiscrete
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    from mappo_b200.algorithms.r_mappo.r_mappo import R_MAPPO
    a = fx_args(fx)
    args = types.SimpleNamespace(**a)
    args.use_recurrent_policy = bool(fx["recurrent"]); args.use_naive_recurrent_policy = False
    args.recurrent_N = int(fx["sample_rnn_states"].shape[1]) if (bool(fx["recurrent"]) and "sample_rnn_states" in fx.files) else 1
    args.use_popart = bool(a.get("use_popart", False)); args.use_orthogonal = True; args.seed = seed
    args.ppo_epoch = 1; args.num_mini_batch = 1; args.data_chunk_length = int(fx["chunk"])
    heads = [int(h) for h in fx["act_heads"]]
    act = Discrete(heads[0]) if len(heads) == 1 else MultiDiscrete([[0, h - 1] for h in heads])
    pol = R_MAPPOPolicy(args, make_space(int(fx["D"])), make_space(int(fx["S"])), act, device="cuda")
    pol.actor.load_state_dict(sd_from(fx, "actor0."))
    pol.critic.load_state_dict(sd_from(fx, "critic0."))
    tr = R_MAPPO(args, pol, device="cuda")
    if "vn0" in fx.files:
        tr.value_normalizer.stats.copy_(torch.as_tensor(fx["vn0"]))
    return args, pol, tr


def sample_of(fx):
    s = {k[7:]: fx[k] for k in fx.files if k.startswith("sample_")}
    B = int(fx["B"])
    z = np.zeros((B, 1, 64), np.float32)
    return (s["share_obs"], s["obs"], s.get("rnn_states", z), s.get("rnn_states_critic", z), s["actions"],
            s["value_preds"], s["returns"], s["masks"], s["active_masks"], s["old_action_log_probs"], s["adv_targ"],
            s.get("available_actions"))


@pytest.mark.parametrize("tag", CASES)
def test_state_dict_keys_and_forward_vs_oracle(tag, golden_dir):
    from oracle import policy as op
    fx = np.load(os.path.join(golden_dir, "ppo_%s.npz" % tag))
    args, pol, tr = build(fx)
    want_keys = [k[7:] for k in fx.files if k.startswith("actor0.")]
    assert list(pol.actor.state_dict().keys()) == want_keys
    assert list(pol.critic.state_dict().keys()) == [k[8:] for k in fx.files if k.startswith("critic0.")]
    s = sample_of(fx)
    share_obs, obs, avail = s[0], s[1], s[11]
    asd, csd = sd_from(fx, "actor0."), sd_from(fx, "critic0.")
    feats, _ = op.actor_features(asd, obs, s[2], s[7], args.use_ReLU, args.use_feature_normalization, False)
    logits = op.actor_logits(asd, feats, avail)
    v_ref, _ = op.critic_values(csd, share_obs, s[3], s[7], args.use_ReLU, args.use_feature_normalization, False)
    # critic values
    values = pol.get_values(share_obs, s[3], s[7])
    np.testing.assert_allclose(values.cpu().numpy(), v_ref.numpy(), rtol=1e-4, atol=2e-5)
    # deterministic actions = argmax, log-probs
    a_ref, lp_ref, _ = op.act_forward(logits, True)
    act, lp, _ = pol.actor(obs, s[2], s[7], avail, deterministic=True)
    assert (act.cpu().numpy() == a_ref.numpy()).mean() > 0.995
    same = act.cpu().numpy() == a_ref.numpy()
    np.testing.assert_allclose(lp.cpu().numpy()[same], lp_ref.numpy()[same], rtol=1e-4, atol=2e-5)
    # sampling: same Philox stream and inverse CDF as the oracle; ties at a CDF boundary may differ
    pol.actor.sample_seed = 99
    acts = torch.empty(obs.shape[0], len(logits), device="cuda"); lps = torch.empty_like(acts)
    pol.actor.forward_into(torch.as_tensor(obs).cuda(), acts, lps,
                           available_actions=None if avail is None else torch.as_tensor(avail).cuda(), step=5, row_offset=1000)
    a_s, lp_s, margin = op.act_forward(logits, False, seed=99, step=5, row_offset=1000)
    diff = acts.cpu().numpy() != a_s.numpy()
    assert not (diff & (margin.numpy() > 1e-5)).any(), "sampled action differs away from a CDF boundary"
    np.testing.assert_allclose(lps.cpu().numpy()[~diff], lp_s.numpy()[~diff], rtol=1e-4, atol=2e-5)
    counts = np.bincount(acts.cpu().numpy()[:, 0].astype(int), minlength=logits[0].shape[1])
    assert (counts > 0).sum() >= 2
    # evaluate_actions
    vals, elp, ent = pol.evaluate_actions(share_obs, obs, s[2], s[3], s[4], s[7], avail, s[8])
    lp_e, ent_e = op.evaluate_actions(logits, s[4], s[8] if args.use_policy_active_masks else None)
    np.testing.assert_allclose(elp.cpu().numpy(), lp_e.numpy(), rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(ent.item(), ent_e.item(), rtol=1e-5)


@pytest.mark.parametrize("tag", CASES + ["gru_reference", "gru2_reference"])   # the last: recurrent_N = 2
def test_ppo_update_matches_reference(tag, golden_dir):
    """Two consecutive ppo_updates on the reference's own minibatch: losses, PRE-clip gradients, post-step
    parameters, Adam moments, ValueNorm statistics."""
    fx = np.load(os.path.join(golden_dir, "ppo_%s.npz" % tag))
    args, pol, tr = build(fx)
    s = sample_of(fx)
    Pa = pol.actor.flat.numel()
    for u in range(2):
        vl, cgn, pl, ent, agn, ratio = tr.ppo_update(s)
        t = "u%d_" % u
        got = dict(value_loss=vl, critic_grad_norm=cgn, policy_loss=pl, dist_entropy=ent, actor_grad_norm=agn,
                   ratio_mean=ratio)
        for k, v in got.items():
            np.testing.assert_allclose(v.item(), fx[t + k].item(), rtol=1e-4, atol=1e-6, err_msg=k)
        ws = next(iter(tr._ws.values()))
        for net, name, g in ((pol.actor, "actor", ws["grad"][:Pa]), (pol.critic, "critic", ws["grad"][Pa:])):
            for k, gv in net.unflatten(g).items():
                want = fx[t + name + "_grad." + k]
                np.testing.assert_allclose(gv.cpu().numpy(), want, rtol=1e-4, atol=1e-4 * np.abs(want).max() + 1e-9,
                                           err_msg="%s grad %s (update %d)" % (name, k, u))
            for k, p in net.state_dict().items():
                if "fc_h" in k:
                    continue
                np.testing.assert_allclose(p.cpu().numpy(), fx[t + name + "." + k], rtol=1e-5, atol=3e-6, err_msg=k)
            opt = pol.actor_optimizer if name == "actor" else pol.critic_optimizer
            for k, m in net.unflatten(opt.exp_avg).items():
                want = fx[t + name + "_exp_avg." + k]
                np.testing.assert_allclose(m.cpu().numpy(), want, rtol=1e-4, atol=1e-4 * np.abs(want).max() + 1e-10)
            for k, m in net.unflatten(opt.exp_avg_sq).items():
                want = fx[t + name + "_exp_avg_sq." + k]
                np.testing.assert_allclose(m.cpu().numpy(), want, rtol=2e-4, atol=2e-4 * np.abs(want).max() + 1e-12)
        if "vn0" in fx.files:
            np.testing.assert_allclose(tr.value_normalizer.stats.cpu().numpy(), fx[t + "vn"], rtol=1e-6)


def test_reference_init_is_bit_identical(golden_dir):
    """Same torch seed -> same initial weights as the reference constructors (the fixture was built after
    torch.manual_seed(3); its LayerNorm/bias perturbation does not touch the matrices of fc1/fc2)."""
    from mappo_b200.envs.spaces import Discrete
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    fx = np.load(os.path.join(golden_dir, "ppo_mlp_tanh.npz"))
    a = fx_args(fx)
    args = types.SimpleNamespace(**a)
    args.use_recurrent_policy = args.use_naive_recurrent_policy = False
    args.recurrent_N = 1; args.use_popart = False; args.use_orthogonal = True; args.seed = 1
    # orthogonal_ runs LAPACK's QR, whose blocking (and so the last bits) follows the intra-op thread count: use the
    # multi-threaded default the fixture was recorded with, not the single thread the oracle checks run on
    from conftest import DEFAULT_TORCH_THREADS
    torch.set_num_threads(max(2, DEFAULT_TORCH_THREADS))
    try:
        torch.manual_seed(3)
        pol = R_MAPPOPolicy(args, make_space(18), make_space(54), Discrete(5), device="cuda")
    finally:
        torch.set_num_threads(1)
    for net, pre in ((pol.actor, "actor0."), (pol.critic, "critic0.")):
        sd = net.state_dict()
        for k in ("base.mlp.fc1.0.weight", "base.mlp.fc2.0.0.weight", "base.mlp.fc_h.0.weight"):
            np.testing.assert_array_equal(sd[k].cpu().numpy(), fx[pre + k])


def test_gradient_is_sum_over_shards_and_tiles():
    """Size-independent property at a large batch: the unnormalised gradient sums of two half batches add up to
    the full batch's (what the multi-GPU all-reduce relies on), and a row gather by a permutation leaves the
    reduced sums unchanged up to summation order."""
    from mappo_b200 import ops
    from mappo_b200.envs.spaces import Discrete
    from mappo_b200.algorithms.r_mappo.algorithm.r_actor_critic import R_Actor
    args = types.SimpleNamespace(hidden_size=64, layer_N=1, use_ReLU=False, use_feature_normalization=True,
                                 use_orthogonal=True, use_recurrent_policy=False, use_naive_recurrent_policy=False,
                                 recurrent_N=1, use_popart=False, gain=0.01, use_policy_active_masks=True, seed=1)
    torch.manual_seed(0)
    actor = R_Actor(args, make_space(18), Discrete(5), "cuda")
    B = 200_000
    g = torch.Generator(device="cuda").manual_seed(1)
    obs = torch.randn(B, 18, device="cuda", generator=g)
    acts = torch.randint(0, 5, (B, 1), device="cuda", generator=g).float()
    old = torch.full((B, 1), -1.6, device="cuda") + 0.1 * torch.randn(B, 1, device="cuda", generator=g)
    adv = torch.randn(B, device="cuda", generator=g)
    loss = ops.LossConfig(0.2, 0.01, 1.0, 10.0, 1, 1, 0, 1, 1, 0)
    P = actor.flat.numel()

    def run(sl, index=None):
        rows = (sl.stop - sl.start) if index is None else index.numel()
        nb = ops.update_num_blocks(actor.cfg, rows)
        gp = torch.empty(nb, P, device="cuda"); lp = torch.empty(nb, 4, dtype=torch.float64, device="cuda")
        if index is None:
            ops.actor_update(actor.cfg, loss, actor.flat, obs[sl], acts[sl], old[sl], adv[sl], None, None, gp, lp)
        else:
            ops.actor_update(actor.cfg, loss, actor.flat, obs, acts, old, adv, None, None, gp, lp, index=index)
        gs = torch.empty(P, device="cuda"); ls = torch.empty(4, dtype=torch.float64, device="cuda")
        ops.grad_reduce(gp, lp, nb, gs, ls)
        return gs.double(), ls
    g_all, l_all = run(slice(0, B))
    g_a, l_a = run(slice(0, B // 2 + 37)); g_b, l_b = run(slice(B // 2 + 37, B))
    scale = g_all.abs().max().item()
    assert (g_all - (g_a + g_b)).abs().max().item() < 2e-5 * scale
    np.testing.assert_allclose((l_a + l_b).cpu().numpy(), l_all.cpu().numpy(), rtol=1e-9)
    perm = torch.randperm(B, device="cuda")
    g_p, l_p = run(None, index=perm)
    assert (g_all - g_p).abs().max().item() < 2e-5 * scale
    np.testing.assert_allclose(l_p.cpu().numpy(), l_all.cpu().numpy(), rtol=1e-9)


def test_recurrent_forward_and_evaluate_vs_oracle(golden_dir):
    """GRU networks (rnn.py:24-77): one rollout step (values, argmax actions, log-probs, next rnn states) and the chunked
    evaluate_actions against the oracle on the reference fixture's weights and minibatch."""
    from oracle import policy as op
    fx = np.load(os.path.join(golden_dir, "ppo_gru_reference.npz"))
    args, pol, tr = build(fx)
    assert list(pol.actor.state_dict().keys()) == [k[7:] for k in fx.files if k.startswith("actor0.")]
    s = sample_of(fx)
    share_obs, obs, rnn_a, rnn_c, actions, masks, active = s[0], s[1], s[2], s[3], s[4], s[7], s[8]
    asd, csd = sd_from(fx, "actor0."), sd_from(fx, "critic0.")
    N = rnn_a.shape[0]
    # one step on the first N rows
    feats, h1 = op.actor_features(asd, obs[:N], rnn_a, masks[:N], args.use_ReLU, args.use_feature_normalization, True)
    a_ref, lp_ref, _ = op.act_forward(op.actor_logits(asd, feats, None), True)
    act, lp, h_new = pol.actor(obs[:N], rnn_a, masks[:N], None, deterministic=True)
    np.testing.assert_allclose(h_new.cpu().numpy(), h1.numpy(), rtol=1e-4, atol=2e-5)
    same = act.cpu().numpy() == a_ref.numpy()
    assert same.mean() > 0.99
    np.testing.assert_allclose(lp.cpu().numpy()[same], lp_ref.numpy()[same], rtol=1e-4, atol=2e-5)
    v_ref, hc1 = op.critic_values(csd, share_obs[:N], rnn_c, masks[:N], args.use_ReLU, args.use_feature_normalization, True)
    v, hc = pol.critic(share_obs[:N], rnn_c, masks[:N])
    np.testing.assert_allclose(v.cpu().numpy(), v_ref.numpy(), rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(hc.cpu().numpy(), hc1.numpy(), rtol=1e-4, atol=2e-5)
    # the whole chunked minibatch
    feats, _ = op.actor_features(asd, obs, rnn_a, masks, args.use_ReLU, args.use_feature_normalization, True)
    lp_e, ent_e = op.evaluate_actions(op.actor_logits(asd, feats, None), actions, active if args.use_policy_active_masks else None)
    elp, ent = pol.actor.evaluate_actions(obs, rnn_a, actions, masks, None, active)
    np.testing.assert_allclose(elp.cpu().numpy(), lp_e.numpy(), rtol=1e-4, atol=3e-5)
    np.testing.assert_allclose(ent.item(), ent_e.item(), rtol=1e-5)


def test_critic_on_the_buffers_replicated_share_obs_view(golden_dir):
    """R_Critic.forward given the buffer's share_obs slot [N, M, S] (M stride-0 copies of one stored row per env,
    utils/shared_buffer.py) evaluates each env once and repeats the value: identical to the reference's call on the
    materialised [N*M, S] concatenation (shared/mpe_runner.py:99,133-137)."""
    from mappo_b200.envs.spaces import Discrete
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    fx = np.load(os.path.join(golden_dir, "ppo_mlp_tanh.npz"))
    args = types.SimpleNamespace(**fx_args(fx))
    args.use_recurrent_policy = args.use_naive_recurrent_policy = False
    args.recurrent_N = 1; args.use_popart = False; args.use_orthogonal = True; args.seed = 1
    torch.manual_seed(5)
    pol = R_MAPPOPolicy(args, make_space(18), make_space(54), Discrete(5), device="cuda")
    N, M, S = 333, 3, 54
    uniq = torch.randn(N, S, device="cuda")
    view = uniq.unsqueeze(1).expand(N, M, S)
    assert view.stride(1) == 0
    v_view = pol.get_values(view, None, None)
    v_full = pol.get_values(view.reshape(N * M, S), None, None)
    assert v_view.shape == (N * M, 1)
    torch.testing.assert_close(v_view, v_full, rtol=0, atol=0)
    # a genuinely different row per agent is not mistaken for the view
    full3 = torch.randn(N, M, S, device="cuda")
    torch.testing.assert_close(pol.get_values(full3, None, None), pol.get_values(full3.reshape(N * M, S), None, None),
                               rtol=0, atol=0)


def test_raw_inputs_beyond_the_fp16_range_do_not_poison_the_tensor_core_path():
    """--use_feature_normalization off: raw observations are the one operand the network does not bound.  The fp16 `hi`
    halves of the tensor-core operands (input rows carry the scale 16) saturate at +-4094 instead of turning into
    inf -> NaN (DESIGN.md 4.1 "operand range"): an outlier row stays finite and the other rows are untouched."""
    from mappo_b200.envs.spaces import Discrete
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    args = types.SimpleNamespace(hidden_size=64, layer_N=1, use_ReLU=False, use_feature_normalization=False,
                                 use_orthogonal=True, use_recurrent_policy=False, use_naive_recurrent_policy=False,
                                 recurrent_N=1, use_popart=False, gain=0.01, use_policy_active_masks=True, seed=1,
                                 lr=7e-4, critic_lr=7e-4, opti_eps=1e-5, weight_decay=0.0)
    torch.manual_seed(0)
    pol = R_MAPPOPolicy(args, make_space(18), make_space(54), Discrete(5), device="cuda")
    obs = torch.randn(300, 18, device="cuda")
    share = torch.randn(300, 54, device="cuda")
    a0, lp0, _ = pol.actor(obs, None, None, None, deterministic=True)
    v0 = pol.get_values(share, None, None)
    obs2, share2 = obs.clone(), share.clone()
    obs2[7, 3] = 1e6; obs2[9, 0] = -3e38; share2[11, 5] = 1e7
    a1, lp1, _ = pol.actor(obs2, None, None, None, deterministic=True)
    v1 = pol.get_values(share2, None, None)
    assert torch.isfinite(lp1).all() and torch.isfinite(v1).all()
    keep = torch.ones(300, dtype=torch.bool, device="cuda"); keep[[7, 9]] = False
    assert torch.equal(lp1[keep], lp0[keep]) and torch.equal(a1[keep], a0[keep])
    keepv = torch.ones(300, dtype=torch.bool, device="cuda"); keepv[11] = False
    assert torch.equal(v1[keepv], v0[keepv])
