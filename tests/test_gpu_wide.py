"""GPU: the wide-input centralised critic (BASELINE config C5: share_obs = 10 agents x 46 = 460 columns) -- the first
layer runs as two streaming tensor-core GEMMs (csrc/mlp_wide.cu) around the fused 64-wide kernel.  Values, value
loss and PRE-clip gradients against the oracle (oracle/policy.py, torch CPU fp32 + autograd) on the same inputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(S, B, flags="", seed=0, D=46, A=5):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.spaces import Box, Discrete
    from mappo_b200.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    from mappo_b200.algorithms.r_mappo.r_mappo import R_MAPPO
    a = parse_args(("--scenario_name simple_spread --num_agents 10 --ppo_epoch 1 " + flags).split())
    torch.manual_seed(seed)
    box = lambda n: Box(-np.inf, np.inf, (n,), np.float32)
    pol = R_MAPPOPolicy(a, box(D), box(S), Discrete(A), device="cuda")
    g = torch.Generator(device="cuda").manual_seed(seed + 1)
    for net in (pol.actor, pol.critic):                     # LayerNorm affines and biases away from (1, 0)
        net.flat.add_(0.05 * torch.randn(net.flat.shape, device="cuda", generator=g))
    tr = R_MAPPO(a, pol, device="cuda")
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    share = r(B, S) * (0.5 + r(1, S).abs()) + 0.3 * r(1, S)
    sample = (share, r(B, D) * (0.5 + r(1, D).abs()), None, None,
              torch.randint(0, A, (B, 1), device="cuda", generator=g).float(),
              0.5 * r(B, 1), 0.5 * r(B, 1) + 0.2, torch.ones(B, 1, device="cuda"),
              (torch.rand(B, 1, device="cuda", generator=g) > 0.1).float(), -1.6 + 0.1 * r(B, 1), r(B, 1), None)
    return a, pol, tr, sample


def _oracle_critic(a, pol, tr, sample):
    from oracle import policy as op
    from oracle.buffer import OracleValueNorm
    sd = {k: v.detach().cpu().clone().requires_grad_("fc_h" not in k) for k, v in pol.critic.state_dict().items()}
    share, vp, ret, am = (sample[i].cpu().numpy() for i in (0, 5, 6, 8))
    B = share.shape[0]
    z = np.zeros((B, 1, 64), np.float32)
    values, _ = op.critic_values(sd, share, z, np.ones((B, 1), np.float32), a.use_ReLU, a.use_feature_normalization, False)
    vn = None
    if tr.value_normalizer is not None:
        vn = OracleValueNorm(); vn.load(tr.value_normalizer.stats.cpu().numpy())
    vl = op.value_loss(values, vp, ret, am, a.clip_param, vn, a.use_huber_loss, a.huber_delta, a.use_clipped_value_loss,
                       a.use_value_active_masks)
    keys = [k for k in sd if "fc_h" not in k]
    grads = torch.autograd.grad(vl * a.value_loss_coef, [sd[k] for k in keys])
    return values.detach().numpy(), vl.item(), dict(zip(keys, grads))


@pytest.mark.parametrize("S,B,flags", [(460, 1000, ""), (460, 333, "--use_ReLU --use_valuenorm"), (600, 700, ""),
                                       (100, 129, "--use_feature_normalization"),
                                       # beyond ~630 columns W1' is staged 32 output columns at a time (one pass over x per part):
                                       (990, 700, ""), (1120, 300, "--use_ReLU --use_feature_normalization"),
                                       (2000, 300, ""),                 # 16 columns at a time
                                       (64, 500, "")])                  # exactly 64: no padding column left in the fused kernel
def test_wide_critic_values_loss_and_gradients(S, B, flags):
    a, pol, tr, sample = _setup(S, B, flags)
    want_v, want_vl, want_g = _oracle_critic(a, pol, tr, sample)
    got_v, _ = pol.critic(sample[0], None, None)
    np.testing.assert_allclose(got_v.cpu().numpy(), want_v, rtol=1e-4, atol=2e-5)
    vl, cgn, pl, ent, agn, ratio = tr.ppo_update(sample)
    np.testing.assert_allclose(vl.item(), want_vl, rtol=1e-4, atol=1e-6)
    ws = next(iter(tr._ws.values()))
    Pa = pol.actor.flat.numel()
    for k, gv in pol.critic.unflatten(ws["grad"][Pa:]).items():
        w = want_g[k].numpy()
        np.testing.assert_allclose(gv.cpu().numpy(), w, rtol=1e-4, atol=1e-4 * np.abs(w).max() + 1e-9, err_msg=k)


def test_wide_critic_strided_rows_and_row_gather():
    """Rows addressed through an index (minibatch gather, several samples per share_obs row) and rows that are column
    slices of a wider storage (unaligned: scalar loads) give the same sums as the plain layout."""
    from mappo_b200 import ops
    a, pol, tr, sample = _setup(460, 900, "")
    cfg, flat = pol.critic.cfg, pol.critic.flat
    share, vp, ret, am = sample[0], sample[5].reshape(-1), sample[6].reshape(-1), sample[8].reshape(-1)
    loss = tr._loss_config()
    vn = tr.value_normalizer.stats
    P = flat.numel()

    def run(x, rows, **kw):
        nb = ops.update_num_blocks(cfg, rows)
        gp = torch.empty(nb, P, device="cuda"); lp = torch.empty(nb, 4, dtype=torch.float64, device="cuda")
        ops.critic_update(cfg, loss, flat, x, kw.pop("vp", vp), kw.pop("ret", ret), kw.pop("am", am), vn, gp, lp, **kw)
        gs = torch.empty(P, device="cuda"); ls = torch.empty(4, dtype=torch.float64, device="cuda")
        ops.grad_reduce(gp, lp, nb, gs, ls)
        return gs.double(), ls
    g0, l0 = run(share, 900)
    scale = g0.abs().max().item()
    wide = torch.zeros(900, 460 + 7, device="cuda")
    wide[:, 3:463] = share
    g1, l1 = run(wide[:, 3:463], 900)
    assert (g0 - g1).abs().max().item() < 2e-5 * scale
    np.testing.assert_allclose(l1.cpu().numpy(), l0.cpu().numpy(), rtol=1e-9)
    # 3 samples per share row, visited through a permutation
    rep = lambda t: t.repeat_interleave(3)
    perm = torch.randperm(2700, device="cuda")
    g2, l2 = run(share, 2700, vp=rep(vp), ret=rep(ret), am=rep(am), index=perm, share_rows_div=3)
    assert (3 * g0 - g2).abs().max().item() < 6e-5 * scale
    np.testing.assert_allclose(l2.cpu().numpy(), 3 * l0.cpu().numpy(), rtol=1e-9)
    # group = 3: the same thing without an index
    g3, l3 = run(share, 900, vp=rep(vp), ret=rep(ret), am=rep(am), group=3)
    assert (3 * g0 - g3).abs().max().item() < 6e-5 * scale


@pytest.mark.parametrize("S,B,flags", [(460, 1000, ""), (460, 12800 + 77, "--use_ReLU"), (100, 129, "--use_feature_normalization"),
                                       (96, 300, ""), (460, 51205, ""), (300, 3000, "")])
def test_packed_inputs_give_the_same_update(S, B, flags):
    """num_mini_batch == 1: the share_obs rows are packed once per iteration into fp16-pair operand tiles
    (mappo_critic_pack_inputs); every epoch's forward (fc1_packed_kernel) and first-layer weight gradient
    (wgrad1_packed_kernel, 2-stage ring: 51205 rows give 11 tiles per CTA, in_dim 300 three column groups, the last 48 columns wide) stream them.  Loss sums and gradient sums
    equal the unpacked path's to the precision of the operand pairs (2^-22 relative per product), over several tiles per
    CTA, a ragged last tile and strided rows."""
    from mappo_b200 import ops
    a, pol, tr, sample = _setup(S, B, flags)
    cfg, flat = pol.critic.cfg, pol.critic.flat
    share, vp, ret, am = sample[0], sample[5].reshape(-1), sample[6].reshape(-1), sample[8].reshape(-1)
    loss, vn, P = tr._loss_config(), tr.value_normalizer.stats, flat.numel()

    def run(x, packed):
        nb = ops.update_num_blocks(cfg, B)
        gp = torch.empty(nb, P, device="cuda"); lp = torch.empty(nb, 4, dtype=torch.float64, device="cuda")
        ops.critic_update(cfg, loss, flat, x, vp, ret, am, vn, gp, lp, packed=packed)
        gs = torch.empty(P, device="cuda"); ls = torch.empty(4, dtype=torch.float64, device="cuda")
        ops.grad_reduce(gp, lp, nb, gs, ls)
        torch.cuda.synchronize()
        return gs.double(), ls
    g0, l0 = run(share, None)
    packed = ops.critic_pack_inputs(cfg, share)
    assert packed is not None and packed.dtype == torch.uint8
    for _ in range(2):                                                   # the packed tiles survive an update (epochs)
        g1, l1 = run(share, packed)
        scale = g0.abs().max().item()
        assert (g0 - g1).abs().max().item() < 2e-5 * scale, (g0 - g1).abs().max().item() / scale
        np.testing.assert_allclose(l1.cpu().numpy(), l0.cpu().numpy(), rtol=1e-6)
    wide = torch.zeros(B, S + 8, device="cuda")
    wide[:, 4:4 + S] = share
    packed = ops.critic_pack_inputs(cfg, wide[:, 4:4 + S])
    g2, l2 = run(wide[:, 4:4 + S], packed)
    assert (g0 - g2).abs().max().item() < 2e-5 * scale


def test_packing_is_refused_with_a_row_gather():
    from mappo_b200 import ops
    a, pol, tr, sample = _setup(460, 600, "")
    cfg, flat = pol.critic.cfg, pol.critic.flat
    share, vp, ret, am = sample[0], sample[5].reshape(-1), sample[6].reshape(-1), sample[8].reshape(-1)
    packed = ops.critic_pack_inputs(cfg, share)
    nb, P = ops.update_num_blocks(cfg, 600), flat.numel()
    gp = torch.empty(nb, P, device="cuda"); lp = torch.empty(nb, 4, dtype=torch.float64, device="cuda")
    with pytest.raises(AssertionError):
        ops.critic_update(cfg, tr._loss_config(), flat, share, vp, ret, am, tr.value_normalizer.stats, gp, lp,
                          index=torch.randperm(600, device="cuda"), packed=packed)



@pytest.mark.parametrize("D,A,B,flags", [(64, 5, 300, ""), (70, 5, 1000, ""), (118, 5, 12800 + 77, "--use_ReLU"), (70, 12, 700, ""),
                                         (100, 5, 129, "--use_feature_normalization")])
def test_wide_actor_logp_loss_and_gradients(D, A, B, flags):
    """Observations wider than 64 columns (simple_spread from 12 agents up: 4 + 2 L + 4 (M - 1)): the actor takes the same
    three-kernel chain as the wide critic (ABI v4).  Log-probs / entropy of given actions, the policy loss and the
    PRE-clip gradients against the oracle on the same inputs; A = 12 takes the generic head path."""
    from oracle import policy as op
    a, pol, tr, sample = _setup(100, B, flags, D=D, A=A)
    obs, actions, am, old_lp, adv = (sample[i] for i in (1, 4, 8, 9, 10))
    old_lp = old_lp - (np.log(A) - 1.6)
    sample = sample[:9] + (old_lp, adv, None)
    sd = {k: v.detach().cpu().clone().requires_grad_("fc_h" not in k) for k, v in pol.actor.state_dict().items()}
    z, ones = np.zeros((B, 1, 64), np.float32), np.ones((B, 1), np.float32)
    feats, _ = op.actor_features(sd, obs.cpu().numpy(), z, ones, a.use_ReLU, a.use_feature_normalization, False)
    logits = op.actor_logits(sd, feats, None)
    logp, ent = op.evaluate_actions(logits, actions.cpu().numpy(), am.cpu().numpy() if a.use_policy_active_masks else None)
    pl, ratio = op.policy_loss(logp, old_lp.cpu().numpy(), adv.cpu().numpy(), am.cpu().numpy(), a.clip_param,
                               a.use_policy_active_masks)
    keys = [k for k in sd if "fc_h" not in k]
    want_g = dict(zip(keys, torch.autograd.grad(pl - ent * a.entropy_coef, [sd[k] for k in keys])))
    got_lp, got_ent = pol.actor.evaluate_actions(obs, None, actions, None, None, am)
    np.testing.assert_allclose(got_lp.cpu().numpy(), logp.detach().numpy(), rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(got_ent.item(), ent.item(), rtol=1e-4)
    acts, lps, _ = pol.actor(obs, None, None, deterministic=True)      # rollout kernel: mode() of the same logits
    assert (acts.cpu().numpy()[:, 0] == logits[0].detach().numpy().argmax(-1)).mean() > 0.999
    vl, cgn, got_pl, got_entropy, agn, got_ratio = tr.ppo_update(sample)
    np.testing.assert_allclose(got_pl.item(), pl.item(), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(got_ratio.item(), ratio.mean().item(), rtol=1e-4)
    ws = next(iter(tr._ws.values()))
    Pa = pol.actor.flat.numel()
    for k, gv in pol.actor.unflatten(ws["grad"][:Pa]).items():
        w = want_g[k].numpy()
        np.testing.assert_allclose(gv.cpu().numpy(), w, rtol=1e-4, atol=1e-4 * np.abs(w).max() + 1e-9, err_msg=k)
