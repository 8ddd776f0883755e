"""GPU: the fused update kernels across input widths, head sizes, batch sizes and option flags.

The reference-recorded fixtures pin the shapes of the MPE scenarios (tests/test_gpu_policy.py); the benchmark-size tests pin
C2 / C5 / C4 (tests/test_gpu_large_batch.py).  The kernels, however, take their shapes from template parameters and
padding rules -- input tile 32 or 64 columns wide with ONE padding column reserved for the bias-gradient ones column
(in_dim 31 -> 32-wide tile, in_dim 32 -> 64-wide), heads of <= 8 logits on the register loss path and the fused
LayerNorm / logit exchange, wider heads on the row functions, critic `group` <= 4 through the bulk-copied loss inputs
and above through plain loads, ragged last tiles, a single row.  Every case runs one actor and one critic update and
compares losses and PRE-clip gradients with the float64 autograd oracle of tests/test_gpu_large_batch.py
(rtol 1e-4, atol 1e-4 max|g| per parameter tensor)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

# (in_dim, critic in_dim, logits, critic group, actor rows per group row count, flags)
CASES = [
    (1, 2, 2, 1, 1, ""),                                            # the smallest everything, ONE row
    (7, 14, 3, 2, 127, "--use_ReLU"),                               # Tanh (store_false), one ragged tile
    (8, 24, 5, 3, 128, ""),                                         # ReLU, exactly one tile
    (17, 31, 8, 4, 129, "--use_ReLU"),                              # 8 logits: the widest register head; group 4: bulk limit
    (31, 32, 9, 5, 1000, "--use_ReLU"),                             # 31 / 32: last 32-wide tile / first 64-wide; 9 logits: row path
    (32, 63, 16, 1, 777, ""),                                       # 64-wide actor tile; 63: last fused critic width; 16 logits
    (47, 50, 5, 2, 5000, "--use_ReLU --use_feature_normalization"),  # raw (un-normalised) inputs
    (63, 54, 6, 3, 4097, "--use_ReLU --use_huber_loss"),            # mse value loss (store_false)
    (18, 54, 5, 3, 2048, "--use_ReLU --use_clipped_value_loss --use_valuenorm"),   # unclipped loss, no ValueNorm
    (18, 54, 5, 3, 2049, "--use_ReLU --use_policy_active_masks --use_value_active_masks"),
]


@pytest.mark.parametrize("D,S,A,group,rows_c,flags", CASES)
def test_update_gradients_across_shapes(D, S, A, group, rows_c, flags):
    from test_gpu_large_batch import (_policy, _mlp_batch, _mask_kinks, _actor_oracle64, _critic_oracle64, _check, _threads)
    from mappo_b200.envs.spaces import Discrete
    from mappo_b200 import ops
    B = rows_c * group
    a, pol, tr, g = _policy(flags, D, S, Discrete(A), seed=D + A)
    # the fused tcgen05 kernels -- except the one shape whose shared-memory map outgrows the SM (64-wide input tile AND a
    # 16-logit head: 229 KB), which the library sends to its fp32 tile kernels; same parity bar there
    paths = (ops.kernel_path(pol.actor.cfg), ops.kernel_path(pol.critic.cfg))
    assert paths == ((0, 1) if (D >= 32 and A > 8) else (1, 1)), paths
    obs, share, acts, old_logp, adv, active, vp, ret = _mlp_batch(pol, g, B, D, A, rows_c, S)
    active = _mask_kinks(a, pol, obs, share, group, active) if B > 64 else torch.ones_like(active)
    vn_before = None
    if tr.value_normalizer is not None:
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
    _check(pol.actor, ws["grad"][:Pa], ga, "actor update D=%d A=%d rows=%d" % (D, A, B))
    _check(pol.critic, ws["grad"][Pa:], gc, "critic update S=%d group=%d rows=%d" % (S, group, rows_c))


@pytest.mark.parametrize("D,S,A,group,rows_c,flags", CASES)
def test_rollout_forward_across_shapes(D, S, A, group, rows_c, flags):
    """The rollout kernels (MODE 0 / 1) at the same shapes: values, deterministic and sampled actions, their log-probs --
    against the float32 oracle forward (oracle/policy.py), with the tolerances of tests/test_gpu_policy.py."""
    from test_gpu_large_batch import _policy
    from mappo_b200.envs.spaces import Discrete
    from oracle import policy as op
    B = rows_c * group
    a, pol, tr, g = _policy(flags, D, S, Discrete(A), seed=D + A)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    obs = r(B, D) * (0.4 + r(1, D).abs()) + 0.3 * r(1, D)
    share = r(rows_c, S) * (0.4 + r(1, S).abs()) + 0.3 * r(1, S)
    cpu = lambda net: {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    asd, csd = cpu(pol.actor), cpu(pol.critic)
    z = lambda n: np.zeros((n, 1, 64), np.float32)
    ones = lambda n: np.ones((n, 1), np.float32)
    feats, _ = op.actor_features(asd, obs.cpu().numpy(), z(B), ones(B), a.use_ReLU, a.use_feature_normalization, False)
    logits = op.actor_logits(asd, feats, None)
    v_ref, _ = op.critic_values(csd, share.cpu().numpy(), z(rows_c), ones(rows_c), a.use_ReLU, a.use_feature_normalization, False)
    values = pol.get_values(share, None, None)
    np.testing.assert_allclose(values.cpu().numpy().reshape(-1), np.asarray(v_ref).reshape(-1), rtol=1e-4, atol=2e-5)
    a_ref, lp_ref, _ = op.act_forward(logits, True)
    act, lp, _ = pol.actor(obs, None, None, None, deterministic=True)
    same = act.cpu().numpy() == a_ref.numpy()
    assert same.mean() > 0.995 or B < 200
    np.testing.assert_allclose(lp.cpu().numpy()[same], lp_ref.numpy()[same], rtol=1e-4, atol=2e-5)
    pol.actor.sample_seed = 7
    acts = torch.empty(B, 1, device="cuda"); lps = torch.empty_like(acts)
    pol.actor.forward_into(obs, acts, lps, step=3, row_offset=11)
    a_s, lp_s, margin = op.act_forward(logits, False, seed=7, step=3, row_offset=11)
    diff = acts.cpu().numpy() != a_s.numpy()
    assert not (diff & (margin.numpy() > 1e-5)).any(), "sampled action differs away from a CDF boundary"
    np.testing.assert_allclose(lps.cpu().numpy()[~diff], lp_s.numpy()[~diff], rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("S,group,rows_c,packed", [(64, 1, 300, False), (65, 3, 1000, True), (100, 2, 129, False),
                                                    (250, 3, 1500, False), (256, 1, 1500, True), (257, 4, 700, False),
                                                    (481, 5, 900, True), (513, 2, 640, False), (1000, 3, 333, True)])
def test_wide_first_layer_across_widths(S, group, rows_c, packed):
    """The wide first layer (csrc/mlp_wide.cu) at widths on both sides of its column-group boundaries (one weight-gradient
    column group holds <= 256 columns) and of the packed tiles' 16-column chunks: critic update vs the float64 oracle."""
    from test_gpu_large_batch import _policy, _critic_oracle64, _check, _threads
    from mappo_b200 import ops
    from mappo_b200.envs.spaces import Discrete
    B = rows_c * group
    a, pol, tr, g = _policy("--use_ReLU", 18, S, Discrete(5), seed=S)
    assert ops.kernel_path(pol.critic.cfg) == 2
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    share = r(rows_c, S) * (0.4 + r(1, S).abs()) + 0.3 * r(1, S)
    obs = r(B, 18)
    acts = torch.randint(0, 5, (B, 1), device="cuda", generator=g).float()
    old_logp = -1.6 + 0.1 * r(B, 1)
    adv, vp, ret = r(B), 0.5 * r(B), -5.0 + 3.0 * r(B)
    active = (torch.rand(B, device="cuda", generator=g) > 0.1).float()
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    vn_before = tr.value_normalizer.stats.cpu().numpy().copy()
    torch.set_num_threads(_threads())
    try:
        vl, gc = _critic_oracle64(a, pol, tr, share, group, vp.view(B, 1), ret.view(B, 1), active.view(B, 1), vn_before)
    finally:
        torch.set_num_threads(1)
    pk = ops.critic_pack_inputs(pol.critic.cfg, share) if packed else None
    ws = tr._update(obs, share, group, acts, old_logp, adv, vp, ret, active, None, B, packed=pk)
    np.testing.assert_allclose(tr._info[4].item(), vl, rtol=1e-4, atol=1e-6)
    _check(pol.critic, ws["grad"][pol.actor.flat.numel():], gc, "wide critic S=%d group=%d rows=%d" % (S, group, rows_c))


@pytest.mark.parametrize("D,S,L,Nc,heads,relu,rN", [
    (7, 14, 1, 77, [5], False, 1),              # one-step chunks, a ragged half tile of chunks
    (21, 42, 10, 129, [5, 10], True, 1),        # C4's networks, one chunk over a tile
    (21, 42, 25, 64, [5, 10], False, 2),        # two stacked GRU layers, whole-episode chunks
    (40, 54, 5, 1000, [3], False, 1),           # 64-wide input tiles on both base kernels
    (66, 132, 10, 200, [5], True, 1),           # wide first layer in front of the GRU
])
def test_recurrent_update_across_shapes(D, S, L, Nc, heads, relu, rN):
    """The recurrent chain (base forward, GRU sweeps, head, ONE dual weight-gradient pass per layer, base backward) at chunk
    counts / lengths / widths off the benchmark's: losses and PRE-clip gradients vs float64 autograd through the oracle."""
    from test_gpu_large_batch import _policy, _mask_kinks, _actor_oracle64, _critic_oracle64, _check, _threads
    from mappo_b200.envs.spaces import Discrete, MultiDiscrete
    B = L * Nc
    space = Discrete(heads[0]) if len(heads) == 1 else MultiDiscrete([[0, h - 1] for h in heads])
    a, pol, tr, g = _policy("--algorithm_name rmappo --data_chunk_length %d --recurrent_N %d%s" % (L, rN, "" if relu else " --use_ReLU"),
                            D, S, space, seed=D + L)
    assert bool(a.use_ReLU) == relu
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    obs = r(B, D) * (0.4 + r(1, D).abs())
    share = r(B, S) * (0.4 + r(1, S).abs())
    rnn_a, rnn_c = 0.5 * r(Nc, rN, 64), 0.5 * r(Nc, rN, 64)
    masks = (torch.rand(B, 1, device="cuda", generator=g) > 0.08).float()
    acts = torch.stack([torch.randint(0, h, (B,), device="cuda", generator=g) for h in heads], -1).float()
    old_logp = torch.stack([-np.log(h) + 0.1 * r(B) for h in heads], -1)
    adv, vp, ret = r(B, 1), 0.5 * r(B, 1), -5.0 + 3.0 * r(B, 1)
    active = (torch.rand(B, 1, device="cuda", generator=g) > 0.1).float()
    if D < 64:
        active = _mask_kinks(a, pol, obs, share, 1, active)
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    vn_before = tr.value_normalizer.stats.cpu().numpy().copy()
    torch.set_num_threads(_threads())
    try:
        pl, ent, ratio, ga = _actor_oracle64(a, pol, obs, acts, old_logp, adv, active, rnn=rnn_a, masks=masks)
        vl, gc = _critic_oracle64(a, pol, tr, share, 1, vp, ret, active, vn_before, rnn=rnn_c, masks=masks)
    finally:
        torch.set_num_threads(1)
    got = tr.ppo_update((share, obs, rnn_a, rnn_c, acts, vp, ret, masks, active, old_logp, adv, None))
    np.testing.assert_allclose(got[2].item(), pl, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(got[3].item(), ent, rtol=1e-5)
    np.testing.assert_allclose(got[0].item(), vl, rtol=1e-4, atol=1e-6)
    ws = next(iter(tr._ws.values()))
    Pa = pol.actor.flat.numel()
    _check(pol.actor, ws["grad"][:Pa], ga, "recurrent actor D=%d L=%d Nc=%d rN=%d" % (D, L, Nc, rN))
    _check(pol.critic, ws["grad"][Pa:], gc, "recurrent critic S=%d L=%d Nc=%d rN=%d" % (S, L, Nc, rN))
