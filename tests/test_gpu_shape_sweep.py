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
