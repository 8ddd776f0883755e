"""GPU: the update kernels' bulk-copy paths against their per-thread load paths (csrc/mlp_tc.cu).

With contiguous, 16-byte aligned rows the fused update kernels bring a tile's input rows (`raw_ok`) and its per-row loss
scalars (`rowp_ok`: actions / old log-probs / advantages / active masks for the actor; returns / old values / active
masks of the `group` terms for the critic, group <= 4) into shared memory by bulk copies a tile ahead; anything else
-- a minibatch index, a misaligned view, a larger group, the ragged last tile -- takes plain loads.  Both are the
same arithmetic on the same numbers in the same order, so the gradient sums and loss sums must be IDENTICAL, bit for
bit.  (Parity of each path with the oracle: tests/test_gpu_large_batch.py runs the bulk paths, tests/test_gpu_policy.py
the indexed ones.)"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(group, B, seed=0, flags=""):
    from test_gpu_large_batch import _policy, _mlp_batch
    from mappo_b200.envs.spaces import Discrete
    D, S, A = 18, 54, 5
    rows_c = B // group
    a, pol, tr, g = _policy(flags, D, S, Discrete(A), seed=seed)
    obs, share, acts, old_logp, adv, active, vp, ret = _mlp_batch(pol, g, rows_c * group, D, A, rows_c, S)
    tr.value_normalizer.update(-4.0 + 2.0 * torch.randn(300, 1, device="cuda", generator=g))
    return a, pol, tr, (obs, share, acts, old_logp, adv, active, vp, ret)


def _shift(x):
    """the same values at an address that is 4 bytes past a 16-byte boundary: no bulk copy may touch it"""
    flat = torch.empty(x.numel() + 1, dtype=x.dtype, device=x.device)
    flat[1:] = x.reshape(-1)
    y = flat[1:].view(x.shape)
    assert y.data_ptr() % 16 == 4
    return y


def _actor(pol, tr, obs, acts, old_logp, adv, active, index=None):
    from mappo_b200 import ops
    rows = obs.shape[0] if index is None else index.numel()
    ws = tr._workspace(rows, rows)
    Pa = pol.actor.flat.numel()
    ops.actor_update(pol.actor.cfg, tr._loss_config(), pol.actor.flat, obs, acts, old_logp, adv, active, None, ws["gpa"],
                     ws["lpa"], index=index)
    ops.grad_reduce(ws["gpa"], ws["lpa"], ws["nba"], ws["gsum"][:Pa], ws["lsum"][:4])
    return ws["gsum"][:Pa].clone(), ws["lsum"][:4].clone()


def _critic(pol, tr, share, vp, ret, active, group, index=None):
    from mappo_b200 import ops
    rows = share.shape[0] if index is None else index.numel()
    ws = tr._workspace(rows * group, rows)
    Pa = pol.actor.flat.numel()
    ops.critic_update(pol.critic.cfg, tr._loss_config(), pol.critic.flat, share, vp, ret, active, tr.value_normalizer.stats,
                      ws["gpc"], ws["lpc"], group=group, index=index)
    ops.grad_reduce(ws["gpc"], ws["lpc"], ws["nbc"], ws["gsum"][Pa:], ws["lsum"][4:])
    return ws["gsum"][Pa:].clone(), ws["lsum"][4:].clone()


def _same(x, y, what):
    for a, b, name in ((x[0], y[0], "gradient sums"), (x[1], y[1], "loss sums")):
        assert torch.isfinite(a).all() and a.abs().max() > 0, what
        assert torch.equal(a, b), "%s: %s differ by %.3e (max |.| %.3e)" % (
            what, name, (a.double() - b.double()).abs().max().item(), a.abs().max().item())


@pytest.mark.parametrize("B", [40 * 128, 40 * 128 + 77])
def test_actor_update_bulk_copies_equal_plain_loads(B):
    """full tiles only / a ragged last tile (which always takes plain loads) x {bulk, identity index, misaligned views}"""
    a, pol, tr, (obs, share, acts, old_logp, adv, active, vp, ret) = _setup(1, B)
    bulk = _actor(pol, tr, obs, acts, old_logp, adv, active)
    idx = torch.arange(B, device="cuda", dtype=torch.int64)
    _same(bulk, _actor(pol, tr, obs, acts, old_logp, adv, active, index=idx), "identity index")
    _same(bulk, _actor(pol, tr, obs, _shift(acts), old_logp, adv, active), "misaligned actions (rows still by bulk copy)")
    _same(bulk, _actor(pol, tr, _shift(obs), acts, old_logp, _shift(adv), active), "misaligned rows and advantages")
    tr._use_policy_active_masks = False
    _same(_actor(pol, tr, obs, acts, old_logp, adv, None), _actor(pol, tr, obs, acts, old_logp, adv, None, index=idx),
          "no active masks")


@pytest.mark.parametrize("group", [1, 3, 4, 5])
def test_critic_update_bulk_copies_equal_plain_loads(group):
    """group <= 4: the loss inputs of a tile arrive by bulk copy; group 5 never does (same kernel, plain loads)"""
    rows_c = 24 * 128 + 31
    a, pol, tr, (obs, share, acts, old_logp, adv, active, vp, ret) = _setup(group, rows_c * group)
    bulk = _critic(pol, tr, share, vp, ret, active, group)
    _same(bulk, _critic(pol, tr, share, _shift(vp), ret, active, group), "misaligned old values")
    _same(bulk, _critic(pol, tr, _shift(share), vp, _shift(ret), _shift(active), group), "misaligned rows, returns, masks")
    if group == 1:
        idx = torch.arange(rows_c, device="cuda", dtype=torch.int64)
        _same(bulk, _critic(pol, tr, share, vp, ret, active, 1, index=idx), "identity index")
