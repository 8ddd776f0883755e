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


def test_update_kernels_are_bit_reproducible_run_to_run():
    """Every buffer of a tile is reused by the next one under mbarrier / bulk-copy ordering (next tile's rows, its loss
    scalars, the operand tiles the background weight-gradient MMAs still read): an ordering hole shows up as run-to-run
    differences.  Six runs of both update kernels at ~17 tiles per CTA, and of the recurrent chain, must agree bit for bit."""
    B = 148 * 17 * 128 + 55
    a, pol, tr, (obs, share, acts, old_logp, adv, active, vp, ret) = _setup(3, B - B % 3)
    Bc = (B - B % 3) // 3
    first_a = first_c = None
    for _ in range(6):
        ga = _actor(pol, tr, obs, acts, old_logp, adv, active)
        gc = _critic(pol, tr, share, vp, ret, active, 3)
        if first_a is None:
            first_a, first_c = ga, gc
        else:
            _same(first_a, ga, "actor update, repeated run")
            _same(first_c, gc, "critic update (%d rows x group 3), repeated run" % Bc)


def test_recurrent_update_is_bit_reproducible_run_to_run():
    from test_gpu_large_batch import _policy
    from mappo_b200.envs.spaces import MultiDiscrete
    D, S, L, Nc = 21, 42, 10, 148 * 128 * 3 + 17
    Bn = L * Nc
    a, pol, tr, g = _policy("--algorithm_name rmappo --data_chunk_length 10 --use_ReLU", D, S, MultiDiscrete([[0, 4], [0, 9]]))
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    obs, share = r(Bn, D), r(Bn, S)
    rnn_a, rnn_c = 0.5 * r(Nc, 1, 64), 0.5 * r(Nc, 1, 64)
    masks = (torch.rand(Bn, 1, device="cuda", generator=g) > 0.08).float()
    acts = torch.stack([torch.randint(0, 5, (Bn,), device="cuda", generator=g),
                        torch.randint(0, 10, (Bn,), device="cuda", generator=g)], -1).float()
    old_logp = torch.stack([-1.6 + 0.1 * r(Bn), -2.3 + 0.1 * r(Bn)], -1)
    adv, vp, ret = r(Bn, 1), 0.5 * r(Bn, 1), -5.0 + 3.0 * r(Bn, 1)
    active = (torch.rand(Bn, 1, device="cuda", generator=g) > 0.1).float()
    sample = (share, obs, rnn_a, rnn_c, acts, vp, ret, masks, active, old_logp, adv, None)
    flat0 = (pol.actor.flat.clone(), pol.critic.flat.clone())
    vn0 = tr.value_normalizer.stats.clone()
    first = None
    for _ in range(4):
        pol.actor.flat.copy_(flat0[0]); pol.critic.flat.copy_(flat0[1]); tr.value_normalizer.stats.copy_(vn0)
        tr.ppo_update(sample, update_actor=False)
        ws = next(iter(tr._ws.values()))
        grad = ws["grad"].clone()
        if first is None:
            first = grad
        else:
            assert torch.equal(first, grad), "recurrent update differs run to run by %.3e" % (first - grad).abs().max().item()
