"""GPU: returns / GAE / ValueNorm / advantage kernels against the reference fixtures (bit-exact) and the oracle."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev(x, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype).cuda().contiguous()


@pytest.mark.parametrize("tag", ["g%dp%dn%d" % (g, p, n) for g in (1, 0) for p in (0, 1) for n in (1, 0)])
def test_returns_bit_exact_vs_reference(tag, golden_dir):
    from mappo_b200 import ops
    fx = np.load(os.path.join(golden_dir, "gae.npz"))
    use_gae, proper, norm = int(tag[1]), int(tag[3]), int(tag[5])
    vp = dev(fx[tag + "_value_preds_in"]); ret = torch.zeros_like(vp)
    vn = dev(fx[tag + "_vn"]) if norm else None
    raw, s3 = ops.gae_returns(dev(fx[tag + "_rewards"]), vp, ret, dev(fx[tag + "_masks"]), dev(fx[tag + "_next_value"]),
                              float(fx["gamma"]), float(fx["gae_lambda"]), use_gae=use_gae,
                              use_proper_time_limits=proper, bad_masks=dev(fx[tag + "_bad_masks"]),
                              active_masks=dev(fx[tag + "_active_masks"]), vn_stats=vn)
    adv = ops.adv_normalize(raw, s3)
    np.testing.assert_array_equal(ret[:-1].cpu().numpy(), fx[tag + "_returns"][:-1])      # north_star: 1e-6; here exact
    if use_gae:
        np.testing.assert_array_equal(vp.cpu().numpy(), fx[tag + "_value_preds_out"])
    np.testing.assert_array_equal(raw.cpu().numpy(), fx[tag + "_adv_raw"])
    act = fx[tag + "_active_masks"][:-1] != 0
    np.testing.assert_allclose(adv.cpu().numpy(), fx[tag + "_adv"], rtol=1e-6, atol=1e-6)
    assert int(s3[2].item()) == int(act.sum())


def test_valuenorm_matches_reference(golden_dir):
    from mappo_b200.utils.valuenorm import ValueNorm
    fx = np.load(os.path.join(golden_dir, "vnorm.npz"))
    vn = ValueNorm(1)
    np.testing.assert_array_equal(vn.denormalize(fx["x"]).cpu().numpy(), fx["fresh_denormalized"])
    for x, want in zip(fx["batches"], fx["stats_seq"]):
        vn.update(x)
        np.testing.assert_allclose(vn.stats.cpu().numpy(), want, rtol=2e-7, atol=0)
    vn.stats.copy_(torch.as_tensor(fx["stats_seq"][-1]))
    np.testing.assert_array_equal(vn.normalize(fx["x"]).cpu().numpy(), fx["normalized"])
    np.testing.assert_array_equal(vn.denormalize(fx["x"]).cpu().numpy(), fx["denormalized"])
    sd = vn.state_dict()
    assert set(sd) == {"running_mean", "running_mean_sq", "debiasing_term"} and sd["running_mean"].shape == (1,)


def test_gae_large_vs_oracle_and_linearity():
    """Full C2 size (T=25, R=196608): oracle equality on a slice, plus a size-independent property:
    with lambda=1, no masks zero and no normaliser, returns[t] = sum_k gamma^k r[t+k] + gamma^(T-t) V_T."""
    from mappo_b200 import ops
    from oracle import buffer as ob
    T, R = 25, 65536 * 3
    g = torch.Generator().manual_seed(0)
    rew = (torch.randn(T, R, 1, generator=g) * 2 - 5).cuda()
    vp = torch.randn(T + 1, R, 1, generator=g).cuda()
    masks = torch.ones(T + 1, R, 1).cuda()
    masks[3, ::7] = 0
    nv = torch.randn(R, 1, generator=g).cuda()
    vn = torch.tensor([-0.3, 2.5, 0.7]).cuda()
    ret = torch.zeros_like(vp)
    raw, s3 = ops.gae_returns(rew, vp, ret, masks, nv, 0.99, 0.95, vn_stats=vn)
    o = ob.OracleValueNorm(); o.load(vn.cpu().numpy())
    sl = slice(0, 4096)
    vpo, reto = ob.compute_returns(rew[:, sl].cpu().numpy(), vp[:, sl].cpu().numpy() * 0 + vp[:, sl].cpu().numpy(),
                                   masks[:, sl].cpu().numpy(), nv[sl].cpu().numpy(), 0.99, 0.95, normalizer=o)
    np.testing.assert_array_equal(ret[:-1, sl].cpu().numpy(), reto[:-1])
    adv = ops.adv_normalize(raw, s3)
    a = adv.double()
    assert abs(a.mean().item()) < 1e-4 and abs(a.std(unbiased=False).item() - 1.0) < 1e-3
    # discounted-sum property
    masks.fill_(1.0)
    ret2 = torch.zeros_like(vp)
    ops.gae_returns(rew, vp, ret2, masks, nv, 0.99, 1.0)
    want = nv.double().clone()
    for t in reversed(range(T)):
        want = rew[t].double() + 0.99 * want
    np.testing.assert_allclose(ret2[0].double().cpu().numpy(), want.cpu().numpy(), rtol=2e-5, atol=2e-5)
