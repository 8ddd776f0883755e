"""GPU: --use_popart against fixtures recorded from the UNMODIFIED reference running on cuda on the B200 box
(tests/golden/ppo_popart*_cuda.npz, made there by tests/golden/make_golden.py::gen_popart from baseline/_ref).

What the reference does on cuda (mappo_b200/algorithms/utils/popart.py has the analysis): PopArt's weight / bias /
statistics are plain tensors -- not in critic.state_dict(), not optimised, not part of the clipped gradient norm -- and
the value head only moves through the rescale at the end of PopArt.update; the value loss uses values formed with the
head BEFORE that rescale and targets normalised with the statistics AFTER the EMA step."""
import os

import numpy as np
import pytest
import torch

from ppo_common import sd_from
from test_gpu_policy import build, sample_of

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["popart_cuda", "popart_mse_cuda"])
def test_popart_update_matches_the_reference_on_cuda(tag, golden_dir):
    fx = np.load(os.path.join(golden_dir, "ppo_%s.npz" % tag))
    args, pol, tr = build(fx)
    assert args.use_popart and tr.value_normalizer is pol.critic.v_out
    # the reference's critic.state_dict() on cuda has no v_out entry at all, and its optimiser never sees the head
    assert sorted(pol.critic.state_dict().keys()) == list(fx["critic_state_dict_keys"])
    assert not any(k.startswith("v_out") for k in dict(pol.critic.named_parameters()))
    assert pol.critic_optimizer.frozen_head and not pol.actor_optimizer.frozen_head
    v = pol.critic.v_out
    v.weight.copy_(torch.as_tensor(fx["popart0.weight"])); v.bias.copy_(torch.as_tensor(fx["popart0.bias"]))
    s = sample_of(fx)
    values = pol.get_values(s[0], s[3], s[7])
    np.testing.assert_allclose(values.cpu().numpy(), fx["values0"], rtol=1e-4, atol=2e-5)
    Pa = pol.actor.flat.numel()
    n_updates = len([k for k in fx.files if k.startswith("u") and k.endswith("_value_loss")])
    for u in range(n_updates):
        vl, cgn, pl, ent, agn, ratio = tr.ppo_update(s)
        t = "u%d_" % u
        for k, got in dict(value_loss=vl, critic_grad_norm=cgn, policy_loss=pl, dist_entropy=ent, actor_grad_norm=agn,
                           ratio_mean=ratio).items():
            np.testing.assert_allclose(got.item(), fx[t + k].item(), rtol=1e-4, atol=1e-6, err_msg="%s (update %d)" % (k, u))
        ws = next(iter(tr._ws.values()))
        grads = pol.critic.unflatten(ws["grad"][Pa:])
        for k, gv in grads.items():
            if k.startswith("v_out."):
                assert (gv == 0).all()                       # outside the optimiser: no gradient reported, no step
                continue
            want = fx[t + "critic_grad." + k]
            np.testing.assert_allclose(gv.cpu().numpy(), want, rtol=1e-4, atol=1e-4 * np.abs(want).max() + 1e-9,
                                       err_msg="critic grad %s (update %d)" % (k, u))
        for k, p in pol.critic.state_dict().items():
            if "fc_h" not in k:
                np.testing.assert_allclose(p.cpu().numpy(), fx[t + "critic." + k], rtol=1e-5, atol=3e-6, err_msg=k)
        # PopArt state: statistics, rescaled head, stddev
        np.testing.assert_allclose(v.stats.cpu().numpy(),
                                   [fx[t + "popart.mean"][0], fx[t + "popart.mean_sq"][0], fx[t + "popart.debiasing_term"]],
                                   rtol=1e-6)
        np.testing.assert_allclose(v.weight.cpu().numpy(), fx[t + "popart.weight"], rtol=2e-6, atol=1e-9)
        np.testing.assert_allclose(v.bias.cpu().numpy(), fx[t + "popart.bias"], rtol=2e-6, atol=1e-9)
        np.testing.assert_allclose(v.stddev.cpu().numpy(), fx[t + "popart.stddev"], rtol=1e-4)
        # what the next rollout / GAE will see
        np.testing.assert_allclose(pol.get_values(s[0], s[3], s[7]).cpu().numpy(), fx[t + "values_after"], rtol=1e-4, atol=2e-5)
        np.testing.assert_allclose(v.denormalize(s[5]).cpu().numpy(), fx[t + "denorm_value_preds"], rtol=1e-5,
                                   atol=1e-6 * float(np.abs(fx[t + "denorm_value_preds"]).max()))   # x * std + mean cancels
