"""CPU: the oracle's policy / PPO-update restatement against fixtures recorded from the reference's own
R_MAPPO.ppo_update (tests/golden/ppo_*.npz)."""
import os

import numpy as np
import pytest
import torch

from ppo_common import oracle_update, sd_from

CASES = ["mlp_tanh", "mlp_relu_opts", "mlp_listener", "gru_reference", "mlp_layer2", "mlp_layer3_relu", "gru2_reference"]


@pytest.mark.parametrize("tag", CASES)
def test_oracle_ppo_update_matches_reference(tag, golden_dir):
    fx = np.load(os.path.join(golden_dir, "ppo_%s.npz" % tag))
    actor_sd, critic_sd = sd_from(fx, "actor0."), sd_from(fx, "critic0.")
    vn = fx["vn0"] if "vn0" in fx.files else None
    state = {"actor": {}, "critic": {}}
    for u in range(2):
        r = oracle_update(fx, actor_sd, critic_sd, vn, state, u + 1)
        t = "u%d_" % u
        for k in ("policy_loss", "dist_entropy", "value_loss", "ratio_mean", "actor_grad_norm", "critic_grad_norm"):
            np.testing.assert_allclose(r[k], fx[t + k].item(), rtol=2e-5, atol=1e-7, err_msg=k)
        for net in ("actor", "critic"):
            for k, g in r[net + "_grads"].items():
                want = fx[t + net + "_grad." + k]
                np.testing.assert_allclose(g.numpy(), want, rtol=1e-4, atol=1e-6 * max(1.0, np.abs(want).max()),
                                           err_msg="%s grad %s" % (net, k))
            for k, p in r[net + "_sd"].items():
                if "fc_h" in k:
                    continue
                np.testing.assert_allclose(p.numpy(), fx[t + net + "." + k], rtol=1e-5, atol=2e-6, err_msg=k)
        if vn is not None:
            np.testing.assert_allclose(r["vn"], fx[t + "vn"], rtol=1e-6)
            vn = r["vn"]
        actor_sd, critic_sd = r["actor_sd"], r["critic_sd"]
        state = {"actor": r["actor_state"], "critic": r["critic_state"]}
