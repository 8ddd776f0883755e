"""GPU: the device-resident MPERunner + R_MAPPO.train replayed on the fixtures recorded from the reference's OWN
MPERunner.run() over two consecutive iterations (tests/golden/runner_*.npz; tests/runner_common.py).  The recorded
actions are evaluated instead of sampled and the recorded reset states are installed at the auto-resets, so every
buffer field, the returns, the normalised advantages and the six train_info scalars can be compared with what the
reference itself produced -- including the second iteration (masks[0] = 0 / zeroed rnn states after after_update, GAE
and advantages under the ValueNorm state left by the first train())."""
import numpy as np
import pytest
import torch

import runner_common as rc

pytestmark = pytest.mark.gpu

# north_star: env state/obs/reward 1e-5 relative, returns/advantages 1e-6 given identical weights (iteration 0: the
# values themselves come from a tensor-core forward within 1e-5 of torch's, so returns follow at 1e-5), losses and
# gradients 1e-4; iteration 1 starts from weights two Adam steps away from the reference's (1e-5 relative apart).
TOL0 = dict(obs=1e-6, logp=2e-5, value=2e-5, ret=5e-5, adv=2e-4)
TOL1 = dict(obs=1e-6, logp=2e-4, value=2e-4, ret=5e-4, adv=2e-3)


def _runner(fx, extra=""):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.shared.mpe_runner import MPERunner
    a = parse_args(rc.flags(fx, extra))
    envs = MPEVecEnv(a, device="cuda:0")
    r = MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda:0"), "num_agents": int(fx["M"]), "run_dir": None})
    r.policy.actor.load_state_dict(rc.sd_from(fx, "actor0."))
    r.policy.critic.load_state_dict(rc.sd_from(fx, "critic0."))
    return r


def _buffers(b):
    n = lambda t: t.detach().cpu().numpy()
    return dict(obs=n(b.obs), rewards=n(b.rewards), masks=n(b.masks), actions=n(b.actions),
                action_log_probs=n(b.action_log_probs), value_preds=n(b.value_preds),
                rnn_states=n(b.rnn_states) if b._recurrent else None,
                rnn_states_critic=n(b.rnn_states_critic) if b._recurrent else None)


@pytest.mark.parametrize("tag", rc.CASES)
def test_two_iterations_of_the_reference_runner(tag, golden_dir):
    from mappo_b200 import ops
    fx = rc.load(golden_dir, tag)
    r = _runner(fx)
    b, tr = r.buffer, r.trainer
    ws = rc.warm_state(fx)
    r.envs.set_state(p_pos=ws["p_pos"], p_vel=np.zeros_like(ws["p_pos"]), lm_pos=ws["lm_pos"], goal=ws["goal"])
    r.envs.reset(out_obs=b.obs[0], use_given_state=True)
    np.testing.assert_allclose(b.obs[0].cpu().numpy(), fx["warm_obs"].astype(np.float32), rtol=1e-6, atol=1e-7)
    for k in range(int(fx["iters"])):
        tol = TOL0 if k == 0 else TOL1
        ga = torch.as_tensor(rc.action_ids(fx, k), dtype=torch.float32, device="cuda:0")
        r._rollout(given_actions=ga, reset_states=rc.reset_states(fx, k))
        rc.check_rollout(fx, k, _buffers(b), tol)
        vn = tr.value_normalizer
        if vn is not None:
            np.testing.assert_allclose(vn.stats.cpu().numpy(), fx["it%d_vn_before" % k], rtol=1e-5)
        r.compute()
        raw, s3 = ops.advantages(b.returns[:-1], b.value_preds[:-1], b.active_masks[:-1], None if vn is None else vn.stats)
        adv = ops.adv_normalize(raw, s3)
        rc.check_compute(fx, k, b.returns.cpu().numpy(), b.value_preds.cpu().numpy(), adv.cpu().numpy(), tol)
        torch.manual_seed(100 + k)
        info = r.train()
        rc.check_info(fx, k, info, rtol=3e-4 if k == 0 else 3e-3)
        if vn is not None:
            np.testing.assert_allclose(vn.stats.cpu().numpy(), fx["it%d_vn_after" % k], rtol=1e-5)
        for name, net in (("actor", r.policy.actor), ("critic", r.policy.critic)):
            for key, p in net.state_dict().items():
                if "fc_h" in key:
                    continue
                np.testing.assert_allclose(p.cpu().numpy(), fx["it%d_%s.%s" % (k, name, key)], rtol=1e-3, atol=2e-5,
                                           err_msg="it%d %s %s" % (k, name, key))
        # after_update (shared_buffer.py:149-159): slot T -> slot 0
        assert (b.masks[0].cpu().numpy() == fx["it%d_after_masks0" % k]).all()
        np.testing.assert_allclose(b.obs[0].cpu().numpy(), fx["it%d_after_obs0" % k], rtol=1e-6, atol=1e-7)
        if b._recurrent:
            np.testing.assert_allclose(b.rnn_states[0].cpu().numpy(), fx["it%d_after_rnn_states0" % k], atol=1e-6)
