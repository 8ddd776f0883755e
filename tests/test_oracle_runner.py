"""CPU: oracle/iteration.py (the runner composition: rollout, insert, compute, train, after_update) replayed on the
fixtures recorded from the reference's own MPERunner.run() over TWO consecutive iterations -- the cross-iteration
effects (masks[0] = 0 and zeroed rnn states from the second rollout on; GAE and the advantages using the ValueNorm state
the previous train() left behind) and train()'s six averaged scalars are pinned here, not only single calls."""
import numpy as np
import pytest
import torch

import runner_common as rc
from mappo_b200.config import parse_args
from oracle import buffer as ob
from oracle.iteration import OracleMAPPO

TOL0 = dict(obs=1e-6, logp=2e-6, value=2e-6, ret=1e-5, adv=2e-5)       # first iteration: same weights, same torch ops
TOL1 = dict(obs=1e-6, logp=1e-4, value=1e-4, ret=2e-4, adv=1e-3)       # second: weights after two Adam steps each


@pytest.mark.parametrize("tag", rc.CASES)
def test_oracle_iteration_replays_the_reference_runner(tag, golden_dir):
    fx = rc.load(golden_dir, tag)
    a = parse_args(rc.flags(fx))
    o = OracleMAPPO(a, rc.sd_from(fx, "actor0."), rc.sd_from(fx, "critic0."))
    o.warmup(rc.warm_state(fx))
    np.testing.assert_allclose(o.obs[0], fx["warm_obs"].astype(np.float32), rtol=1e-6, atol=1e-7)
    for k in range(int(fx["iters"])):
        tol = TOL0 if k == 0 else TOL1
        o.rollout(given_actions=rc.action_ids(fx, k), reset_states=rc.reset_states(fx, k))
        got = dict(obs=o.obs, rewards=o.rewards, masks=o.masks, actions=o.actions, action_log_probs=o.action_log_probs,
                   value_preds=o.value_preds, rnn_states=o.rnn_states, rnn_states_critic=o.rnn_states_critic)
        rc.check_rollout(fx, k, got, tol)
        if o.vn is not None:
            np.testing.assert_allclose(o.vn.stats(), fx["it%d_vn_before" % k], rtol=1e-5)
        o.compute()
        _, adv = ob.advantages(o.returns, o.value_preds, o.active_masks, o.vn)
        rc.check_compute(fx, k, o.returns, o.value_preds, adv, tol)
        torch.manual_seed(100 + k)                         # the recurrent generator's randperm (order of a mean only)
        info = o.train()
        rc.check_info(fx, k, info, rtol=2e-4 if k == 0 else 2e-3)
        if o.vn is not None:
            np.testing.assert_allclose(o.vn.stats(), fx["it%d_vn_after" % k], rtol=1e-5)
        for key in o.akeys:
            np.testing.assert_allclose(o.actor[key].numpy(), fx["it%d_actor.%s" % (k, key)], rtol=1e-3, atol=2e-5, err_msg=key)
        for key in o.ckeys:
            np.testing.assert_allclose(o.critic[key].numpy(), fx["it%d_critic.%s" % (k, key)], rtol=1e-3, atol=2e-5, err_msg=key)
        # after_update (shared_buffer.py:149-159)
        assert (o.masks[0] == fx["it%d_after_masks0" % k]).all() and (o.masks[0] == 0).all()
        np.testing.assert_allclose(o.obs[0], fx["it%d_after_obs0" % k], rtol=1e-6, atol=1e-7)
