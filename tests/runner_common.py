"""Shared by tests/test_oracle_runner.py (CPU) and tests/test_gpu_runner.py (GPU): the fixtures recorded from the
reference's OWN MPERunner.run() + R_MAPPO.train() over two consecutive iterations (tests/golden/runner_*.npz, made by
tests/golden/make_golden.py::make_runner_fixture) and what to compare after each stage.

Reference lines pinned here: runner/shared/mpe_runner.py:16-139 (run / warmup / collect / insert), shared/base_runner.py:
110-125 (compute / train), r_mappo.py:166-219 (train: advantage normalisation, epochs, the six averaged scalars),
shared_buffer.py:149-159 (after_update: slot T -> slot 0, so masks[0] = 0 and rnn_states[0] = 0 from the second
rollout on)."""
import os

import numpy as np
import torch

CASES = ["spread_mappo", "reference_rmappo"]
INFO_KEYS = ("value_loss", "policy_loss", "dist_entropy", "actor_grad_norm", "critic_grad_norm", "ratio")


def load(golden_dir, tag):
    return np.load(os.path.join(golden_dir, "runner_%s.npz" % tag))


def flags(fx, extra=""):
    """The command line that rebuilds the recorded run's namespace for mappo_b200.config.parse_args."""
    a = {k[5:]: fx[k].item() for k in fx.files if k.startswith("args_")}
    f = ["--env_name", "MPE", "--algorithm_name", str(fx["algo"]), "--scenario_name", str(fx["scenario"]),
         "--num_agents", str(int(fx["M"])), "--num_landmarks", str(int(fx["L"])), "--n_rollout_threads", str(int(fx["N"])),
         "--episode_length", str(int(fx["T"])), "--ppo_epoch", str(int(a["ppo_epoch"])), "--num_mini_batch",
         str(int(a["num_mini_batch"])), "--lr", repr(float(a["lr"])), "--critic_lr", repr(float(a["critic_lr"])),
         "--gain", repr(float(a["gain"])), "--data_chunk_length", str(int(fx["chunk"])), "--seed", "1"]
    if not a["use_ReLU"]:
        f.append("--use_ReLU")                               # store_false in the reference's parser (config.py:203)
    return f + extra.split()


def sd_from(fx, prefix):
    return {k[len(prefix):]: torch.tensor(fx[k]) for k in fx.files if k.startswith(prefix)}


def action_ids(fx, k):
    """Sampled action ids of iteration k, [T, N, M, heads] (what the reference's buffer stored)."""
    return fx["it%d_buf_actions" % k].astype(np.int64)


def reset_states(fx, k):
    """{step: state} for the steps of iteration k at which the envs auto-reset (dones all True)."""
    out = {}
    dones = fx["it%d_dones" % k]
    for s in range(dones.shape[0]):
        if dones[s].all():
            out[s] = dict(p_pos=fx["it%d_p_pos" % k][s], lm_pos=fx["it%d_lm_pos" % k][s], goal=fx["it%d_goal" % k][s])
        else:
            assert not dones[s].any()
    return out


def warm_state(fx):
    return dict(p_pos=fx["warm_p_pos"], lm_pos=fx["warm_lm_pos"], goal=fx["warm_goal"])


def check_rollout(fx, k, got, tol):
    """got: dict of numpy arrays named like the reference buffer's fields, after the rollout of iteration k."""
    T = int(fx["T"])
    np.testing.assert_allclose(got["obs"], fx["it%d_buf_obs" % k], rtol=tol["obs"], atol=tol["obs"], err_msg="it%d obs" % k)
    np.testing.assert_allclose(got["rewards"], fx["it%d_buf_rewards" % k], rtol=tol["obs"], atol=0, err_msg="it%d rewards" % k)
    assert (got["masks"] == fx["it%d_buf_masks" % k]).all(), "it%d masks" % k
    # masks[0] is 1 only in the very first rollout
    assert (got["masks"][0] == (1.0 if k == 0 else 0.0)).all()
    assert (got["masks"][T] == 0.0).all()
    assert (got["actions"] == fx["it%d_buf_actions" % k]).all()
    np.testing.assert_allclose(got["action_log_probs"], fx["it%d_buf_action_log_probs" % k], rtol=tol["logp"],
                               atol=tol["logp"], err_msg="it%d action_log_probs" % k)
    np.testing.assert_allclose(got["value_preds"][:T], fx["it%d_value_preds_rollout" % k][:T], rtol=tol["value"],
                               atol=tol["value"], err_msg="it%d value_preds" % k)
    if int(fx["recurrent"]):
        np.testing.assert_allclose(got["rnn_states"], fx["it%d_buf_rnn_states" % k], rtol=tol["value"], atol=tol["value"],
                                   err_msg="it%d rnn_states" % k)
        np.testing.assert_allclose(got["rnn_states_critic"], fx["it%d_buf_rnn_states_critic" % k], rtol=tol["value"],
                                   atol=tol["value"], err_msg="it%d rnn_states_critic" % k)
        if k > 0:
            assert (got["rnn_states"][0] == 0.0).all() and (got["rnn_states_critic"][0] == 0.0).all()


def check_compute(fx, k, returns, value_preds, adv, tol):
    np.testing.assert_allclose(value_preds, fx["it%d_value_preds" % k], rtol=tol["value"], atol=tol["value"],
                               err_msg="it%d value_preds after compute" % k)
    np.testing.assert_allclose(returns, fx["it%d_returns" % k], rtol=tol["ret"], atol=tol["ret"], err_msg="it%d returns" % k)
    if adv is not None:
        np.testing.assert_allclose(adv, fx["it%d_adv" % k], rtol=tol["adv"], atol=tol["adv"], err_msg="it%d advantages" % k)


def check_info(fx, k, info, rtol, atol=2e-6):
    for name in INFO_KEYS:
        np.testing.assert_allclose(info[name], float(fx["it%d_info_%s" % (k, name)]), rtol=rtol, atol=atol,
                                   err_msg="it%d train_info[%s]" % (k, name))
