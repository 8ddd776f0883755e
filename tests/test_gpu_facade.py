"""GPU: (1) the single-env MultiAgentEnv / World facade (envs/mpe/environment.py; reference environment.py:17-163,
core.py:114-140) replayed on the env fixtures recorded from the reference's own single envs; (2) Runner.save ->
restore round trip and loading a checkpoint WRITTEN BY THE REFERENCE (tests/golden/ref_ckpt, base_runner.py:127-146)."""
import os
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(scenario, M, L, **kw):
    return types.SimpleNamespace(scenario_name=scenario, num_agents=M, num_landmarks=L, episode_length=25,
                                 n_rollout_threads=1, seed=1, **kw)


def _one_hot(sizes, ids):
    return np.concatenate([np.eye(int(s))[int(i)] for s, i in zip(sizes, ids)])


@pytest.mark.parametrize("tag", ["spread_m3", "reference", "speaker_listener"])
def test_single_env_facade_replays_the_reference_env(tag, golden_dir):
    """One reference env of the fixture, driven through MPEEnv(args).step(action_n) with one-hot actions and the
    per-agent-list returns; state injected / read through env.world.agents[i].state."""
    from mappo_b200.envs.mpe.MPE_env import MPEEnv
    fx = np.load(os.path.join(golden_dir, "env_%s.npz" % tag))
    M, L, n = int(fx["M"]), int(fx["L"]), 1
    env = MPEEnv(_args(str(fx["scenario"]), M, L))
    assert env.n == M and len(env.action_space) == M and env.world_length == 25
    env.reset()
    for i, ag in enumerate(env.world.agents):                    # install the fixture's initial state through the view
        ag.state.p_pos = fx["init_p_pos"][n, i]; ag.state.p_vel = np.zeros(2)
    for j, lm in enumerate(env.world.landmarks):
        lm.state.p_pos = fx["init_lm_pos"][n, j]
    env._vec.goal[:, 0] = torch.as_tensor(fx["init_goal"][n]).to(env._vec.goal)
    np.testing.assert_array_equal(env.world.agents[0].state.p_pos, fx["init_p_pos"][n, 0])
    sizes = env._vec.head_sizes
    first_reset = int(fx["reset_steps"][0])
    for s in range(first_reset + 3):                             # past the episode end: the bare env does NOT reset itself
        action_n = [_one_hot(sizes[i], fx["actions"][s, n, i, :len(sizes[i])]) for i in range(M)]
        obs_n, reward_n, done_n, info_n = env.step(action_n)
        assert isinstance(obs_n, list) and len(obs_n) == M and obs_n[0].dtype == np.float64
        assert done_n == [s + 1 >= 25] * M and env.current_step == s + 1
        if s < first_reset:                                      # same trajectory as the reference env up to its reset
            for i in range(M):
                np.testing.assert_allclose(obs_n[i], fx["obs_%d" % i][s, n].astype(np.float32), rtol=1e-6, atol=1e-7)
                np.testing.assert_allclose(reward_n[i][0], np.float32(fx["rewards"][s, n, i, 0]), rtol=1e-6)
                np.testing.assert_allclose(info_n[i]["individual_reward"], np.float32(fx["indiv"][s, n, i]), rtol=1e-6)
            np.testing.assert_allclose(env.world.agents[M - 1].state.p_pos, fx["post_p_pos"][s, n, M - 1], rtol=1e-12, atol=1e-14)
            np.testing.assert_allclose(env.world.agents[0].state.p_vel, fx["post_p_vel"][s, n, 0], rtol=1e-12, atol=1e-14)
    obs_n = env.reset()
    assert env.current_step == 0 and len(obs_n) == M


def test_view_of_one_env_of_a_batch():
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.envs.mpe.environment import MultiAgentEnv
    a = _args("simple_spread", 3, 3)
    a.n_rollout_threads = 8
    vec = MPEVecEnv(a, device="cuda")
    vec.reset()
    view = MultiAgentEnv.view(vec, 5)
    before = vec.get_state()["p_pos"].copy()
    obs_n, reward_n, done_n, info_n = view.step([np.eye(5)[1]] * 3)         # +x for every agent of env 5
    after = vec.get_state()
    assert (after["p_vel"][5, :, 0] > 0).all() and np.abs(after["p_vel"][[0, 1, 2, 3, 4, 6, 7]]).max() < 1.0
    np.testing.assert_allclose(view.world.agents[2].state.p_pos, after["p_pos"][5, 2])
    assert np.abs(after["p_pos"][5] - before[5]).max() > 0


def _runner(tmp, extra=""):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.shared.mpe_runner import MPERunner
    a = parse_args(("--scenario_name simple_spread --num_agents 3 --num_landmarks 3 --n_rollout_threads 64 "
                    "--episode_length 25 --ppo_epoch 2 --use_ReLU --seed 3 " + extra).split())
    envs = MPEVecEnv(a, device="cuda:0")
    return MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda:0"), "num_agents": 3, "run_dir": tmp})


def test_save_restore_round_trip(tmp_path):
    torch.manual_seed(5)
    r = _runner(str(tmp_path))
    r.warmup(); r.run_episode(); r.run_episode()
    r.save()
    files = sorted(os.listdir(os.path.join(str(tmp_path), "models")))
    assert files == ["actor.pt", "critic.pt", "vnorm.pt"]
    torch.manual_seed(99)                                        # different initial weights
    r2 = _runner(None, "--model_dir %s" % os.path.join(str(tmp_path), "models"))
    for a, b in ((r.policy.actor, r2.policy.actor), (r.policy.critic, r2.policy.critic)):
        assert torch.equal(a.flat, b.flat)
        for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
            assert ka == kb and torch.equal(va, vb)
    assert torch.equal(r.trainer.value_normalizer.stats, r2.trainer.value_normalizer.stats)
    # the restored policy acts identically
    obs = torch.randn(50, 18, device="cuda")
    a1, l1, _ = r.policy.actor(obs, None, None, None, deterministic=True)
    a2, l2, _ = r2.policy.actor(obs, None, None, None, deterministic=True)
    assert torch.equal(a1, a2) and torch.equal(l1, l2)


def test_restore_a_checkpoint_written_by_the_reference(golden_dir):
    d = os.path.join(golden_dir, "ref_ckpt")
    io = np.load(os.path.join(d, "io.npz"))
    r = _runner(None, "--model_dir %s" % d)
    values, actions, logp, _, _ = r.policy.get_actions(io["share_obs"], io["obs"], None, None, None, None, deterministic=True)
    assert (actions.cpu().numpy() == io["actions"]).all()
    np.testing.assert_allclose(logp.cpu().numpy(), io["log_probs"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(values.cpu().numpy(), io["values"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(r.trainer.value_normalizer.stats.cpu().numpy(), io["vn"], rtol=0, atol=0)
    np.testing.assert_allclose(r.trainer.value_normalizer.denormalize(io["values"]).cpu().numpy(), io["denorm"], rtol=1e-6)
    # and the file it writes back can be read by torch.load with the reference's key set
    want = torch.load(os.path.join(d, "actor.pt"), map_location="cpu")
    assert list(r.policy.actor.state_dict().keys()) == list(want.keys())
