"""GPU: the CUDA env (through the C ABI) against (a) fixtures recorded from the unmodified reference and
(b) the oracle at larger sizes, including Philox auto-resets bit-for-bit."""
import os
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TAGS = ["spread_m3", "spread_m10", "speaker_listener", "reference", "adversary", "push", "tag", "crypto", "world_comm"]


def make_args(scenario, M, L, T=25, N=4, seed=1, **kw):
    return types.SimpleNamespace(scenario_name=scenario, num_agents=M, num_landmarks=L, episode_length=T,
                                 n_rollout_threads=N, seed=seed, **kw)


def f32(x):
    return np.asarray(x, dtype=np.float64).astype(np.float32)


@pytest.mark.parametrize("tag", TAGS)
def test_env_matches_reference_fixture(tag, golden_dir):
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    fx = np.load(os.path.join(golden_dir, "env_%s.npz" % tag))
    N, M, L = int(fx["N"]), int(fx["M"]), int(fx["L"])
    kw = dict(num_adversaries=M - 1, num_good_agents=1) if str(fx["scenario"]) == "simple_tag" else {}
    if str(fx["scenario"]) == "simple_world_comm":
        kw = dict(num_adversaries=M - 2, num_good_agents=2)
    env = MPEVecEnv(make_args(str(fx["scenario"]), M, L, int(fx["episode_length"]), N, **kw), device="cuda")
    env.set_state(p_pos=fx["init_p_pos"], p_vel=np.zeros_like(fx["init_p_pos"]), lm_pos=fx["init_lm_pos"],
                  goal=fx["init_goal"])
    obs = env.reset(use_given_state=True)
    obs = [obs[:, i] for i in range(M)] if env.homogeneous else obs
    for i in range(M):
        np.testing.assert_array_equal(obs[i].cpu().numpy(), f32(fx["init_obs_%d" % i]))
    resets = {int(s): k for k, s in enumerate(fx["reset_steps"])}
    worst = 0.0
    for s in range(int(fx["steps"])):
        rs = None
        if s in resets:
            k = resets[s]
            rs = dict(p_pos=fx["reset_p_pos"][k], lm_pos=fx["reset_lm_pos"][k], goal=fx["reset_goal"][k])
        ids = torch.as_tensor(fx["actions"][s], dtype=torch.float32, device="cuda").contiguous()
        o, r, m = env.step_ids(ids, reset_state=rs)
        o = o.view(N, env.obs_width).cpu().numpy()
        off = 0
        for i in range(M):
            d = env.obs_dims[i]
            want = f32(fx["obs_%d" % i][s])
            # north_star tolerance: 1e-5 relative per step; the f64 kernel is far inside it
            np.testing.assert_allclose(o[:, off:off + d], want, rtol=1e-6, atol=1e-7,
                                       err_msg="%s obs step %d agent %d" % (tag, s, i))
            worst = max(worst, float(np.abs(o[:, off:off + d] - want).max()))
            off += d
        np.testing.assert_allclose(r.view(N, M).cpu().numpy(), f32(fx["rewards"][s][..., 0]), rtol=1e-6)
        np.testing.assert_allclose(env._indiv.cpu().numpy(), f32(fx["indiv"][s]), rtol=1e-6)
        assert (env._dones.cpu().numpy().astype(bool) == fx["dones"][s]).all()
        assert (m.view(N, M).cpu().numpy() == 1.0 - fx["dones"][s]).all()
        if s not in resets:
            st = env.get_state()
            np.testing.assert_allclose(st["p_pos"], fx["post_p_pos"][s], rtol=1e-12, atol=1e-14)
            np.testing.assert_allclose(st["p_vel"], fx["post_p_vel"][s], rtol=1e-12, atol=1e-14)
    print(tag, "worst abs obs diff", worst)


@pytest.mark.parametrize("scenario,M,L,nadv", [("simple_spread", 3, 3, None), ("simple_spread", 10, 3, None),
                                               ("simple_speaker_listener", 2, 3, None), ("simple_reference", 2, 3, None),
                                               ("simple_adversary", 3, 2, None), ("simple_adversary", 5, 4, None),
                                               ("simple_push", 2, 2, None), ("simple_push", 4, 2, None),
                                               ("simple_tag", 4, 2, 3), ("simple_tag", 6, 3, 2),
                                               ("simple_crypto", 3, 2, None), ("simple_crypto_display", 3, 4, None),
                                               ("simple_world_comm", 4, 1, 2), ("simple_world_comm", 6, 2, 4)])
def test_env_matches_oracle_with_philox_resets(scenario, M, L, nadv):
    """Free-running 60 steps (two auto-resets) at N=1000 with an env-id offset: states equal the oracle's."""
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from oracle.mpe import OracleMPE
    N, T, off = 1000, 25, 12345
    kw = dict(num_adversaries=nadv) if nadv else {}
    env = MPEVecEnv(make_args(scenario, M, L, T, N, seed=7, **kw), device="cuda", env_id_offset=off)
    orc = OracleMPE(scenario, N, M, L, T, seed=7, env_id_offset=off, **kw)
    o_dev = env.reset()
    o_ref = orc.reset()
    st = env.get_state()
    assert (st["p_pos"] == orc.p_pos).all() and (st["lm_pos"] == orc.lm_pos).all()      # Philox: bit-exact
    assert (st["goal"] == orc.goal).all()
    rng = np.random.RandomState(0)
    heads = env.head_sizes
    for s in range(60):
        ids = np.zeros((N, M, env.act_width), dtype=np.int64)
        for i in range(M):
            for h, k in enumerate(heads[i]):
                ids[:, i, h] = rng.randint(0, k, N)
        o, r, m = env.step_ids(torch.as_tensor(ids, dtype=torch.float32, device="cuda"))
        obs, rew, done, indiv = orc.step(ids)
        want = np.concatenate(obs, axis=-1).astype(np.float32)
        np.testing.assert_allclose(o.view(N, -1).cpu().numpy(), want, rtol=1e-6, atol=1e-7, err_msg="step %d" % s)
        np.testing.assert_allclose(r.view(N, M).cpu().numpy(), rew[..., 0].astype(np.float32), rtol=1e-6, atol=1e-7)
        st = env.get_state()
        np.testing.assert_allclose(st["p_pos"], orc.p_pos, rtol=1e-11, atol=1e-13)
        assert (st["step_count"] == orc.step_count).all()
    assert (env.episode.cpu().numpy() == orc.episode).all() and orc.episode[0] == 3


def test_env_rejects_bad_config():
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    with pytest.raises(NotImplementedError):
        MPEVecEnv(make_args("simple_attack", 4, 4))     # broken in the reference itself (NameError: bound), not built
    with pytest.raises(AssertionError):
        MPEVecEnv(make_args("simple_reference", 3, 3))
