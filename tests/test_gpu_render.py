"""GPU: the rasteriser kernel (mappo_env_render, csrc/render.cu) through MPEVecEnv.render against oracle/render.py,
bit for bit, for every scenario's colour / size / communication-circle rules."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [("simple_spread", 3, 3, {}), ("simple_spread", 10, 3, {}), ("simple_speaker_listener", 2, 3, {}),
         ("simple_reference", 2, 3, {}), ("simple_adversary", 3, 2, {}), ("simple_push", 2, 2, {}),
         ("simple_tag", 4, 2, dict(num_adversaries=3, num_good_agents=1)), ("simple_crypto", 3, 2, {}),
         ("simple_world_comm", 4, 1, dict(num_adversaries=2, num_good_agents=2))]


def make_env(scenario, M, L, N, **kw):
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    a = types.SimpleNamespace(scenario_name=scenario, num_agents=M, num_landmarks=L, episode_length=25,
                              n_rollout_threads=N, seed=7, **kw)
    return MPEVecEnv(a, device="cuda")


def oracle_frames(env, env_ids, w, h):
    from mappo_b200.envs.mpe import rendering
    from oracle import render as R
    st = env.get_state()
    size, cdim = rendering.entity_sizes(env), rendering.comm_dims(env)
    rgba = rendering.entity_rgba(env, env_ids).cpu().numpy()
    out = []
    for k, n in enumerate(env_ids):
        pos = np.concatenate([st["p_pos"][n], st["lm_pos"][n]], 0)
        comm = np.concatenate([st["comm"][n], -np.ones(env.num_landmarks_total, np.int64)])
        out.append(R.render_env(pos, size, rgba[k], cdim, comm, width=w, height=h))
    return np.stack(out)


@pytest.mark.parametrize("scenario,M,L,kw", CASES)
def test_render_matches_oracle(scenario, M, L, kw):
    N = 6
    env = make_env(scenario, M, L, N, **kw)
    env.reset()
    rng = np.random.RandomState(3)
    for _ in range(4):                                                  # move, and let the speakers say something
        ids = np.stack([rng.randint(0, h[0], (N,)) for h in env.head_sizes], 1)[..., None].astype(np.float32)
        ids = np.repeat(ids, env.act_width, -1)
        if env.act_width == 2:
            ids[..., 1] = rng.randint(0, min(h[1] for h in env.head_sizes if len(h) > 1), (N, env.num_agents))
        env.step_ids(torch.as_tensor(ids, device="cuda").contiguous())
    sel = [4, 0, 5]
    w, h = 232, 200                                                     # ragged against the 128 x 8 tile
    got = env.render("rgb_array", env_ids=sel, width=w, height=h)
    assert got.shape == (3, 1, h, w, 3) and got.dtype == np.uint8
    want = oracle_frames(env, sel, w, h)
    assert (got[:, 0] != 255).any(), "nothing was drawn"
    np.testing.assert_array_equal(got[:, 0], want)


def test_render_full_size_all_envs_and_properties():
    env = make_env("simple_spread", 3, 3, 4)
    env.reset()
    frames = env.render("rgb_array", as_numpy=False)
    assert tuple(frames.shape) == (4, 1, 700, 700, 3)
    f = frames[:, 0].cpu().numpy()
    np.testing.assert_array_equal(f[2], oracle_frames(env, [2], 700, 700)[0])
    # each agent centre that no other entity overlaps shows the agent colour blended once over white
    st = env.get_state()
    px = lambda p: (int((p[0] + 2.0) * 175.0), 699 - int((p[1] + 2.0) * 175.0))
    for n in range(4):
        ents = np.concatenate([st["p_pos"][n], st["lm_pos"][n]], 0)
        for i in range(3):
            d = np.linalg.norm(ents - ents[i], axis=1)
            d[i] = 9.0
            if d.min() > 0.32:
                x, y = px(ents[i])
                assert tuple(f[n, y, x]) == (159, 159, 223)             # (0.25, 0.25, 0.75) * 0.5 + 0.5
    # idempotent, and rendering does not touch the world state
    again = env.render("rgb_array", as_numpy=False)
    assert torch.equal(frames, again)
    st2 = env.get_state()
    for k in st:
        np.testing.assert_array_equal(st[k], st2[k])


def test_render_rejects_human_mode_and_oversize():
    env = make_env("simple_spread", 3, 3, 2)
    with pytest.raises(NotImplementedError):
        env.render("human")
    with pytest.raises(AssertionError):
        env.render("rgb_array", width=230, height=64)                   # width must be a multiple of 4


@pytest.mark.parametrize("algo,scenario", [("mappo", "simple_spread"), ("rmappo", "simple_reference")])
def test_runner_render_episode_frames_and_gif(tmp_path, algo, scenario):
    """runner.render (shared/mpe_runner.py:185-248): one frame after the reset and one per step, env 0; GIF on request."""
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.shared.mpe_runner import MPERunner
    M = 3 if scenario == "simple_spread" else 2
    a = parse_args(("--algorithm_name %s --scenario_name %s --num_agents %d --num_landmarks 3 --n_rollout_threads 1 "
                    "--episode_length 6 --render_episodes 2 --save_gifs --use_render --ifi 0.05 --seed 3"
                    % (algo, scenario, M)).split())
    envs = MPEVecEnv(a, device="cuda")
    runner = MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda"), "num_agents": M,
                        "run_dir": str(tmp_path)})
    frames, avg = runner.render()
    assert frames.shape == (2 * 7, 700, 700, 3) and frames.dtype == np.uint8 and np.isfinite(avg)
    assert (frames[0] != frames[7]).any(), "the second episode starts from a new reset state"
    # the last frame is the state the env holds now
    np.testing.assert_array_equal(frames[-1], envs.render("rgb_array", env_ids=[0])[0, 0])
    from PIL import Image
    gif = Image.open(str(tmp_path / "gifs" / "render.gif"))
    assert gif.n_frames == 14 and gif.size == (700, 700)
