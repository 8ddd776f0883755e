"""CPU: the separated-policy oracle (oracle/iteration.py:OracleSeparatedMAPPO) -- structure, determinism and its
agreement with the shared oracle where the two settings coincide."""
import numpy as np
import torch

from mappo_b200.algorithms.r_mappo.algorithm.r_actor_critic import _reference_init
from mappo_b200.config import parse_args
from oracle.iteration import OracleMAPPO, OracleSeparatedMAPPO


def make_args(**kw):
    flags = []
    for k, v in kw.items():
        if isinstance(v, bool):
            flags.append("--" + k)                  # store_true / store_false flags: presence flips the default
        else:
            flags += ["--" + k, str(v)]
    return parse_args(flags)


def random_state_dicts(args, d, s, heads, seed):
    torch.manual_seed(seed)
    return _reference_init(args, d, heads, False), _reference_init(args, s, [1], True)


def _sds(args, dims, S_list, acts, seed):
    out_a, out_c = [], []
    for m, (d, s, a) in enumerate(zip(dims, S_list, acts)):
        asd, csd = random_state_dicts(args, d, s, a, seed + m)
        out_a.append(asd); out_c.append(csd)
    return out_a, out_c


def test_speaker_listener_iteration_runs_and_is_deterministic():
    a = make_args(scenario_name="simple_speaker_listener", num_agents=2, num_landmarks=3, n_rollout_threads=8,
                  episode_length=25, ppo_epoch=2, share_policy=True, use_ReLU=True)
    asd, csd = _sds(a, [3, 11], [14, 14], [[3], [5]], 5)
    infos = []
    for _ in range(2):
        o = OracleSeparatedMAPPO(a, asd, csd)
        o.warmup()
        infos.append(o.iteration())
        assert o.agents[0].obs.shape == (26, 8, 1, 3) and o.agents[1].share.shape == (26, 8, 1, 14)
        np.testing.assert_array_equal(o.agents[0].share, o.agents[1].share)
        np.testing.assert_array_equal(o.agents[0].share[..., :3], o.agents[0].obs)
        np.testing.assert_array_equal(o.agents[1].share[..., 3:], o.agents[1].obs)
        np.testing.assert_array_equal(o.agents[0].rewards, o.agents[1].rewards)       # collaborative reward
        assert o.agents[0].actions.max() <= 2 and o.agents[1].actions.max() <= 4
    for m in range(2):
        for k, v in infos[0][m].items():
            assert np.isfinite(v) and v == infos[1][m][k]


def test_one_agent_separated_equals_shared_on_that_agent():
    """With identical networks for every agent and a decentralised critic, agent m of the separated setting sees
    exactly the rows (t, n, m) of the shared setting: same values / log-probs on the same observations."""
    a = make_args(scenario_name="simple_spread", num_agents=3, num_landmarks=3, n_rollout_threads=4,
                  episode_length=5, ppo_epoch=1, use_centralized_V=True, use_ReLU=True)
    asd, csd = random_state_dicts(a, 18, 18, [5], 3)
    sh = OracleMAPPO(a, asd, csd)
    sep = OracleSeparatedMAPPO(a, [asd] * 3, [csd] * 3)
    sep.warmup()
    obs = np.stack([ag.obs[0, :, 0] for ag in sep.agents], axis=1)        # [N, M, D]
    sh.obs[0] = obs
    sh.share_obs_of = lambda o: o                                          # decentralised: share_obs is obs
    v_sh, _, _ = sh.collect(0)
    for m, ag in enumerate(sep.agents):
        v, _, _ = ag.collect(0)
        np.testing.assert_allclose(v[:, 0], v_sh[:, m], rtol=1e-6, atol=1e-7)
