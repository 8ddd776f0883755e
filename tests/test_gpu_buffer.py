"""GPU: the device-resident SharedReplayBuffer (utils/shared_buffer.py) against fixtures recorded from the unmodified
reference (tests/golden/generators.npz): the three minibatch generators yield the reference's 12-tuples BIT FOR BIT
from the same torch seed (north_star level 1: minibatch indexing and bookkeeping), and insert / after_update / step keep
the reference's slot bookkeeping (shared_buffer.py:79-159)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FIELDS = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "value_preds", "returns", "masks",
          "active_masks", "old_action_log_probs", "adv_targ", "available_actions"]
BUF = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "action_log_probs", "value_preds", "returns",
       "masks", "active_masks"]


def _buffer(fx, space):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.spaces import Box, Discrete, MultiDiscrete
    from mappo_b200.utils.shared_buffer import SharedReplayBuffer
    T, N, M = int(fx["T"]), int(fx["N"]), int(fx["M"])
    a = parse_args(("--algorithm_name rmappo --scenario_name simple_spread --num_agents %d --episode_length %d "
                    "--n_rollout_threads %d --hidden_size 8" % (M, T, N)).split())
    act = Discrete(5) if space == "disc" else MultiDiscrete([[0, 4], [0, 9]])
    p = "shared_%s_buf_" % space
    D, S = fx[p + "obs"].shape[-1], fx[p + "share_obs"].shape[-1]
    b = SharedReplayBuffer(a, M, Box(-np.inf, np.inf, (D,), np.float32), Box(-np.inf, np.inf, (S,), np.float32), act,
                           device="cuda:0")
    for name in BUF:
        getattr(b, name).copy_(torch.from_numpy(fx[p + name]))
    if space == "disc":
        b._materialise_avail()
        b.available_actions.copy_(torch.from_numpy(fx[p + "available_actions"]))
    return b, torch.from_numpy(fx[p + "advantages"]).cuda()


@pytest.mark.parametrize("space", ["disc", "multi"])
def test_shared_generators_equal_the_reference_bit_for_bit(space, golden_dir):
    fx = np.load(os.path.join(golden_dir, "generators.npz"))
    b, adv = _buffer(fx, space)
    gens = (("ff", lambda: b.feed_forward_generator(adv, 2)), ("rec", lambda: b.recurrent_generator(adv, 2, 10)),
            ("naive", lambda: b.naive_recurrent_generator(adv, 2)))
    for tag, make in gens:
        torch.manual_seed(123)                      # the fixture drew torch.randperm right after this seed
        batches = list(make())
        assert len(batches) == 2
        for k, sample in enumerate(batches):
            assert len(sample) == 12
            for f, got in zip(FIELDS, sample):
                key = "shared_%s_%s_mb%d_%s" % (space, tag, k, f)
                if key not in fx.files:
                    assert got is None, key
                    continue
                np.testing.assert_array_equal(got.cpu().numpy().reshape(fx[key].shape), fx[key], err_msg=key)


def test_insert_after_update_and_step_bookkeeping():
    """shared_buffer.py:79-112 (slot step + 1 for obs / share_obs / rnn / masks, slot step for the rest, step wraps at
    episode_length) and :149-159 (after_update copies slot T to slot 0), with HOST arrays as the reference passes."""
    from mappo_b200.config import parse_args
    from mappo_b200.envs.spaces import Box, Discrete
    from mappo_b200.utils.shared_buffer import SharedReplayBuffer
    T, N, M, D, S = 3, 5, 2, 4, 9
    a = parse_args(("--algorithm_name rmappo --scenario_name simple_spread --num_agents %d --episode_length %d "
                    "--n_rollout_threads %d --hidden_size 64" % (M, T, N)).split())
    b = SharedReplayBuffer(a, M, Box(-np.inf, np.inf, (D,), np.float32), Box(-np.inf, np.inf, (S,), np.float32),
                           Discrete(5), device="cuda:0")
    rng = np.random.RandomState(0)
    rec = []
    for t in range(T):
        d = dict(share_obs=rng.randn(N, M, S), obs=rng.randn(N, M, D), rs=rng.randn(N, M, 1, 64), rsc=rng.randn(N, M, 1, 64),
                 actions=rng.randint(0, 5, (N, M, 1)).astype(np.float64), logp=rng.randn(N, M, 1), values=rng.randn(N, M, 1),
                 rewards=rng.randn(N, M, 1), masks=(rng.rand(N, M, 1) > 0.3).astype(np.float32),
                 active=(rng.rand(N, M, 1) > 0.1).astype(np.float32), avail=(rng.rand(N, M, 5) > 0.2).astype(np.float32))
        assert b.step == t
        b.insert(d["share_obs"], d["obs"], d["rs"], d["rsc"], d["actions"], d["logp"], d["values"], d["rewards"], d["masks"],
                 None, d["active"], d["avail"])
        rec.append(d)
    assert b.step == 0                                                   # (step + 1) % episode_length
    f32 = lambda x: np.asarray(x, dtype=np.float32)
    for t, d in enumerate(rec):
        for name, key, slot in (("share_obs", "share_obs", t + 1), ("obs", "obs", t + 1), ("rnn_states", "rs", t + 1),
                                ("rnn_states_critic", "rsc", t + 1), ("masks", "masks", t + 1),
                                ("active_masks", "active", t + 1), ("available_actions", "avail", t + 1),
                                ("actions", "actions", t), ("action_log_probs", "logp", t), ("value_preds", "values", t),
                                ("rewards", "rewards", t)):
            np.testing.assert_array_equal(getattr(b, name)[slot].cpu().numpy(), f32(d[key]), err_msg="%s[%d]" % (name, slot))
    assert bool((b.bad_masks == 1).all())                                # untouched when bad_masks is None
    last = {n: getattr(b, n)[-1].clone() for n in ("share_obs", "obs", "rnn_states", "rnn_states_critic", "masks",
                                                   "bad_masks", "active_masks", "available_actions")}
    b.after_update()
    for n, want in last.items():
        assert torch.equal(getattr(b, n)[0], want), n
