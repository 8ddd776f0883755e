"""GPU: the host-array boundary (what bench.py's `e2e` drives): pass-through of arrays the library itself brought out
of the device (mappo_b200/host_mirror.py), the ordinary host -> device path for any other array, and the one-launch
one-hot -> id decode (mappo_onehot_to_ids) against the reference's np.eye(n)[actions] convention
(runner/shared/mpe_runner.py:110-121)."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _env_and_buffer(N=64, scenario="simple_spread", M=3):
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.utils.shared_buffer import SharedReplayBuffer
    a = parse_args(("--scenario_name %s --num_agents %d --num_landmarks 3 --n_rollout_threads %d --episode_length 25"
                    % (scenario, M, N)).split())
    env = MPEVecEnv(a, device="cuda:0", host_io=True)
    buf = SharedReplayBuffer(a, M, env.observation_space[0], env.share_observation_space[0], env.action_space[0], device="cuda:0")
    return a, env, buf


def test_pass_through_and_ordinary_host_arrays_give_the_same_buffer():
    from mappo_b200 import host_mirror
    a, env, buf = _env_and_buffer()
    N, M = 64, 3
    env.reset()
    rng = np.random.RandomState(0)
    ids = rng.randint(0, 5, (N, M))
    obs, rew, dones, infos = env.step(np.eye(5)[ids])                        # float64 one-hot, like the reference
    assert isinstance(obs, np.ndarray) and host_mirror.lookup(obs) is not None and host_mirror.lookup(rew) is not None
    zeros = np.zeros((N, M, 1), np.float32)
    share = np.broadcast_to(obs.reshape(N, 1, -1), (N, M, obs.shape[-1] * M))
    # (1) the registered arrays themselves: served device -> device
    buf.insert(share, obs, None, None, zeros, zeros, zeros, rew, 1.0 - dones[..., None].astype(np.float32))
    got1, r1 = buf.obs[1].cpu().numpy().copy(), buf.rewards[0].cpu().numpy().copy()
    np.testing.assert_array_equal(got1, obs); np.testing.assert_array_equal(r1, rew)
    # (2) a DIFFERENT host array with different contents takes the host -> device path and round-trips
    other = obs.copy() + 1.0
    assert host_mirror.lookup(other) is None
    buf.step = 0
    buf.insert(share, other, None, None, zeros, zeros, zeros, rew.copy() * 2, zeros + 1)
    np.testing.assert_array_equal(buf.obs[1].cpu().numpy(), other)
    np.testing.assert_array_equal(buf.rewards[0].cpu().numpy(), rew * 2)
    # (3) an in-place edit of a registered array must be declared: forget() sends it over PCIe again
    obs += 3.0
    host_mirror.forget(obs)
    buf.step = 0
    buf.insert(share, obs, None, None, zeros, zeros, zeros, rew, zeros + 1)
    np.testing.assert_array_equal(buf.obs[1].cpu().numpy(), obs)


def test_t2n_twin_feeds_insert_and_env_step():
    from mappo_b200 import host_mirror
    a, env, buf = _env_and_buffer()
    N, M = 64, 3
    env.reset()
    dev_actions = torch.randint(0, 5, (N, M, 1), device="cuda:0")
    dev_vals = torch.randn(N, M, 1, device="cuda:0")
    acts, vals = host_mirror.t2n(dev_actions, "actions"), host_mirror.t2n(dev_vals, "values")
    assert acts.dtype == np.int64 and (acts == dev_actions.cpu().numpy()).all()
    assert host_mirror.lookup(acts).data_ptr() == dev_actions.data_ptr()
    assert host_mirror.lookup(vals).data_ptr() == dev_vals.data_ptr()
    obs, rew, dones, _ = env.step(acts)                                     # ids form, served by the twin
    env2_ids = env._ids_scratch if env._ids_scratch is not None else None
    zeros = np.zeros((N, M, 1), np.float32)
    buf.insert(None, obs, None, None, acts, zeros, vals, rew, zeros + 1)
    np.testing.assert_array_equal(buf.actions[0].cpu().numpy(), acts.astype(np.float32))
    np.testing.assert_array_equal(buf.value_preds[0].cpu().numpy(), vals)


@pytest.mark.parametrize("dtype", [np.uint8, np.float32, np.float64])
def test_onehot_decode_equals_argmax(dtype):
    from mappo_b200 import gather
    rng = np.random.RandomState(1)
    rows = 100_003
    for sizes in ([5], [5, 10], [3]):
        ids = np.stack([rng.randint(0, s, rows) for s in sizes], -1)
        oh = np.concatenate([np.eye(s)[ids[:, h]] for h, s in enumerate(sizes)], -1).astype(dtype)
        out = torch.full((rows, len(sizes) + 1), -1.0, device="cuda:0")
        gather.onehot_to_ids(torch.from_numpy(oh).cuda(), sizes, out)
        np.testing.assert_array_equal(out[:, :len(sizes)].cpu().numpy(), ids.astype(np.float32))
        assert (out[:, len(sizes)] == -1).all()


def test_env_step_accepts_every_one_hot_dtype_and_ids():
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    a = types.SimpleNamespace(scenario_name="simple_reference", num_agents=2, num_landmarks=3, episode_length=25,
                              n_rollout_threads=32, seed=4)
    rng = np.random.RandomState(2)
    ids = np.stack([rng.randint(0, 5, (32, 2)), rng.randint(0, 10, (32, 2))], -1)           # MultiDiscrete [5, 10]
    oh = np.concatenate([np.eye(5)[ids[..., 0]], np.eye(10)[ids[..., 1]]], -1)
    outs = []
    for arg in (ids.astype(np.float32), oh, oh.astype(np.float32), oh.astype(np.uint8), torch.as_tensor(oh).cuda()):
        env = MPEVecEnv(a, device="cuda:0", host_io=True)
        env.reset()
        outs.append(env.step(arg)[0].copy())
    for o in outs[1:]:
        np.testing.assert_array_equal(o, outs[0])


def test_copy_segments_and_batched_insert_equal_plain_copies():
    """mappo_copy_segments: up to 8 device copies per launch (SharedReplayBuffer.insert batches its stores through it);
    aligned (16-byte pieces) and 4-byte-aligned segments, more than 8 pairs, odd sizes."""
    from mappo_b200 import gather
    g = torch.Generator(device="cuda").manual_seed(3)
    sizes = [1, 3, 4, 5, 1023, 4096, 65536 * 54 + 3, 7, 128, 129, 1000001]
    base = [torch.randn(n + 1, device="cuda", generator=g) for n in sizes]
    srcs = [b[1:] if i % 2 else b[:-1] for i, b in enumerate(base)]                 # every other one 4 bytes off a 16-byte boundary
    dsts = [torch.full((n + 2,), -7.0, device="cuda") for n in sizes]
    gather.copy_segments([(d[1:-1] if i % 3 == 0 else d[:n], s) for i, (d, s, n) in enumerate(zip(dsts, srcs, sizes))])
    for i, (d, s, n) in enumerate(zip(dsts, srcs, sizes)):
        got = d[1:-1] if i % 3 == 0 else d[:n]
        assert torch.equal(got, s), (i, n)
        assert (d[-1] == -7.0) and (i % 3 != 0 or d[0] == -7.0), "wrote outside segment %d" % i


def test_host_onehot_from_float_ids_equals_the_int64_form():
    from mappo_b200 import host_mirror
    rng = np.random.default_rng(0)
    for heads in ([5], [3, 5], [5, 10]):
        ids = np.stack([rng.integers(0, h, size=(64, 3)) for h in heads], axis=-1)
        a = host_mirror.onehot(ids.astype(np.int64), heads, tag="oh_i").copy()
        b = host_mirror.onehot(ids.astype(np.float32), heads, tag="oh_f").copy()
        want = np.concatenate([np.eye(h, dtype=np.uint8)[ids[..., k]] for k, h in enumerate(heads)], axis=-1)
        assert a.dtype == b.dtype == np.uint8 and (a == want).all() and (b == want).all()


def test_onehot_of_pass_through_ids_is_built_on_the_device_and_feeds_the_env():
    """ids that are a host copy the library made (t2n) -> host_mirror.onehot builds the one-hot on the device, returns the
    same bytes as the host loop, and registers the device copy so that MPEVecEnv.step uploads nothing."""
    from mappo_b200 import host_mirror
    g = torch.Generator(device="cuda").manual_seed(5)
    for heads in ([5], [3, 5]):
        ids_dev = torch.stack([torch.randint(0, h, (96, 3), device="cuda", generator=g) for h in heads], -1)
        if len(heads) == 2:
            ids_dev = ids_dev.float()                        # int64 (what the policy returns) and float32 ids
        ids_host = host_mirror.t2n(ids_dev, tag="ids%d" % len(heads))
        assert host_mirror.lookup(ids_host) is not None
        oh = host_mirror.onehot(ids_host, heads, tag="ohd%d" % len(heads))
        want = np.concatenate([np.eye(h, dtype=np.uint8)[ids_host[..., k].astype(int)] for k, h in enumerate(heads)], -1)
        assert oh.dtype == np.uint8 and oh.shape == want.shape and (oh == want).all()
        twin = host_mirror.lookup(oh)
        assert twin is not None and twin.is_cuda and (twin.cpu().numpy() == want).all()
        # an ordinary array of the same values takes the host loop and is NOT registered
        oh2 = host_mirror.onehot(ids_host.copy(), heads, tag="ohh%d" % len(heads))
        assert (oh2 == want).all() and host_mirror.lookup(oh2) is None


def test_onehot_of_pass_through_ids_refuses_an_id_outside_its_head():
    """np.eye(5)[7] raises IndexError in the reference's runner: so do both forms of the drop-in."""
    import pytest as _pt
    from mappo_b200 import host_mirror
    ids_dev = torch.tensor([[0.0], [4.0], [7.0]], device="cuda")
    ids_host = host_mirror.t2n(ids_dev, tag="ids_bad")
    with _pt.raises(IndexError):
        host_mirror.onehot(ids_host, [5], tag="oh_bad")
    ok = host_mirror.onehot(host_mirror.t2n(torch.tensor([[0.0], [4.0]], device="cuda"), tag="ids_ok"), [5], tag="oh_bad2")
    assert (ok == np.eye(5, dtype=np.uint8)[[0, 4]]).all()
    with _pt.raises(Exception):
        host_mirror.onehot(np.array([[0], [7]], dtype=np.int64), [5], tag="oh_bad3")
