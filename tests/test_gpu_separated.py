"""GPU: separated policies (BASELINE config C3, runner/separated/*): SeparatedReplayBuffer generators against the
reference's index fixtures, and whole iterations on simple_speaker_listener / simple_reference beside the CPU oracle
(oracle/iteration.py:OracleSeparatedMAPPO) from identical weights, reset streams and sampling streams."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(extra):
    from mappo_b200.config import parse_args
    return parse_args(extra.split())


def _buffer(T=6, N=5, D=4, S=7, A=3, extra=""):
    from mappo_b200.envs.spaces import Box, Discrete
    from mappo_b200.utils.separated_buffer import SeparatedReplayBuffer
    a = _args("--scenario_name simple_spread --episode_length %d --n_rollout_threads %d %s" % (T, N, extra))
    b = SeparatedReplayBuffer(a, Box(-np.inf, np.inf, (D,), np.float32), Box(-np.inf, np.inf, (S,), np.float32),
                              Discrete(A), device="cuda:0")
    g = torch.Generator().manual_seed(3)
    for name in ("obs", "share_obs", "value_preds", "returns", "actions", "action_log_probs", "rewards", "masks",
                 "active_masks"):
        t = getattr(b, name)
        t.copy_(torch.randn(t.shape, generator=g))
    return a, b


def test_generators_follow_oracle_rows():
    """separated_buffer.py:162-393: every yielded field equals the flat [T*N] array indexed by the oracle's rows."""
    from oracle import buffer as ob
    a, b = _buffer()
    T, N = 6, 5
    adv = torch.randn(T, N, 1, generator=torch.Generator().manual_seed(4)).cuda()
    flat = lambda x: x.reshape(T * N, -1).cpu().numpy()
    names = {0: "share_obs", 1: "obs", 4: "actions", 5: "value_preds", 6: "returns", 7: "masks", 8: "active_masks",
             9: "action_log_probs"}
    src = {i: flat(getattr(b, n)[:T]) for i, n in names.items()}
    src[10] = flat(adv)

    def check(gen, rows_list):
        got = list(gen)
        assert len(got) == len(rows_list)
        for sample, rows in zip(got, rows_list):
            rows = rows[0] if isinstance(rows, tuple) else rows
            for i, ref in src.items():
                np.testing.assert_array_equal(sample[i].cpu().numpy().reshape(len(rows), -1), ref[rows], err_msg=str(i))
            assert sample[11].shape == (len(rows), 3) and bool((sample[11] == 1).all())

    torch.manual_seed(9); perm = torch.randperm(T * N).numpy()
    torch.manual_seed(9); check(b.feed_forward_generator(adv, 2), ob.feed_forward_rows(T, N, 1, perm, 2))
    torch.manual_seed(9); perm = torch.randperm(T * N // 3).numpy()
    torch.manual_seed(9); check(b.recurrent_generator(adv, 2, 3), ob.separated_recurrent_rows(T, N, perm, 2, 3))
    torch.manual_seed(9); perm = torch.randperm(N).numpy()
    torch.manual_seed(9); check(b.naive_recurrent_generator(adv, 2), ob.naive_recurrent_rows(T, N, perm, 2))


def test_recurrent_rows_address_what_the_generators_yield():
    """The index forms the fused kernels consume (recurrent_rows / naive_recurrent_rows) against the arrays the
    reference-shaped generators yield under the same torch.randperm draw: every field row for row, and the initial
    states at the seed rows -- including the (Nc, L)-stack-flattened-as-(L, Nc) order of separated_buffer.py:361-391."""
    T, N = 6, 5
    a, b = _buffer(T, N, extra="--algorithm_name rmappo")
    assert b._recurrent
    g = torch.Generator().manual_seed(5)
    b.rnn_states.copy_(torch.randn(b.rnn_states.shape, generator=g))
    b.rnn_states_critic.copy_(torch.randn(b.rnn_states_critic.shape, generator=g))
    adv = torch.randn(T, N, 1, generator=g).cuda()
    v = b.flat_views()
    for gen, rows_fn in ((lambda: b.recurrent_generator(adv, 2, 3), lambda: b.recurrent_rows(2, 3)),
                         (lambda: b.naive_recurrent_generator(adv, 2), lambda: b.naive_recurrent_rows(2))):
        torch.manual_seed(9); samples = list(gen())
        torch.manual_seed(9); batches = list(rows_fn())
        assert len(samples) == len(batches) == 2
        for sample, (rows, nc, L, seed) in zip(samples, batches):
            assert rows.numel() == nc * L
            seed = rows[:nc] if seed is None else seed
            eq = lambda x, y: np.testing.assert_array_equal(x.cpu().numpy().reshape(y.shape[0], -1),
                                                            y.cpu().numpy().reshape(y.shape[0], -1))
            eq(sample[1], v["obs"][rows]); eq(sample[0], v["share"][rows]); eq(sample[4], v["actions"][rows])
            eq(sample[7], v["masks"][rows]); eq(sample[6], v["returns"][rows]); eq(sample[10], adv.view(-1)[rows])
            eq(sample[2], v["rnn"][seed]); eq(sample[3], v["rnn_critic"][seed])


def test_insert_after_update_roundtrip():
    """separated_buffer.py:72-110"""
    a, b = _buffer()
    T, N = 6, 5
    z = lambda *s: torch.full(s, 2.5)
    b.step = T - 1
    b.insert(z(N, 7), z(N, 4), None, None, z(N, 1), z(N, 1), z(N, 1), z(N, 1), torch.zeros(N, 1))
    assert b.step == 0
    assert bool((b.obs[T] == 2.5).all()) and bool((b.share_obs[T] == 2.5).all()) and bool((b.masks[T] == 0).all())
    assert bool((b.actions[T - 1] == 2.5).all()) and bool((b.rewards[T - 1] == 2.5).all())
    b.after_update()
    assert bool((b.obs[0] == 2.5).all()) and bool((b.share_obs[0] == 2.5).all()) and bool((b.masks[0] == 0).all())


@pytest.mark.parametrize("scenario,extra", [
    ("simple_speaker_listener", ""),
    ("simple_reference", ""),
    ("simple_speaker_listener", "--use_centralized_V"),      # store_false: decentralised critic
    ("simple_spread", "--num_mini_batch 2"),
    # the reference's own train_mpe_comm.sh: rmappo with separated policies (GRU, chunks of 10)
    ("simple_speaker_listener", "--algorithm_name rmappo --data_chunk_length 10"),
    ("simple_reference", "--algorithm_name rmappo --data_chunk_length 5 --num_mini_batch 2"),
    # whole-trajectory recurrent minibatches (ippo keeps the recurrent flags as given; decentralised critic)
    ("simple_speaker_listener", "--algorithm_name ippo --use_recurrent_policy --use_naive_recurrent_policy --num_mini_batch 2"),
])
def test_separated_iteration_matches_oracle(scenario, extra):
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.separated.mpe_runner import MPERunner
    from oracle.iteration import OracleSeparatedMAPPO
    N = 64
    M = 3 if scenario == "simple_spread" else 2
    a = _args("--scenario_name %s --num_agents %d --num_landmarks 3 --n_rollout_threads %d --episode_length 25 "
              "--ppo_epoch 2 --use_ReLU --share_policy --seed 7 %s" % (scenario, M, N, extra))
    assert a.share_policy is False
    torch.manual_seed(21)
    envs = MPEVecEnv(a, device="cuda:0")
    runner = MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda:0"), "num_agents": M, "run_dir": None})
    cpu = lambda sd: {k: v.detach().cpu().clone() for k, v in sd.items()}
    orc = OracleSeparatedMAPPO(a, [cpu(p.actor.state_dict()) for p in runner.policy],
                               [cpu(p.critic.state_dict()) for p in runner.policy],
                               sample_seeds=[p.actor.sample_seed for p in runner.policy])
    runner.warmup(); orc.warmup()
    recurrent = a.use_recurrent_policy or a.use_naive_recurrent_policy
    permute = a.num_mini_batch > 1 and not recurrent
    if permute:
        for b in runner.buffer:
            b.force_permutation = True
    # rollout + returns first (before training rolls the last slot to the front)
    for step in range(a.episode_length):
        values, actions, logps, rs, rsc, actions_env = runner.collect(step)
        obs, rew, masks = envs.step_ids(actions_env, out_obs=runner.joint_obs[step + 1], out_rewards=runner._rew_env,
                                        out_masks=runner._mask_env)
        runner.insert((obs, rew, masks, None, values, actions, logps, rs, rsc))
    runner.compute()
    orc.rollout(); orc.compute()
    for m in range(M):
        b, o = runner.buffer[m], orc.agents[m]
        same = (b.actions.cpu().numpy() == o.actions[:, :, 0]).mean()
        assert same > 0.998, "agent %d: sampled actions diverge from the oracle: %.4f" % (m, same)
        if same == 1.0:
            np.testing.assert_allclose(b.obs.cpu().numpy(), o.obs[:, :, 0], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(b.share_obs.cpu().numpy(), o.share[:, :, 0], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(b.rewards.cpu().numpy(), o.rewards[:, :, 0], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(b.returns.cpu().numpy(), o.returns[:, :, 0], rtol=1e-4, atol=1e-5)
            np.testing.assert_array_equal(b.masks.cpu().numpy(), o.masks[:, :, 0])
    torch.manual_seed(33)
    infos = runner.train()
    torch.manual_seed(33)
    refs = orc.train(permute=permute)
    for m in range(M):
        for k in ("value_loss", "policy_loss", "dist_entropy", "actor_grad_norm", "critic_grad_norm", "ratio"):
            np.testing.assert_allclose(infos[m][k], refs[m][k], rtol=3e-3, atol=3e-5, err_msg="agent %d %s" % (m, k))
        # after_update: slot 0 is the last observation
        np.testing.assert_allclose(runner.buffer[m].obs[0].cpu().numpy(), orc.agents[m].obs[0, :, 0], rtol=1e-5, atol=1e-6)
