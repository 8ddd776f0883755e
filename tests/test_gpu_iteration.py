"""GPU: one whole MAPPO iteration (device env rollout, GAE + ValueNorm, PPO epochs on the tensor-core kernels) beside
the CPU oracle iteration (oracle/iteration.py) from identical weights, reset streams and sampling streams."""
import pytest

pytestmark = pytest.mark.gpu


def test_full_iteration_matches_oracle():
    import __graft_entry__ as g
    g._smoke_iteration()


def test_full_iteration_ten_agents_wide_critic():
    """BASELINE config C5 shape: M = 10 (obs 46, share_obs 460 -> the critic's wide first layer)."""
    import __graft_entry__ as g
    g._smoke_iteration(M=10, N=48)


def test_full_iteration_recurrent_gru_reference_scenario():
    """BASELINE config C4 shape: simple_reference, shared GRU actor / critic (rmappo), MultiDiscrete heads, chunks of 10
    steps cut from the [N, M, T] stream (straddling trajectories, as in the reference)."""
    import __graft_entry__ as g
    g._smoke_iteration(M=2, N=64, scenario="simple_reference", extra="--algorithm_name rmappo --data_chunk_length 10")


def test_full_iteration_naive_recurrent():
    import __graft_entry__ as g
    g._smoke_iteration(M=3, N=32, extra="--algorithm_name ippo --use_recurrent_policy --use_naive_recurrent_policy --num_mini_batch 2")


def test_full_iteration_ippo_decentralised_critic():
    """ippo with a shared MLP policy: the critic reads each agent's own observation (share_obs is obs,
    mpe_runner.py:136-137)."""
    import __graft_entry__ as g
    g._smoke_iteration(M=3, N=64, extra="--algorithm_name ippo --use_recurrent_policy")


def test_full_iteration_wide_observations():
    """15 agents: obs_dim = 4 + 2 L + 4 (M - 1) = 66 > 64, so actor and (decentralised, ippo) critic both take the wide
    first-layer chain, rollout and update, beside the oracle."""
    import __graft_entry__ as g
    g._smoke_iteration(M=15, N=24, extra="--algorithm_name ippo --use_recurrent_policy")
    # centralised critic over 15 x 66 = 990 columns (first layer in two 32-column halves)
    g._smoke_iteration(M=15, N=24)


def test_full_iteration_recurrent_wide_inputs():
    """GRU policies whose base reads more than 64 columns: rmappo with 5 agents (centralised critic over 5 x 26 = 130
    columns) and ippo's default recurrent policy with 15 agents (actor and critic over 66 columns)."""
    import __graft_entry__ as g
    g._smoke_iteration(M=5, N=32, extra="--algorithm_name rmappo --data_chunk_length 10")
    g._smoke_iteration(M=15, N=16, extra="--algorithm_name ippo --data_chunk_length 5")


def test_full_iteration_two_gru_layers():
    """--recurrent_N 2 (config.py:219, rnn.py:13): stacked GRU layers through rollout (states [N, M, 2, 64]), chunked
    update with back-propagation through both layers, beside the oracle iteration."""
    import __graft_entry__ as g
    g._smoke_iteration(M=2, N=64, scenario="simple_reference",
                       extra="--algorithm_name rmappo --data_chunk_length 10 --recurrent_N 2")
    g._smoke_iteration(M=3, N=32, extra="--algorithm_name rmappo --data_chunk_length 5 --recurrent_N 2")


def test_full_iteration_two_and_three_hidden_blocks():
    """--layer_N 2 / 3 (config.py:201, mlp.py:20-28): rollout and update through the fp32 tile kernel with extra hidden
    blocks, beside the oracle iteration."""
    import __graft_entry__ as g
    g._smoke_iteration(M=3, N=64, extra="--layer_N 2")
    g._smoke_iteration(M=3, N=48, extra="--layer_N 3 --algorithm_name ippo --use_recurrent_policy")


def _make_runner(extra, N=192, M=3):
    import torch
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.shared.mpe_runner import MPERunner
    a = parse_args(("--scenario_name simple_spread --num_agents %d --num_landmarks 3 --n_rollout_threads %d "
                    "--episode_length 25 --ppo_epoch 3 --lr 7e-4 --critic_lr 7e-4 --seed 5 %s" % (M, N, extra)).split())
    torch.manual_seed(11)
    envs = MPEVecEnv(a, device="cuda:0")
    r = MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda:0"), "num_agents": M, "run_dir": None})
    r.warmup()
    return r


@pytest.mark.parametrize("M,extra", [(3, ""), (10, ""), (3, "--algorithm_name rmappo --data_chunk_length 10")])
def test_cuda_graph_iteration_equals_eager(M, extra):
    """--use_cuda_graph: iteration 1 eager with device-resident counters, iteration 2 captured,
    iterations 3+ replayed.  Same Philox streams and the same kernels as the eager runner, so the sampled actions are
    identical; the Adam bias corrections come from pow() on the device instead of the host (<= 1 ulp of float32 apart),
    hence a tolerance on the parameters instead of equality.  rmappo: with num_mini_batch == 1 the chunks are taken in
    storage order (their permutation only reorders the sums), so the recurrent update is capturable too."""
    import numpy as np
    import torch
    eager, graph = _make_runner(extra, M=M), _make_runner("--use_cuda_graph " + extra, M=M)
    assert graph._graph_eligible() and not eager._graph_eligible()
    for it in range(5):
        ie, ig = eager.run_episode(), graph.run_episode()
        same = (eager.buffer.actions == graph.buffer.actions).float().mean().item()
        assert same > 0.999, "iteration %d: sampled actions diverge (%.5f)" % (it, same)
        for k in ie:
            np.testing.assert_allclose(ig[k], ie[k], rtol=2e-4, atol=1e-6, err_msg="iteration %d %s" % (it, k))
    assert graph._graph_state == 2 and graph._graph is not None
    pe, pg = eager.policy, graph.policy
    torch.testing.assert_close(pg.actor.flat, pe.actor.flat, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(pg.critic.flat, pe.critic.flat, rtol=1e-4, atol=1e-6)
    # host mirrors of the device counters stay in step
    assert pg.actor._calls == pe.actor._calls == int(pg.actor.step_dev.item())
    assert pg.actor_optimizer.step_count == pe.actor_optimizer.step_count == int(pg.actor_optimizer.step_dev.item())
    assert pg.critic_optimizer.step_count == int(pg.critic_optimizer.step_dev.item())
    torch.testing.assert_close(graph.trainer.value_normalizer.stats, eager.trainer.value_normalizer.stats, rtol=1e-5, atol=0)


def test_cuda_graph_not_used_where_the_host_draws_permutations():
    r = _make_runner("--use_cuda_graph --num_mini_batch 2")
    assert not r._graph_eligible()
    r.run_episode()
    assert r._graph_state == 0
