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
    g._smoke_iteration(M=3, N=32, extra="--use_naive_recurrent_policy --num_mini_batch 2")
