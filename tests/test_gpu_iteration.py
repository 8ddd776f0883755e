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
