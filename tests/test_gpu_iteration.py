"""GPU: one whole MAPPO iteration (device env rollout, GAE + ValueNorm, PPO epochs on the tensor-core kernels) beside
the CPU oracle iteration (oracle/iteration.py) from identical weights, reset streams and sampling streams."""
import pytest

pytestmark = pytest.mark.gpu


def test_full_iteration_matches_oracle():
    import __graft_entry__ as g
    g._smoke_iteration()
