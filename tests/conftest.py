import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# The oracle (torch CPU fp32) is the checker: keep it single-threaded.  With the intra-op pool of the B200 box's host
# (16 threads) one worker's ~32-row slice of a 700-row batch came out ~1e-4 off in about one process in eight
# (tools/flake_probe.py: the GPU output hash is identical in every process, the oracle's is not).
import torch  # noqa: E402

DEFAULT_TORCH_THREADS = torch.get_num_threads()
torch.set_num_threads(1)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # a gpu-marked test that is selected on a machine without CUDA must fail loudly, not skip:
    # the product has no CPU fallback.
    pass


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
