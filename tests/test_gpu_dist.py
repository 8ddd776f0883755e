"""GPU, 2 ranks (skipped with fewer than 2 GPUs): the data-parallel update on hardware.  torchrun starts two ranks of
`bench.py --dp-check-only`: each rank rolls out and updates on its own 512 envs with the NCCL hooks attached; rank 0
then repeats the same iteration on ONE GPU with the union of the envs.  SURVEY.md section 8(e): the reduced gradient
equals the single-GPU gradient up to summation order, and the replicas stay bit-identical."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_update_equals_single_gpu_update_on_the_union_of_envs():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2); bench.py prints the same check as `dp_check` at every N > 1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "bench.py"), "--gpus", "2",
           "--dp-check-only"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    dp = json.loads(line)["dp_check"]
    assert dp["ranks"] == 2 and dp["replicas_identical"], dp
    assert dp["grad_vs_single_gpu_max_rel"] < 1e-5, dp
    assert dp["info_vs_single_gpu_max_rel"] < 1e-5, dp
    assert dp["params_vs_single_gpu_max_abs"] < 1e-6, dp
