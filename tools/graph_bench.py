"""Eager vs captured-graph iteration time (CUDA events, max of nothing: 1 GPU).  Usage: python tools/graph_bench.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mappo_b200.config import parse_args  # noqa: E402
from mappo_b200.envs.mpe.MPE_env import MPEVecEnv  # noqa: E402
from mappo_b200.runner.shared.mpe_runner import MPERunner  # noqa: E402


def run(N, M, graph, steps=10, warm=4, extra=""):
    a = parse_args(("--scenario_name simple_spread --num_agents %d --num_landmarks 3 --n_rollout_threads %d "
                    "--episode_length 25 --ppo_epoch 10 --lr 7e-4 --critic_lr 7e-4 --seed 1 %s %s"
                    % (M, N, "--use_cuda_graph" if graph else "", extra)).split())
    envs = MPEVecEnv(a, device="cuda:0")
    r = MPERunner({"all_args": a, "envs": envs, "device": torch.device("cuda:0"), "num_agents": M, "run_dir": None})
    r.warmup()
    for _ in range(warm):
        r.run_episode()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r.run_episode()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return dict(ms_per_iteration=ms, agent_steps_per_s=N * M * 25 / ms * 1e3)


if __name__ == "__main__":
    out = {}
    for name, N, M in (("C1_128envs", 128, 3), ("4096envs", 4096, 3), ("C2_65536envs", 65536, 3), ("C5_16384x10", 16384, 10)):
        out[name] = {"eager": run(N, M, False), "graph": run(N, M, True)}
    rm = "--algorithm_name rmappo --data_chunk_length 10"
    for name, N, M in (("rmappo_128envs", 128, 3), ("rmappo_4096envs", 4096, 3)):
        out[name] = {"eager": run(N, M, False, extra=rm), "graph": run(N, M, True, extra=rm)}
    print(json.dumps(out))
