#!/usr/bin/env python
"""Short program for ncu captures: ONE C2-shaped iteration (65,536 envs x 3 agents, T = 25) with ONE ppo epoch after
one eager warm-up iteration.  Launch order of the library's kernels inside an iteration (what -k / -s / -c select):
  per rollout step: mlp_tc_kernel<0> (actor), mlp_tc_kernel<1> (critic), env_kernel      x 25
  compute:          mlp_tc_kernel<1> (bootstrap value), gae_kernel, stats_reduce_kernel
  train:            advantages_kernel, ..., mlp_tc_kernel<2>, grad_reduce, mlp_tc_kernel<3>, grad_reduce, adam x 2
    python tools/ncu_target.py [--envs N] [--agents M]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--agents", type=int, default=3)
    ap.add_argument("--epochs", type=int, default=1)
    ns = ap.parse_args()
    import torch
    bench.set_workload(ns.agents, None)
    bench.PPO_EPOCH = ns.epochs
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    _, envs, runner = bench.build_runner(ns.envs, dev, 0, 1)
    runner.warmup()
    runner.run_episode()
    torch.cuda.synchronize()
    info = runner.run_episode()
    torch.cuda.synchronize()
    print("ncu_target ok", {k: round(v, 5) for k, v in info.items()})


if __name__ == "__main__":
    main()
