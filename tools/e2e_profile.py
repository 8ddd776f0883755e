#!/usr/bin/env python
"""cProfile of the host side of bench.py's host-boundary iteration (diagnostic)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench


def main():
    n_envs = 65536
    dev = torch.device("cuda", 0)
    a, envs, runner = bench.build_runner(n_envs, dev, 0, 1)
    runner.warmup(); runner.run_episode()
    envs.set_host_io(True)
    pin = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt).pin_memory()
    hb = {"masks": pin(n_envs, bench.M_AGENTS, 1)}
    bench.host_boundary_iteration(runner, envs, hb)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        bench.host_boundary_iteration(runner, envs, hb)
    torch.cuda.synchronize()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)


if __name__ == "__main__":
    main()
