#!/usr/bin/env python
"""Where the host-boundary iteration (bench.py `e2e`) spends its wall time: per-section host timers with a device
synchronise at each boundary (diagnostic only; not a bench number)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench


def main():
    n_envs = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    dev = torch.device("cuda", 0)
    a, envs, runner = bench.build_runner(n_envs, dev, 0, 1)
    runner.warmup()
    runner.run_episode()
    envs.set_host_io(True)
    M = bench.M_AGENTS
    pin = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt).pin_memory()
    hb = {"onehot": pin(n_envs, M, bench.N_ACT, dt=torch.uint8), "masks": pin(n_envs, M, 1),
          "row5": np.arange(n_envs * M, dtype=np.int64) * bench.N_ACT}
    from mappo_b200 import host_mirror
    bench.host_boundary_iteration(runner, envs, hb)
    b, pol = runner.buffer, runner.trainer.policy
    N = n_envs
    T = {}
    def tick(name, t0):
        torch.cuda.synchronize()
        T[name] = T.get(name, 0.0) + time.perf_counter() - t0
    # pinned-copy bandwidth of this box (what bounds the obs D2H of every step)
    big_d = torch.empty(64 << 20, dtype=torch.uint8, device=dev); big_h = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    for name, (dst, src) in (("d2h", (big_h, big_d)), ("h2d", (big_d, big_h))):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(10):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        print("pinned %s: %.1f GB/s" % (name, 10 * 64 / 1024 / (time.perf_counter() - t0)))
    t_all = time.perf_counter()
    for step in range(b.episode_length):
        t0 = time.perf_counter()
        values, actions, logp, rs, rsc = pol.get_actions(b.share_obs[step], b.obs[step].view(N * M, -1), None, None, None)
        tick("policy", t0); t0 = time.perf_counter()
        out = host_mirror.t2n_many({"values": values.view(N, M, 1), "actions": actions.view(N, M, 1), "logp": logp.view(N, M, 1)})
        values, actions, logp = out["values"], out["actions"], out["logp"]
        tick("d2h policy outputs (t2n)", t0); t0 = time.perf_counter()
        actions_env = host_mirror.onehot(actions, [bench.N_ACT])
        tick("host one-hot (uint8)", t0); t0 = time.perf_counter()
        obs, rewards, dones, infos = envs.step(actions_env)
        tick("envs.step (h2d + kernel + d2h)", t0); t0 = time.perf_counter()
        masks = hb["masks"].numpy(); masks[..., 0] = ~dones
        share_obs = np.broadcast_to(obs.reshape(N, 1, -1), (N, M, obs.shape[-1] * M))
        tick("host masks", t0); t0 = time.perf_counter()
        b.insert(share_obs, obs, None, None, actions, logp, values, rewards, masks)
        tick("buffer.insert", t0)
    t0 = time.perf_counter()
    runner.compute(); runner.train()
    tick("compute + train", t0)
    total = time.perf_counter() - t_all
    for k, v in T.items():
        print("%-34s %8.2f ms  %5.1f%%" % (k, v * 1e3, 100 * v / total))
    print("total %.2f ms" % (total * 1e3))


if __name__ == "__main__":
    main()
