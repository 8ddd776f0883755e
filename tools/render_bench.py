"""Rasteriser throughput (csrc/render.cu): frames/s and written GB/s for n frames of 700 x 700, CUDA events, L2-cold
(the output of one launch, n x 1.47 MB, exceeds the 126 MB L2 for n >= 128).  Usage: python tools/render_bench.py [n]"""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mappo_b200.envs.mpe import rendering  # noqa: E402
from mappo_b200.envs.mpe.MPE_env import MPEVecEnv  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    res = {}
    for scenario, M, L in (("simple_spread", 3, 3), ("simple_spread", 10, 3), ("simple_reference", 2, 3)):
        a = types.SimpleNamespace(scenario_name=scenario, num_agents=M, num_landmarks=L, episode_length=25,
                                  n_rollout_threads=n, seed=1)
        env = MPEVecEnv(a, device="cuda")
        env.reset()
        out = torch.empty(n, 700, 700, 3, dtype=torch.uint8, device="cuda")
        for _ in range(3):
            rendering.render_rgb(env, None, out=out)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 10
        ev[0].record()
        for _ in range(reps):
            rendering.render_rgb(env, None, out=out)
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / reps
        res["%s_M%d" % (scenario, M)] = dict(frames=n, ms_per_launch=ms, frames_per_s=n / ms * 1e3,
                                             written_GBps=out.numel() / ms * 1e-6)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
