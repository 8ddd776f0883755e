"""Harness that runs the UNMODIFIED reference (Chen001117/mappo, package `onpolicy`) on this box's host cores.

The reference lives in the git-ignored `baseline/_ref/` (made by tools/install_ref.sh in the build container; it
travels to the GPU box with the repo snapshot like the built .so) and is imported through the `sys.modules` shims of
SURVEY.md section 8(c) (tests/golden/ref_shims.py: gym, seaborn, imp, tensorboardX, imageio -- none of them is on this
image).  No reference file is edited.  What runs is the reference's own stack exactly as its training script builds
it (onpolicy/scripts/train/train_mpe.py:16-31,52-172): `SubprocVecEnv` of `MPEEnv` worker processes, `MPERunner`,
`R_MAPPOPolicy`, `R_MAPPO`, `SharedReplayBuffer`, torch on the CPU.

Used by `bench.py --impl reference` and bench.py's `cpu_baseline` leg only -- never by the product package."""
import os
import sys
import tempfile
import time
from pathlib import Path

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF_ROOT = os.path.join(HERE, "_ref")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "onpolicy", "runner", "shared"))


def install():
    """Shims + reference on sys.path.  Idempotent."""
    if not available():
        raise RuntimeError("baseline/_ref is missing: run tools/install_ref.sh in the build container")
    g = os.path.join(ROOT, "tests", "golden")
    if g not in sys.path:
        sys.path.insert(0, g)
    import ref_shims
    ref_shims.install(REF_ROOT)
    return ref_shims


def c1_flags(n_envs=128, n_training_threads=1, scenario="simple_spread", num_agents=3, num_landmarks=3,
             algo="mappo", ppo_epoch=10):
    """onpolicy/scripts/train_mpe_scripts/train_mpe_spread.sh:4-17 with BASELINE config C1's MLP policy (`mappo`).
    `--use_ReLU` is a store_false flag there (Tanh); `--cuda` (store_false) keeps the reference on the CPU and
    `--use_wandb` (store_false) selects its tensorboardX branch (shimmed) so nothing touches the network."""
    return ("--env_name MPE --algorithm_name %s --experiment_name check --scenario_name %s --num_agents %d "
            "--num_landmarks %d --seed 1 --n_training_threads %d --n_rollout_threads %d --num_mini_batch 1 "
            "--episode_length 25 --num_env_steps 20000000 --ppo_epoch %d --use_ReLU --gain 0.01 --lr 7e-4 "
            "--critic_lr 7e-4 --cuda --use_wandb" % (algo, scenario, num_agents, num_landmarks, n_training_threads,
                                                      n_envs, ppo_epoch)).split()


class ReferenceRun:
    """The reference's runner, built the way train_mpe.main() builds it, stepped one iteration at a time with the
    body of MPERunner.run() (runner/shared/mpe_runner.py:24-40) so that each iteration can be timed."""

    def __init__(self, flags):
        install()
        import numpy as np
        import torch
        from onpolicy.config import get_config
        from onpolicy.scripts.train import train_mpe
        a = train_mpe.parse_args(flags, get_config())
        if a.algorithm_name == "rmappo":                       # train_mpe.py:68-80
            a.use_recurrent_policy, a.use_naive_recurrent_policy = True, False
        elif a.algorithm_name == "mappo":
            a.use_recurrent_policy, a.use_naive_recurrent_policy = False, False
        elif a.algorithm_name == "ippo":
            a.use_centralized_V = False
        torch.set_num_threads(a.n_training_threads)            # train_mpe.py:94-96 (cpu branch)
        torch.manual_seed(a.seed); np.random.seed(a.seed)      # :131-133
        self.args = a
        self.envs = train_mpe.make_train_env(a)                # SubprocVecEnv x n_rollout_threads (:16-31)
        self.run_dir = Path(tempfile.mkdtemp(prefix="mappo_ref_run_"))
        from onpolicy.runner.shared.mpe_runner import MPERunner
        self.runner = MPERunner({"all_args": a, "envs": self.envs, "eval_envs": None, "num_agents": a.num_agents,
                                 "device": torch.device("cpu"), "run_dir": self.run_dir})
        self.runner.warmup()
        self.agent_steps = a.n_rollout_threads * a.episode_length * a.num_agents

    def iteration(self):
        """One pass of MPERunner.run()'s loop body; returns (seconds, split dict, train_infos)."""
        r = self.runner
        t_collect = t_env = t_insert = 0.0
        t0 = time.perf_counter()
        for step in range(r.episode_length):
            a0 = time.perf_counter()
            values, actions, action_log_probs, rnn_states, rnn_states_critic, actions_env = r.collect(step)
            a1 = time.perf_counter()
            obs, rewards, dones, infos = r.envs.step(actions_env)
            a2 = time.perf_counter()
            r.insert((obs, rewards, dones, infos, values, actions, action_log_probs, rnn_states, rnn_states_critic))
            a3 = time.perf_counter()
            t_collect += a1 - a0; t_env += a2 - a1; t_insert += a3 - a2
        b0 = time.perf_counter()
        r.compute()
        b1 = time.perf_counter()
        info = r.train()
        b2 = time.perf_counter()
        return b2 - t0, dict(collect=t_collect, env_step=t_env, insert=t_insert, compute=b1 - b0, train=b2 - b1), info

    def close(self):
        try:
            self.envs.close()
        except Exception:
            pass


def time_reference(flags, steps, warmup):
    """-> dict(value agent-steps/s, seconds per iteration, split of the median iteration, n_envs, ...)"""
    run = ReferenceRun(flags)
    try:
        for _ in range(warmup):
            run.iteration()
        times, splits, info = [], [], None
        t0 = time.perf_counter()
        for _ in range(steps):
            dt, split, info = run.iteration()
            times.append(dt); splits.append(split)
        total = time.perf_counter() - t0
    finally:
        run.close()
    order = sorted(range(len(times)), key=lambda i: times[i])
    med = order[len(order) // 2]
    return dict(value=run.agent_steps * steps / total, sec_per_iter=total / steps, sec_median=times[med],
                split=splits[med], n_envs=run.args.n_rollout_threads, agent_steps=run.agent_steps,
                train_info={k: float(v) for k, v in (info or {}).items()},
                n_training_threads=run.args.n_training_threads)


if __name__ == "__main__":
    import json
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    thr = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    print(json.dumps(time_reference(c1_flags(n, thr), steps=3, warmup=1)))
