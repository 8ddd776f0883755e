#!/usr/bin/env python
"""Training entry point with the reference's command line (onpolicy/scripts/train/train_mpe.py:17-176):

    python -m mappo_b200.scripts.train_mpe --env_name MPE --algorithm_name rmappo --scenario_name simple_spread \
        --num_agents 3 --num_landmarks 3 --n_rollout_threads 4096 --episode_length 25 --num_env_steps 20000000 \
        --ppo_epoch 10 --num_mini_batch 1 --use_ReLU --gain 0.01 --lr 7e-4 --critic_lr 7e-4

`make_train_env` / `make_eval_env` return the device-resident batched env (one object instead of n_rollout_threads
subprocesses, train_mpe.py:17-56), the runner is chosen by `--share_policy` exactly as there (:152-156), models go
to `<run_dir>/models` under the reference's file names.  Logging: wandb / tensorboardX are not installed here, so
scalars go to `<run_dir>/logs/summary.jsonl` (one JSON object per logged key).  One process drives one GPU; under
`torchrun` every rank takes its shard of the envs and the gradients are all-reduced (mappo_b200/parallel.py)."""
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch

from ..config import parse_args


def make_train_env(all_args, device, env_id_offset=0):
    """train_mpe.py:17-35: MPE only."""
    from ..envs.mpe.MPE_env import MPEVecEnv
    if all_args.env_name != "MPE":
        print("Can not support the " + all_args.env_name + "environment.")
        raise NotImplementedError
    return MPEVecEnv(all_args, device=device, env_id_offset=env_id_offset)


def make_eval_env(all_args, device):
    """train_mpe.py:38-56: n_eval_rollout_threads envs seeded seed * 50000 + rank * 10000."""
    from ..envs.mpe.MPE_env import MPEVecEnv
    import copy
    a = copy.copy(all_args)
    a.n_rollout_threads = all_args.n_eval_rollout_threads
    a.seed = all_args.seed * 50000
    return MPEVecEnv(a, device=device)


def next_run_dir(all_args, root=None):
    """train_mpe.py:98-128: results/<env>/<scenario>/<algorithm>/<experiment>/run<k>."""
    root = Path(root or os.path.join(os.getcwd(), "results"))
    base = root / all_args.env_name / all_args.scenario_name / all_args.algorithm_name / all_args.experiment_name
    base.mkdir(parents=True, exist_ok=True)
    nums = [int(f.name.split("run")[1]) for f in base.iterdir() if f.name.startswith("run") and f.name[3:].isdigit()]
    run_dir = base / ("run%i" % (max(nums) + 1 if nums else 1))
    run_dir.mkdir(parents=True, exist_ok=True)
    return run_dir


def main(argv):
    all_args = parse_args(argv)
    if not torch.cuda.is_available():
        raise RuntimeError("mappo_b200 has no CPU path: a B200 (sm_100a) is required")
    from .. import parallel
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(device)
    rank, world = parallel.init_distributed("nccl", device)  # (0, 1) without torchrun
    run_dir = next_run_dir(all_args) if rank == 0 else None
    torch.manual_seed(all_args.seed)                          # same initial weights on every rank
    torch.cuda.manual_seed_all(all_args.seed)
    np.random.seed(all_args.seed)
    envs = make_train_env(all_args, device, parallel.env_shard(all_args.n_rollout_threads, rank))
    eval_envs = make_eval_env(all_args, device) if all_args.use_eval else None
    log = None
    if run_dir is not None:
        (run_dir / "logs").mkdir(exist_ok=True)
        log = open(run_dir / "logs" / "summary.jsonl", "w")
    config = {"all_args": all_args, "envs": envs, "eval_envs": eval_envs, "num_agents": all_args.num_agents,
              "device": device, "run_dir": run_dir,
              "log_fn": (lambda k, v, step: (log.write(json.dumps({"key": k, "value": v, "step": step}) + "\n"),
                                             log.flush())) if log else None}
    if all_args.share_policy:
        from ..runner.shared.mpe_runner import MPERunner as Runner
    else:
        from ..runner.separated.mpe_runner import MPERunner as Runner
    runner = Runner(config)
    for tr in (runner.trainer if isinstance(runner.trainer, list) else [runner.trainer]):
        parallel.attach(tr, world)
    infos = runner.run()
    envs.close()
    if eval_envs is not None and eval_envs is not envs:
        eval_envs.close()
    if log:
        log.close()
    return infos


if __name__ == "__main__":
    main(sys.argv[1:])
