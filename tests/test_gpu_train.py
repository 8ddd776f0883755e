"""GPU: the training entry point (mappo_b200/scripts/train_mpe.py, the reference's train_mpe.py command line) learns:
simple_spread with 3 agents, 2,048 envs, 60 iterations of the stock hyper-parameters -- the average episode reward
rises clearly above the random policy's, models and the scalar log land under the run directory."""
import json
import os

import pytest

pytestmark = pytest.mark.gpu


def test_train_mpe_learns_simple_spread(tmp_path, monkeypatch):
    from mappo_b200.scripts import train_mpe
    monkeypatch.chdir(tmp_path)
    T, N, iters = 25, 2048, 60
    argv = ("--env_name MPE --algorithm_name mappo --experiment_name t --scenario_name simple_spread --num_agents 3 "
            "--num_landmarks 3 --seed 1 --n_rollout_threads %d --num_mini_batch 1 --episode_length %d "
            "--num_env_steps %d --ppo_epoch 10 --use_ReLU --gain 0.01 --lr 7e-4 --critic_lr 7e-4 --log_interval 1 "
            "--save_interval 1000" % (N, T, T * N * iters)).split()
    train_mpe.main(argv)
    run = tmp_path / "results" / "MPE" / "simple_spread" / "mappo" / "t" / "run1"
    assert (run / "models" / "actor.pt").exists() and (run / "models" / "critic.pt").exists()
    rows = [json.loads(l) for l in open(run / "logs" / "summary.jsonl")]
    rew = [r["value"] for r in rows if r["key"] == "average_episode_rewards"]
    assert len(rew) == iters
    first, last = sum(rew[:3]) / 3, sum(rew[-3:]) / 3
    print("average episode reward: first 3 iterations %.1f, last 3 %.1f" % (first, last))
    assert last > first + 20.0, (first, last)
