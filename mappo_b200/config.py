"""Flags of the hot path, with the reference's names, defaults and (inverted) store_false semantics
(onpolicy/config.py:156-287, scripts/train/train_mpe.py:52-61).  Only what the rollout / GAE / PPO-update path
reads is declared; unknown flags are ignored like the reference's `parse_known_args`."""
import argparse


def get_config():
    p = argparse.ArgumentParser(description="mappo_b200", formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("--algorithm_name", type=str, default="mappo", choices=["rmappo", "mappo", "ippo"])
    p.add_argument("--experiment_name", type=str, default="check")
    p.add_argument("--seed", type=int, default=1)
    p.add_argument("--n_rollout_threads", type=int, default=32)
    p.add_argument("--n_eval_rollout_threads", type=int, default=1)
    p.add_argument("--num_env_steps", type=int, default=int(10e6))
    p.add_argument("--env_name", type=str, default="MPE")
    p.add_argument("--episode_length", type=int, default=200)
    p.add_argument("--share_policy", action="store_false", default=True)
    p.add_argument("--use_centralized_V", action="store_false", default=True)
    p.add_argument("--hidden_size", type=int, default=64)
    p.add_argument("--layer_N", type=int, default=1)
    p.add_argument("--use_ReLU", action="store_false", default=True)
    p.add_argument("--use_popart", action="store_true", default=False)
    p.add_argument("--use_valuenorm", action="store_false", default=True)
    p.add_argument("--use_feature_normalization", action="store_false", default=True)
    p.add_argument("--use_orthogonal", action="store_false", default=True)
    p.add_argument("--gain", type=float, default=0.01)
    p.add_argument("--use_naive_recurrent_policy", action="store_true", default=False)
    p.add_argument("--use_recurrent_policy", action="store_false", default=True)
    p.add_argument("--recurrent_N", type=int, default=1)
    p.add_argument("--data_chunk_length", type=int, default=10)
    p.add_argument("--lr", type=float, default=5e-4)
    p.add_argument("--critic_lr", type=float, default=5e-4)
    p.add_argument("--opti_eps", type=float, default=1e-5)
    p.add_argument("--weight_decay", type=float, default=0)
    p.add_argument("--ppo_epoch", type=int, default=15)
    p.add_argument("--use_clipped_value_loss", action="store_false", default=True)
    p.add_argument("--clip_param", type=float, default=0.2)
    p.add_argument("--num_mini_batch", type=int, default=1)
    p.add_argument("--entropy_coef", type=float, default=0.01)
    p.add_argument("--value_loss_coef", type=float, default=1)
    p.add_argument("--use_max_grad_norm", action="store_false", default=True)
    p.add_argument("--max_grad_norm", type=float, default=10.0)
    p.add_argument("--use_gae", action="store_false", default=True)
    p.add_argument("--gamma", type=float, default=0.99)
    p.add_argument("--gae_lambda", type=float, default=0.95)
    p.add_argument("--use_proper_time_limits", action="store_true", default=False)
    p.add_argument("--use_huber_loss", action="store_false", default=True)
    p.add_argument("--use_value_active_masks", action="store_false", default=True)
    p.add_argument("--use_policy_active_masks", action="store_false", default=True)
    p.add_argument("--huber_delta", type=float, default=10.0)
    p.add_argument("--use_linear_lr_decay", action="store_true", default=False)
    p.add_argument("--save_interval", type=int, default=1)
    p.add_argument("--log_interval", type=int, default=5)
    p.add_argument("--use_eval", action="store_true", default=False)
    p.add_argument("--eval_interval", type=int, default=25)
    p.add_argument("--eval_episodes", type=int, default=32)
    p.add_argument("--model_dir", type=str, default=None)
    # not in the reference: replay the whole iteration as one captured CUDA graph (runner/shared/mpe_runner.py)
    p.add_argument("--use_cuda_graph", action="store_true", default=False)
    # render parameters (config.py:174,278-282)
    p.add_argument("--n_render_rollout_threads", type=int, default=1)
    p.add_argument("--save_gifs", action="store_true", default=False)
    p.add_argument("--use_render", action="store_true", default=False)
    p.add_argument("--render_episodes", type=int, default=5)
    p.add_argument("--ifi", type=float, default=0.1)
    # train_mpe.py:52-61
    p.add_argument("--scenario_name", type=str, default="simple_spread")
    p.add_argument("--num_landmarks", type=int, default=3)
    p.add_argument("--num_agents", type=int, default=2)
    return p


def parse_args(argv):
    """train_mpe.py:64-80: algorithm_name decides the recurrent flags (rmappo = GRU, mappo = MLP, ippo = decentralised
    critic with the recurrent flags left at their defaults)."""
    a = get_config().parse_known_args(argv)[0]
    if a.algorithm_name == "rmappo":
        a.use_recurrent_policy, a.use_naive_recurrent_policy = True, False
    elif a.algorithm_name == "mappo":
        a.use_recurrent_policy, a.use_naive_recurrent_policy = False, False
    elif a.algorithm_name == "ippo":
        a.use_centralized_V = False
    assert not (a.share_policy and a.scenario_name == "simple_speaker_listener"), (
        "The simple_speaker_listener scenario can not use shared policy. Please check the config.py.")
    return a
