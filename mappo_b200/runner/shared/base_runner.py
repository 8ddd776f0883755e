"""Runner mirror (onpolicy/runner/shared/base_runner.py:12-171): builds policy / trainer / buffer from the same
`config` dict and keeps compute / train / save / restore.  Logging writers (wandb / tensorboardX) are the
caller's business: `log_train` / `log_env` call `config['log_fn'](key, value, step)` if given."""
import os

import numpy as np
import torch

from ...algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy as Policy
from ...algorithms.r_mappo.r_mappo import R_MAPPO as TrainAlgo
from ...utils.shared_buffer import SharedReplayBuffer


class Runner(object):
    def __init__(self, config):
        self.all_args = config["all_args"]
        self.envs = config["envs"]
        self.eval_envs = config.get("eval_envs")
        self.device = config["device"]
        self.num_agents = config["num_agents"]
        a = self.all_args
        self.env_name = getattr(a, "env_name", "MPE")
        self.algorithm_name = getattr(a, "algorithm_name", "mappo")
        self.experiment_name = getattr(a, "experiment_name", "check")
        self.use_centralized_V = a.use_centralized_V
        self.num_env_steps = a.num_env_steps
        self.episode_length = a.episode_length
        self.n_rollout_threads = a.n_rollout_threads
        self.n_eval_rollout_threads = getattr(a, "n_eval_rollout_threads", 1)
        self.use_linear_lr_decay = a.use_linear_lr_decay
        self.hidden_size = a.hidden_size
        self.recurrent_N = a.recurrent_N
        self.save_interval = getattr(a, "save_interval", 1)
        self.use_eval = getattr(a, "use_eval", False)
        self.eval_interval = getattr(a, "eval_interval", 25)
        self.log_interval = getattr(a, "log_interval", 5)
        self.model_dir = getattr(a, "model_dir", None)
        self.run_dir = config.get("run_dir")
        self.save_dir = None
        if self.run_dir is not None:
            self.save_dir = os.path.join(str(self.run_dir), "models")
            os.makedirs(self.save_dir, exist_ok=True)
        self.log_fn = config.get("log_fn")
        share_observation_space = self.envs.share_observation_space[0] if self.use_centralized_V \
            else self.envs.observation_space[0]
        self.policy = Policy(a, self.envs.observation_space[0], share_observation_space, self.envs.action_space[0],
                             device=self.device)
        if self.model_dir is not None:
            self.restore()
        self.trainer = TrainAlgo(a, self.policy, device=self.device)
        if self.model_dir is not None and self.trainer._use_valuenorm:
            p = os.path.join(str(self.model_dir), "vnorm.pt")
            if os.path.exists(p):
                self.trainer.value_normalizer.load_state_dict(torch.load(p, map_location="cpu"))
        self.buffer = SharedReplayBuffer(a, self.num_agents, self.envs.observation_space[0], share_observation_space,
                                         self.envs.action_space[0], device=self.device)

    def run(self):
        raise NotImplementedError

    def warmup(self):
        raise NotImplementedError

    def collect(self, step):
        raise NotImplementedError

    def insert(self, data):
        raise NotImplementedError

    @torch.no_grad()
    def compute(self):
        """base_runner.py:110-118: bootstrap value of the last slot, then returns/GAE."""
        self.trainer.prep_rollout()
        b = self.buffer
        nv = torch.empty(b.n_rollout_threads, b.num_agents, 1, dtype=torch.float32, device=b.device)
        uniq = b.share_obs_unique()
        if b._recurrent:
            M = b.num_agents
            scratch = torch.empty_like(b.rnn_states_critic[-1])
            if uniq is not None and self.use_centralized_V:
                self.policy.critic.forward_into(uniq[-1], nv, rnn_in=b.rnn_states_critic[-1], masks=b.masks[-1],
                                                rnn_out=scratch, share_rows_div=M)
            else:
                x = b.share_obs[-1] if self.use_centralized_V else b.obs[-1]
                self.policy.critic.forward_into(x.reshape(-1, x.shape[-1]), nv, rnn_in=b.rnn_states_critic[-1],
                                                masks=b.masks[-1], rnn_out=scratch)
        elif uniq is not None:
            self.policy.critic.forward_into(uniq[-1], nv, group=b.num_agents)
        else:
            self.policy.critic.forward_into(b.share_obs[-1].reshape(-1, b.share_obs.shape[-1]), nv, group=1)
        b.compute_returns(nv, self.trainer.value_normalizer)

    def train(self):
        self.trainer.prep_training()
        train_infos = self.trainer.train(self.buffer)
        self.buffer.after_update()
        return train_infos

    def save(self):
        """Same three files and state_dict keys as base_runner.py:127-135."""
        if self.save_dir is None:
            return
        cpu = lambda sd: {k: v.detach().cpu() for k, v in sd.items()}
        torch.save(cpu(self.trainer.policy.actor.state_dict()), os.path.join(self.save_dir, "actor.pt"))
        torch.save(cpu(self.trainer.policy.critic.state_dict()), os.path.join(self.save_dir, "critic.pt"))
        if self.trainer._use_valuenorm:
            torch.save(cpu(self.trainer.value_normalizer.state_dict()), os.path.join(self.save_dir, "vnorm.pt"))

    def restore(self):
        """base_runner.py:137-146"""
        self.policy.actor.load_state_dict(torch.load(os.path.join(str(self.model_dir), "actor.pt"), map_location="cpu"))
        if not getattr(self.all_args, "use_render", False):
            self.policy.critic.load_state_dict(torch.load(os.path.join(str(self.model_dir), "critic.pt"),
                                                          map_location="cpu"))

    def log_train(self, train_infos, total_num_steps):
        if self.log_fn:
            for k, v in train_infos.items():
                self.log_fn(k, v, total_num_steps)

    def log_env(self, env_infos, total_num_steps):
        if self.log_fn:
            for k, v in env_infos.items():
                if len(v) > 0:
                    self.log_fn(k, float(np.mean(v)), total_num_steps)
