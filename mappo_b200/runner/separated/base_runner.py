"""Separated-policy Runner mirror (onpolicy/runner/separated/base_runner.py:15-180): one Policy / TrainAlgo /
SeparatedReplayBuffer per agent, built from the same `config` dict; `compute`, `train`, `save`, `restore`.

Layout difference: all agents' buffers share one joint observation storage `joint_obs` [T+1, N, W] that the env kernel
writes once per step; agent m's `obs` is the column slice of it and its `share_obs` the whole row (what
runner/separated/mpe_runner.py:86-96 builds by concatenating and copying)."""
import os

import numpy as np
import torch

from ...algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy as Policy
from ...algorithms.r_mappo.r_mappo import R_MAPPO as TrainAlgo
from ...utils.separated_buffer import SeparatedReplayBuffer


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
        T, N, W = self.episode_length, self.n_rollout_threads, self.envs.obs_width
        self.joint_obs = torch.zeros(T + 1, N, W, dtype=torch.float32, device=self.device)
        self.policy, self.trainer, self.buffer = [], [], []
        seed = int(getattr(a, "seed", 1))
        off = 0
        for agent_id in range(self.num_agents):
            obs_space = self.envs.observation_space[agent_id]
            share_space = self.envs.share_observation_space[agent_id] if self.use_centralized_V else obs_space
            po = Policy(a, obs_space, share_space, self.envs.action_space[agent_id], device=self.device)
            # the reference draws every agent's actions from one global torch generator; here each agent's sampler is
            # a counter RNG keyed by (seed, env, step), so give each agent its own key
            po.actor.sample_seed = seed + 0x9E3779B1 * agent_id
            self.policy.append(po)
            off_m = off
            off += obs_space.shape[0]
            self.buffer.append(SeparatedReplayBuffer(a, obs_space, share_space, self.envs.action_space[agent_id],
                                                     device=self.device, joint_obs=self.joint_obs, obs_offset=off_m))
        if self.model_dir is not None:
            self.restore()
        for agent_id in range(self.num_agents):
            self.trainer.append(TrainAlgo(a, self.policy[agent_id], device=self.device))

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
        """base_runner.py:110-119: per agent, bootstrap value of the last slot, then returns/GAE."""
        for agent_id in range(self.num_agents):
            self.trainer[agent_id].prep_rollout()
            b = self.buffer[agent_id]
            nv = torch.empty(b.n_rollout_threads, 1, dtype=torch.float32, device=b.device)
            if b._recurrent:
                self.policy[agent_id].critic.forward_into(b.share_obs[-1], nv, rnn_in=b.rnn_states_critic[-1],
                                                          masks=b.masks[-1],
                                                          rnn_out=torch.empty_like(b.rnn_states_critic[-1]))
            else:
                self.policy[agent_id].critic.forward_into(b.share_obs[-1], nv, group=1)
            b.compute_returns(nv, self.trainer[agent_id].value_normalizer)

    def train(self):
        """base_runner.py:121-130.  The joint observation row is rolled once, by the last agent's after_update."""
        train_infos = []
        for agent_id in range(self.num_agents):
            self.trainer[agent_id].prep_training()
            train_infos.append(self.trainer[agent_id].train(self.buffer[agent_id]))
            self.buffer[agent_id].after_update(copy_obs=(agent_id == self.num_agents - 1))
        return train_infos

    def save(self):
        """Same file names and state_dict keys as base_runner.py:132-142."""
        if self.save_dir is None:
            return
        cpu = lambda sd: {k: v.detach().cpu() for k, v in sd.items()}
        for agent_id in range(self.num_agents):
            tr = self.trainer[agent_id]
            torch.save(cpu(tr.policy.actor.state_dict()), os.path.join(self.save_dir, "actor_agent%d.pt" % agent_id))
            torch.save(cpu(tr.policy.critic.state_dict()), os.path.join(self.save_dir, "critic_agent%d.pt" % agent_id))
            if tr._use_valuenorm:
                torch.save(cpu(tr.value_normalizer.state_dict()),
                           os.path.join(self.save_dir, "vnorm_agent%d.pt" % agent_id))

    def restore(self):
        """base_runner.py:144-153"""
        for agent_id in range(self.num_agents):
            d = str(self.model_dir)
            self.policy[agent_id].actor.load_state_dict(
                torch.load(os.path.join(d, "actor_agent%d.pt" % agent_id), map_location="cpu"))
            self.policy[agent_id].critic.load_state_dict(
                torch.load(os.path.join(d, "critic_agent%d.pt" % agent_id), map_location="cpu"))

    def log_train(self, train_infos, total_num_steps):
        """base_runner.py:155-160: keys `agent%i/<name>`."""
        if self.log_fn:
            for agent_id in range(self.num_agents):
                for k, v in train_infos[agent_id].items():
                    self.log_fn("agent%i/" % agent_id + k, v, total_num_steps)

    def log_env(self, env_infos, total_num_steps):
        if self.log_fn:
            for k, v in env_infos.items():
                if len(v) > 0:
                    self.log_fn(k, float(np.mean(v)), total_num_steps)
