"""Separated MPERunner mirror (onpolicy/runner/separated/mpe_runner.py:13-270), device-resident.

Per step: one actor + one critic launch per agent writing into that agent's buffer slots, the per-agent action ids are
packed into the env's [N, M, act_width] action array, one env-step launch writes the next joint observation row
straight into the buffers' shared storage, and `insert` scatters rewards / masks into the per-agent buffers."""
import time

import numpy as np
import torch

from .base_runner import Runner


class MPERunner(Runner):
    def __init__(self, config):
        super(MPERunner, self).__init__(config)
        N, M = self.n_rollout_threads, self.num_agents
        f = dict(dtype=torch.float32, device=self.device)
        self._act_env = torch.zeros(N, M, self.envs.act_width, **f)
        self._rew_env = torch.zeros(N, M, 1, **f)
        self._mask_env = torch.ones(N, M, 1, **f)

    def run(self):
        self.warmup()
        start = time.time()
        # --n_rollout_threads is PER RANK (each rank owns that many envs, parallel.env_shard): the job advances
        # world_size x n_rollout_threads envs per step, and --num_env_steps / the lr schedule count those
        n_global = self.n_rollout_threads * self.trainer[0].world_size
        episodes = int(self.num_env_steps) // self.episode_length // n_global
        train_infos = []
        for episode in range(episodes):
            train_infos = self.run_episode(episode, episodes)
            total_num_steps = (episode + 1) * self.episode_length * n_global
            if episode % self.save_interval == 0 or episode == episodes - 1:
                self.save()
            if episode % self.log_interval == 0:
                end = time.time()
                print("\n Scenario {} Algo {} Exp {} updates {}/{} episodes, total num timesteps {}/{}, FPS {}.\n".format(
                    self.all_args.scenario_name, self.algorithm_name, self.experiment_name, episode, episodes,
                    total_num_steps, self.num_env_steps, int(total_num_steps / (end - start))))
                indiv = self.envs._indiv.mean(dim=0).cpu().numpy()
                for agent_id in range(self.num_agents):
                    train_infos[agent_id]["individual_rewards"] = float(indiv[agent_id])
                    train_infos[agent_id]["average_episode_rewards"] = \
                        float(self.buffer[agent_id].rewards.mean().item()) * self.episode_length
                self.log_train(train_infos, total_num_steps)
            if episode % self.eval_interval == 0 and self.use_eval:
                self.eval(total_num_steps)
        return train_infos

    def run_episode(self, episode=0, episodes=1):
        """One iteration (mpe_runner.py:24-41)."""
        if self.use_linear_lr_decay:
            for agent_id in range(self.num_agents):
                self.trainer[agent_id].policy.lr_decay(episode, episodes)
        for step in range(self.episode_length):
            values, actions, action_log_probs, rnn_states, rnn_states_critic, actions_env = self.collect(step)
            obs, rewards, masks = self.envs.step_ids(actions_env, out_obs=self.joint_obs[step + 1],
                                                     out_rewards=self._rew_env, out_masks=self._mask_env)
            data = obs, rewards, masks, None, values, actions, action_log_probs, rnn_states, rnn_states_critic
            self.insert(data)
        self.compute()
        return self.train()

    def warmup(self):
        """mpe_runner.py:81-96: reset; every agent's obs[0] / share_obs[0] is (a slice of) the joint row 0."""
        self.envs.reset(out_obs=self.joint_obs[0])

    @torch.no_grad()
    def collect(self, step):
        """mpe_runner.py:98-140.  Returns per-agent lists (the reference transposes them into [N, M, ...] arrays that
        are ragged for heterogeneous agents) and the packed action-id array the env kernel consumes."""
        values, actions, logps, rs, rsc = [], [], [], [], []
        off = self.envs._cfg.env_id_offset
        for agent_id in range(self.num_agents):
            self.trainer[agent_id].prep_rollout()
            b, pol = self.buffer[agent_id], self.trainer[agent_id].policy
            if b._recurrent:
                # rnn.py:24-28: h = GRU(x, h * mask); the new states go straight into slot step + 1
                pol.actor.forward_into(b.obs[step], b.actions[step], b.action_log_probs[step], row_offset=off,
                                       rnn_in=b.rnn_states[step], masks=b.masks[step], rnn_out=b.rnn_states[step + 1])
                pol.critic.forward_into(b.share_obs[step], b.value_preds[step], rnn_in=b.rnn_states_critic[step],
                                        masks=b.masks[step], rnn_out=b.rnn_states_critic[step + 1])
            else:
                pol.actor.forward_into(b.obs[step], b.actions[step], b.action_log_probs[step], row_offset=off)
                pol.critic.forward_into(b.share_obs[step], b.value_preds[step], group=1)
            self._act_env[:, agent_id, :b.actions.shape[2]].copy_(b.actions[step])
            values.append(b.value_preds[step]); actions.append(b.actions[step]); logps.append(b.action_log_probs[step])
            rs.append(b.rnn_states[step + 1]); rsc.append(b.rnn_states_critic[step + 1])
        return values, actions, logps, rs, rsc, self._act_env

    def insert(self, data):
        """mpe_runner.py:142-176: rewards / masks of agent m go to buffer[m]; the rest already sits in its slot."""
        obs, rewards, masks, infos, values, actions, action_log_probs, rnn_states, rnn_states_critic = data
        for agent_id in range(self.num_agents):
            b = self.buffer[agent_id]
            if b._recurrent:
                # mpe_runner.py:145-146: states of finished episodes are zeroed (masks = 1 - done)
                keep = masks[:, agent_id].unsqueeze(-1)
                b.rnn_states[b.step + 1].mul_(keep)
                b.rnn_states_critic[b.step + 1].mul_(keep)
            b.insert(b.share_obs[b.step + 1], b.obs[b.step + 1], None, None, actions[agent_id],
                     action_log_probs[agent_id], values[agent_id], rewards[:, agent_id], masks[:, agent_id])

    @torch.no_grad()
    def eval(self, total_num_steps):
        """mpe_runner.py:178-230: deterministic rollout on eval_envs."""
        envs = self.eval_envs
        N, M = envs.num_envs, self.num_agents
        envs.reset()
        acts = torch.zeros(N, M, envs.act_width, dtype=torch.float32, device=self.device)
        total = torch.zeros(N, M, 1, dtype=torch.float32, device=self.device)
        views = envs._shape_obs(envs._obs)
        recurrent = bool(self.buffer[0]._recurrent)
        f = dict(dtype=torch.float32, device=self.device)
        rnn = [torch.zeros(N, self.recurrent_N, self.hidden_size, **f) for _ in range(M)]
        nxt = [torch.empty_like(x) for x in rnn]
        masks = torch.ones(N, M, 1, **f)
        for _ in range(self.episode_length):
            for agent_id in range(M):
                h = len(envs.head_sizes[agent_id])
                a = torch.empty(N, h, **f)
                kw = dict(rnn_in=rnn[agent_id], masks=masks[:, agent_id].contiguous(), rnn_out=nxt[agent_id]) \
                    if recurrent else {}
                self.trainer[agent_id].policy.actor.forward_into(
                    views[:, agent_id] if torch.is_tensor(views) else views[agent_id], a, torch.empty_like(a),
                    deterministic=True, **kw)
                acts[:, agent_id, :h].copy_(a)
            _, rew, m = envs.step_ids(acts)
            masks = m.view(N, M, 1).clone()                              # mpe_runner.py:213-216: h * mask at the next step
            rnn, nxt = nxt, rnn
            total += rew
        out = []
        for agent_id in range(M):
            r = float(total[:, agent_id].mean().item())
            print("eval average episode rewards of agent%i: " % agent_id + str(r))
            out.append(r)
        return out
