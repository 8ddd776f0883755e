"""MPERunner mirror (onpolicy/runner/shared/mpe_runner.py:11-183), device-resident.

The reference's loop moves every step's tensors host->device->host (mpe_runner.py:95-123); here `collect` writes
actions / log-probs / values straight into the buffer slots, `envs.step` reads the action slot and writes the next
obs / reward / mask slots, so `insert` only advances the step counter.  One rollout step = 3 kernel launches."""
import time

import numpy as np
import torch

from .base_runner import Runner


class MPERunner(Runner):
    def __init__(self, config):
        super(MPERunner, self).__init__(config)
        self._last_infos = None
        # --use_cuda_graph: the whole iteration (T x {actor, critic, env step}, bootstrap value, GAE, every ppo_update,
        # after_update) is captured once and replayed as ONE graph launch per iteration
        self.use_cuda_graph = bool(getattr(self.all_args, "use_cuda_graph", False))
        self._graph = None
        self._graph_state = 0            # 0: counters still on the host; 1: one eager iteration done; 2: captured
        self._graph_updates = 0

    def run(self):
        self.warmup()
        start = time.time()
        # --n_rollout_threads is PER RANK (each rank owns that many envs, parallel.env_shard): the job advances
        # world_size x n_rollout_threads envs per step, and --num_env_steps / the lr schedule count those
        n_global = self.n_rollout_threads * self.trainer.world_size
        episodes = int(self.num_env_steps) // self.episode_length // n_global
        train_infos = {}
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
                env_infos = {}
                indiv = self.envs._indiv.mean(dim=0).cpu().numpy()          # last step's individual rewards
                for agent_id in range(self.num_agents):
                    env_infos["agent%i/individual_rewards" % agent_id] = [float(indiv[agent_id])]
                train_infos["average_episode_rewards"] = float(self.buffer.rewards.mean().item()) * self.episode_length
                print("average episode rewards is {}".format(train_infos["average_episode_rewards"]))
                self.log_train(train_infos, total_num_steps)
                self.log_env(env_infos, total_num_steps)
            if episode % self.eval_interval == 0 and self.use_eval:
                self.eval(total_num_steps)
        return train_infos

    def run_episode(self, episode=0, episodes=1):
        """One iteration: episode_length x (collect, step, insert), compute, train (mpe_runner.py:24-40)."""
        if self.use_linear_lr_decay:
            self.trainer.policy.lr_decay(episode, episodes)
        if self._graph_eligible():
            return self._run_episode_graph()
        self._rollout()
        self.compute()
        return self.train()

    def _graph_eligible(self):
        """A graph replays fixed launches: nothing on the path may depend on host state that changes between
        iterations.  That rules out host-drawn permutations (num_mini_batch > 1, the naive recurrent generator), a learning
        rate schedule, the NCCL hooks and per-kernel event timing; the sampling counter and the Adam step counts move
        into HBM (R_Actor.use_device_step, FusedAdam.use_device_step)."""
        from ... import ops
        t, b = self.trainer, self.buffer
        return (self.use_cuda_graph and not t._use_naive_recurrent and t.num_mini_batch == 1 and not b.force_permutation
                and not self.use_linear_lr_decay and t.all_reduce is None and not ops.TIMERS.enabled)

    def _iteration_on_stream(self):
        self._rollout()
        self.compute()
        self.trainer.prep_training()
        n = self.trainer.train_enqueue(self.buffer)
        self.buffer.after_update()
        return n

    def _run_episode_graph(self):
        pol = self.trainer.policy
        if self._graph_state == 0:
            # iteration 1 runs eagerly with the device counters: every lazily created workspace and kernel attribute
            # exists before the capture
            pol.actor.use_device_step(); pol.actor_optimizer.use_device_step(); pol.critic_optimizer.use_device_step()
            self._graph_state = 1
            n = self._iteration_on_stream()
            return self.trainer.train_finish(n)
        if self._graph_state == 1:
            torch.cuda.synchronize()
            from ... import ops
            ops.pin_workspaces()                  # scratch buffers the graph records must outlive any later growth
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._graph_updates = self._iteration_on_stream()     # records the launches; host counters advance once
            self._graph, self._graph_state = g, 2
        else:
            # host mirrors of the counters the graph advances on the device
            pol.actor._calls += self.episode_length
            pol.actor_optimizer.step_count += self._graph_updates
            pol.critic_optimizer.step_count += self._graph_updates
        self._graph.replay()
        return self.trainer.train_finish(self._graph_updates)

    def _rollout(self, given_actions=None, reset_states=None):
        """given_actions [T, N, M, heads] (device float32 ids) / reset_states {step: state dict in the reference's per-env
        layout}: replay a recorded rollout -- the recorded actions are evaluated instead of sampled and the recorded
        reset states are installed where the envs auto-reset (tests/test_gpu_runner.py: the reference's own runs)."""
        for step in range(self.episode_length):
            ga = None if given_actions is None else given_actions[step]
            values, actions, action_log_probs, rnn_states, rnn_states_critic, actions_env = self.collect(step, ga)
            rs = None if reset_states is None else reset_states.get(step)
            obs, rewards, masks = self.envs.step_ids(actions_env, out_obs=self.buffer.obs[step + 1],
                                                     out_rewards=self.buffer.rewards[step],
                                                     out_masks=self.buffer.masks[step + 1], reset_state=rs)
            data = obs, rewards, masks, None, values, actions, action_log_probs, rnn_states, rnn_states_critic
            self.insert(data)

    def warmup(self):
        """mpe_runner.py:81-93: reset; obs[0] (and through the view share_obs[0]) receive the first observation."""
        b = self.buffer
        if b.share_obs_unique() is None and self.use_centralized_V:
            raise NotImplementedError("shared runner needs homogeneous agents (use the separated runner)")
        self.envs.reset(out_obs=b.obs[0])

    @torch.no_grad()
    def collect(self, step, given_actions=None):
        """mpe_runner.py:95-123 with outputs written in place.  `actions_env` is the action-id slot itself: the env
        kernel decodes ids, so the one-hot conversion (:110-121) disappears."""
        b = self.buffer
        N, M = b.n_rollout_threads, b.num_agents
        pol = self.trainer.policy
        self.trainer.prep_rollout()
        if b._recurrent:
            # rnn.py:24-28: h = GRU(x, h * mask); the new states go straight into slot step + 1
            pol.actor.forward_into(b.obs[step].view(N * M, -1), b.actions[step], b.action_log_probs[step],
                                   row_offset=self.envs._cfg.env_id_offset * M, rnn_in=b.rnn_states[step],
                                   masks=b.masks[step], rnn_out=b.rnn_states[step + 1], given_actions=given_actions)
            uniq = b.share_obs_unique()
            if self.use_centralized_V and uniq is not None:
                pol.critic.forward_into(uniq[step], b.value_preds[step], rnn_in=b.rnn_states_critic[step],
                                        masks=b.masks[step], rnn_out=b.rnn_states_critic[step + 1], share_rows_div=M)
            else:
                x = b.share_obs[step] if self.use_centralized_V else b.obs[step]
                pol.critic.forward_into(x.reshape(N * M, -1), b.value_preds[step], rnn_in=b.rnn_states_critic[step],
                                        masks=b.masks[step], rnn_out=b.rnn_states_critic[step + 1])
            return (b.value_preds[step], b.actions[step], b.action_log_probs[step], b.rnn_states[step + 1],
                    b.rnn_states_critic[step + 1], b.actions[step])
        pol.actor.forward_into(b.obs[step].view(N * M, -1), b.actions[step], b.action_log_probs[step],
                               row_offset=self.envs._cfg.env_id_offset * M, given_actions=given_actions)
        if self.use_centralized_V:
            pol.critic.forward_into(b.share_obs_unique()[step], b.value_preds[step], group=M)
        else:
            pol.critic.forward_into(b.obs[step].view(N * M, -1), b.value_preds[step], group=1)
        return (b.value_preds[step], b.actions[step], b.action_log_probs[step], b.rnn_states[step + 1],
                b.rnn_states_critic[step + 1], b.actions[step])

    def insert(self, data):
        """mpe_runner.py:125-139: everything already sits in its slot (masks = 1 - done written by the env kernel;
        MLP policies carry no rnn state), so this validates the slot identity and advances `buffer.step`."""
        obs, rewards, masks, infos, values, actions, action_log_probs, rnn_states, rnn_states_critic = data
        b = self.buffer
        if b._recurrent:
            # mpe_runner.py:128-129: states of finished episodes are zeroed (masks = 1 - done, written by the env kernel)
            s1 = b.step + 1
            b.rnn_states[s1].mul_(b.masks[s1].unsqueeze(-1))
            b.rnn_states_critic[s1].mul_(b.masks[s1].unsqueeze(-1))
        b.insert(b.share_obs[b.step + 1], obs, rnn_states, rnn_states_critic, actions, action_log_probs, values,
                 rewards, masks)

    @torch.no_grad()
    def eval(self, total_num_steps):
        """mpe_runner.py:141-183: deterministic rollout on eval_envs."""
        envs = self.eval_envs
        N, M = envs.num_envs, self.num_agents
        obs = envs.reset()
        acts = torch.empty(N, M, envs.act_width, dtype=torch.float32, device=self.device)
        lps = torch.empty_like(acts)
        total = torch.zeros(N, M, 1, dtype=torch.float32, device=self.device)
        recurrent = bool(self.buffer._recurrent)
        rnn = torch.zeros(N, M, self.recurrent_N, self.hidden_size, dtype=torch.float32, device=self.device)
        nxt = torch.empty_like(rnn)
        masks = torch.ones(N, M, 1, dtype=torch.float32, device=self.device)
        for _ in range(self.episode_length):
            kw = dict(rnn_in=rnn, masks=masks, rnn_out=nxt) if recurrent else {}
            self.trainer.policy.actor.forward_into(envs._obs.view(N * M, -1), acts, lps, deterministic=True, **kw)
            _, rew, m = envs.step_ids(acts)
            masks = m.view(N, M, 1).clone()                              # mpe_runner.py:166-169: h * mask at the next step
            rnn, nxt = nxt, rnn
            total += rew
        r = float(total.mean().item())
        print("eval average episode rewards of agent: " + str(r))
        self.log_env({"eval_average_episode_rewards": [r]}, total_num_steps)
        return r

    @torch.no_grad()
    def render(self):
        """mpe_runner.py:185-248: deterministic episodes on `self.envs`, one frame of env 0 after the reset and after
        every step (`envs.render('rgb_array')[0][0]`), rasterised on the device (csrc/render.cu) and kept in HBM until
        the episode ends.  With `save_gifs` the frames go to `<run_dir>/gifs/render.gif` (PIL on the host; the
        reference uses imageio, which is not installed).  `human` mode needs a window: unavailable on this box, so
        frames are always produced.  Returns (frames uint8 [F, 700, 700, 3] on the host, average episode reward)."""
        a = self.all_args
        envs = self.envs
        N, M = envs.num_envs, self.num_agents
        pol = self.trainer.policy
        recurrent = bool(self.buffer._recurrent)
        frames, ep_rewards = [], []
        acts = torch.empty(N, M, envs.act_width, dtype=torch.float32, device=self.device)
        lps = torch.empty_like(acts)
        for _ in range(int(getattr(a, "render_episodes", 1))):
            envs.reset()
            frames.append(envs.render("rgb_array", env_ids=[0], as_numpy=False)[0, 0])
            rnn = torch.zeros(N, M, self.recurrent_N, self.hidden_size, dtype=torch.float32, device=self.device)
            rnn_next = torch.zeros_like(rnn)
            masks = torch.ones(N, M, 1, dtype=torch.float32, device=self.device)
            total = torch.zeros(N, M, 1, dtype=torch.float32, device=self.device)
            for _step in range(self.episode_length):
                if recurrent:
                    pol.actor.forward_into(envs._obs.view(N * M, -1), acts, lps, deterministic=True, rnn_in=rnn,
                                           masks=masks, rnn_out=rnn_next)
                    rnn, rnn_next = rnn_next, rnn
                else:
                    pol.actor.forward_into(envs._obs.view(N * M, -1), acts, lps, deterministic=True)
                _, rew, masks = envs.step_ids(acts)
                masks = masks.clone()
                if recurrent:
                    rnn.mul_(masks.unsqueeze(-1))                              # mpe_runner.py:232
                total += rew
                frames.append(envs.render("rgb_array", env_ids=[0], as_numpy=False)[0, 0])
            ep_rewards.append(total)
        avg = float(torch.stack(ep_rewards).mean().item())
        print("average episode rewards is: " + str(avg))
        out = torch.stack(frames).cpu().numpy()
        if getattr(a, "save_gifs", False) and self.run_dir is not None:
            import os
            from PIL import Image
            gif_dir = os.path.join(str(self.run_dir), "gifs")
            os.makedirs(gif_dir, exist_ok=True)
            imgs = [Image.fromarray(f) for f in out]
            imgs[0].save(os.path.join(gif_dir, "render.gif"), save_all=True, append_images=imgs[1:],
                         duration=int(1000 * float(getattr(a, "ifi", 0.1))), loop=0)
        return out, avg
