"""R_MAPPO trainer mirror (onpolicy/algorithms/r_mappo/r_mappo.py:8-227): same constructor, `train`,
`ppo_update`, `cal_value_loss`, `prep_training`, `prep_rollout`, `value_normalizer`.

One `ppo_update` is: fused actor forward+loss+backward kernel, fused critic kernel, a fixed-order reduction of
the per-CTA partial gradients, (multi-GPU: one NCCL all-reduce of the two flat gradient buffers and the loss
sums), and one kernel per network that finishes the chain rule, clips the global norm and applies Adam.
Nothing synchronises with the host until `train` returns its six scalars."""
import os

import numpy as np
import torch

from ... import ops
from ...utils.valuenorm import ValueNorm
from .algorithm.r_actor_critic import check


class R_MAPPO:
    def __init__(self, args, policy, device=torch.device("cuda")):
        self.device = torch.device(device)
        self.tpdv = dict(dtype=torch.float32, device=self.device)
        self.policy = policy
        self.clip_param = args.clip_param
        self.ppo_epoch = args.ppo_epoch
        self.num_mini_batch = args.num_mini_batch
        self.data_chunk_length = args.data_chunk_length
        self.value_loss_coef = args.value_loss_coef
        self.entropy_coef = args.entropy_coef
        self.max_grad_norm = args.max_grad_norm
        self.huber_delta = args.huber_delta
        self._use_recurrent_policy = args.use_recurrent_policy
        self._use_naive_recurrent = args.use_naive_recurrent_policy
        self._use_max_grad_norm = args.use_max_grad_norm
        self._use_clipped_value_loss = args.use_clipped_value_loss
        self._use_huber_loss = args.use_huber_loss
        self._use_popart = args.use_popart
        self._use_valuenorm = args.use_valuenorm
        self._use_value_active_masks = args.use_value_active_masks
        self._use_policy_active_masks = args.use_policy_active_masks
        assert (self._use_popart and self._use_valuenorm) == False, \
            ("self._use_popart and self._use_valuenorm can not be set True simultaneously")
        if self._use_popart:
            self.value_normalizer = self.policy.critic.v_out            # r_mappo.py:45-46
        else:
            self.value_normalizer = ValueNorm(1, device=self.device) if self._use_valuenorm else None
        # multi-GPU hook: callable(tensor) summing it in place over ranks (torch.distributed.all_reduce); None on 1 GPU
        self.all_reduce = None
        self.pack_wide_inputs = os.environ.get("MAPPO_PACK_WIDE", "1") != "0"   # A/B switch for the packed first layer
        self._ws = {}
        self._seed_scratch = {}
        self._info = torch.zeros(8, **self.tpdv)       # actor {loss, gnorm, entropy, ratio}, critic {loss, gnorm, -, -}
        self._info_total = torch.zeros(8, **self.tpdv)

    def set_all_reduce(self, fn, world_size):
        """Data-parallel update: `fn(tensor)` sums a CUDA tensor in place over the ranks (NCCL all-reduce)."""
        self.all_reduce = fn
        self._ws = {}                       # the reduce buffers are laid out differently with a collective attached
        self.world_size = int(world_size)
        if self.value_normalizer is not None:
            self.value_normalizer.all_reduce = fn

    def _loss_config(self):
        return ops.LossConfig(self.clip_param, self.entropy_coef, self.value_loss_coef, self.huber_delta,
                              int(self._use_huber_loss), int(self._use_clipped_value_loss),
                              int(self._use_valuenorm or self._use_popart),
                              int(self._use_policy_active_masks), int(self._use_value_active_masks), 0)

    def _workspace(self, rows_a, rows_c):
        key = (rows_a, rows_c)
        ws = self._ws.get(key)
        if ws is None:
            Pa, Pc = self.policy.actor.flat.numel(), self.policy.critic.flat.numel()
            nba = ops.update_num_blocks(self.policy.actor.cfg, rows_a)
            nbc = ops.update_num_blocks(self.policy.critic.cfg, rows_c)
            dev = self.device
            ws = dict(nba=nba, nbc=nbc,
                      gpa=torch.empty(nba, Pa, dtype=torch.float32, device=dev),
                      gpc=torch.empty(nbc, Pc, dtype=torch.float32, device=dev),
                      lpa=torch.empty(nba, 4, dtype=torch.float64, device=dev),
                      lpc=torch.empty(nbc, 4, dtype=torch.float64, device=dev),
                      grad=torch.empty(Pa + Pc, dtype=torch.float32, device=dev))
            if self.all_reduce is not None:
                # data parallel: [actor grad sums | critic grad sums | 2 x 4 loss sums] in ONE float64 buffer -> one
                # all-reduce per ppo_update (the reduce kernel writes it, mappo_adam_step_f64 reads it)
                ws["red"] = torch.empty(Pa + Pc + 8, dtype=torch.float64, device=dev)
                ws["gsum"], ws["lsum"] = ws["red"][:Pa + Pc], ws["red"][Pa + Pc:]
            else:
                ws.update(gsum=torch.empty(Pa + Pc, dtype=torch.float32, device=dev),
                          lsum=torch.empty(8, dtype=torch.float64, device=dev))
            if ops.GRAPH_PINNED:        # a captured graph replays the pointers of the workspace it was recorded with
                self._ws[key] = ws
            else:
                self._ws = {key: ws}    # keep one (most recent) workspace
        return ws

    # ------------------------------------------------------------------ core: one minibatch on device tensors
    def _update(self, obs, share_obs, group, actions, old_logp, adv, value_preds, returns, active, avail,
                rows_total_a, index=None, share_rows_div=1, update_actor=True, vn_batch=None, vn_moments=None,
                packed=None):
        """obs [rows_a, D] (strided rows ok); share_obs [rows_c, S]; the per-sample arrays are flat and, with
        `index`, addressed through it.  rows_total_a: number of actor samples in this minibatch."""
        pol = self.policy
        rows_a = rows_total_a
        rows_c = rows_a if index is not None else share_obs.shape[0]
        ws = self._workspace(rows_a, rows_c)
        Pa = pol.actor.flat.numel()
        loss = self._loss_config()
        gsa, gsc = ws["gsum"][:Pa], ws["gsum"][Pa:]
        lsa, lsc = ws["lsum"][:4], ws["lsum"][4:]
        ops.actor_update(pol.actor.cfg, loss, pol.actor.flat, obs, actions, old_logp, adv, active, avail,
                         ws["gpa"], ws["lpa"], index=index)
        ops.grad_reduce(ws["gpa"], ws["lpa"], ws["nba"], gsa, lsa)
        # ValueNorm.update(return_batch) happens inside the value loss, before normalize (r_mappo.py:65-67)
        old_stats = None
        if self.value_normalizer is not None:
            kw = dict(defer_rescale=True) if self._use_popart else {}
            old_stats = self.value_normalizer.update(returns if vn_batch is None else vn_batch, moments=vn_moments, **kw)
        ops.critic_update(pol.critic.cfg, loss, pol.critic.flat, share_obs, value_preds, returns, active,
                          None if self.value_normalizer is None else self.value_normalizer.stats,
                          ws["gpc"], ws["lpc"], group=group, index=index, share_rows_div=share_rows_div, packed=packed)

        ops.grad_reduce(ws["gpc"], ws["lpc"], ws["nbc"], gsc, lsc)
        rows_global = rows_a
        if self.all_reduce is not None:
            self.all_reduce(ws["red"])
            rows_global = rows_a * self.world_size
        self._info.zero_()
        aopt, copt = pol.actor_optimizer, pol.critic_optimizer
        ops.adam_step(pol.actor.cfg, aopt.config(self.max_grad_norm, self._use_max_grad_norm), pol.actor.flat,
                      aopt.exp_avg, aopt.exp_avg_sq, gsa, lsa, 3, rows_global, ws["grad"][:Pa], self._info[:4],
                      apply_update=update_actor, step_dev=aopt.step_dev)
        if update_actor:
            aopt.step_count += 1
        ops.adam_step(pol.critic.cfg, copt.config(self.max_grad_norm, self._use_max_grad_norm), pol.critic.flat,
                      copt.exp_avg, copt.exp_avg_sq, gsc, lsc, 1, rows_global, ws["grad"][Pa:], self._info[4:],
                      apply_update=True, step_dev=copt.step_dev)
        copt.step_count += 1
        if self._use_popart:
            # popart.py:69-70: update() rescales the head.  Done last: the loss and its backward pass (including the
            # chain rule through the LayerNorm affine in front of the head, inside mappo_adam_step) belong to the head
            # the forward pass used
            self.value_normalizer.rescale(old_stats)
        self._info_total += self._info
        return ws

    def _seeded_states(self, v, first_rows, seed_rows):
        """The kernels read the initial state of slot c at the sample row of its step 0 (`first_rows[c]`).  Where the
        buffer keeps it elsewhere (SeparatedReplayBuffer.recurrent_rows), copy it there in a scratch array of the same
        shape; only those Nc rows of the scratch are ever read."""
        out = dict(v)
        for k in ("rnn", "rnn_critic"):
            sc = self._seed_scratch.get(k)
            if sc is None or sc.shape != v[k].shape:
                if sc is not None and ops.GRAPH_PINNED:
                    ops._RETIRED.append(sc)
                sc = self._seed_scratch[k] = torch.empty_like(v[k])
            sc.index_copy_(0, first_rows, v[k].index_select(0, seed_rows))
            out[k] = sc
        return out

    def _update_rnn(self, v, idx, n_chunks, L, update_actor=True):
        """One recurrent minibatch: idx int64 [L * n_chunks] (step-major) flat sample rows of the chunks
        (shared_buffer.py:385-494 / :288-383); back-propagation through time inside the kernels (csrc/mlp_gru.cu)."""
        pol = self.policy
        rows = n_chunks * L
        ws = self._workspace(rows, rows)
        Pa = pol.actor.flat.numel()
        loss = self._loss_config()
        gsa, gsc = ws["gsum"][:Pa], ws["gsum"][Pa:]
        lsa, lsc = ws["lsum"][:4], ws["lsum"][4:]
        ops.actor_update_rnn(pol.actor.cfg, loss, pol.actor.flat, v["obs"], n_chunks, L, idx, v["rnn"], v["masks"],
                             v["actions"], v["logp"], v["adv"], v["active"], v["avail"], ws["gpa"], ws["lpa"])
        ops.grad_reduce(ws["gpa"], ws["lpa"], ws["nba"], gsa, lsa)
        old_stats = None
        if self.value_normalizer is not None:
            kw = dict(defer_rescale=True) if self._use_popart else {}
            old_stats = self.value_normalizer.update(v["returns"].index_select(0, idx), **kw)
        ops.critic_update_rnn(pol.critic.cfg, loss, pol.critic.flat, v["share"], n_chunks, L, idx, v["rnn_critic"],
                              v["masks"], v["value_preds"], v["returns"], v["active"],
                              None if self.value_normalizer is None else self.value_normalizer.stats,
                              ws["gpc"], ws["lpc"], share_rows_div=v["M"] if v["share_is_unique"] else 1)
        ops.grad_reduce(ws["gpc"], ws["lpc"], ws["nbc"], gsc, lsc)
        rows_global = rows
        if self.all_reduce is not None:
            self.all_reduce(ws["red"])
            rows_global = rows * self.world_size
        self._info.zero_()
        aopt, copt = pol.actor_optimizer, pol.critic_optimizer
        ops.adam_step(pol.actor.cfg, aopt.config(self.max_grad_norm, self._use_max_grad_norm), pol.actor.flat,
                      aopt.exp_avg, aopt.exp_avg_sq, gsa, lsa, 3, rows_global, ws["grad"][:Pa], self._info[:4],
                      apply_update=update_actor, step_dev=aopt.step_dev)
        if update_actor:
            aopt.step_count += 1
        ops.adam_step(pol.critic.cfg, copt.config(self.max_grad_norm, self._use_max_grad_norm), pol.critic.flat,
                      copt.exp_avg, copt.exp_avg_sq, gsc, lsc, 1, rows_global, ws["grad"][Pa:], self._info[4:],
                      apply_update=True, step_dev=copt.step_dev)
        copt.step_count += 1
        if self._use_popart:
            # popart.py:69-70: update() rescales the head.  Done last: the loss and its backward pass (including the
            # chain rule through the LayerNorm affine in front of the head, inside mappo_adam_step) belong to the head
            # the forward pass used
            self.value_normalizer.rescale(old_stats)
        self._info_total += self._info
        return ws

    world_size = 1

    # ------------------------------------------------------------------ reference API
    def cal_value_loss(self, values, value_preds_batch, return_batch, active_masks_batch):
        """API-compatible helper (r_mappo.py:52-89) on device tensors.  The training path does not call it: the
        value loss lives inside the fused critic kernel."""
        values, value_preds_batch = check(values, self.device), check(value_preds_batch, self.device)
        return_batch, active_masks_batch = check(return_batch, self.device), check(active_masks_batch, self.device)
        clipped = value_preds_batch + (values - value_preds_batch).clamp(-self.clip_param, self.clip_param)
        if self._use_popart or self._use_valuenorm:                  # r_mappo.py:64-67
            self.value_normalizer.update(return_batch)
            target = self.value_normalizer.normalize(return_batch).view_as(return_batch)
        else:
            target = return_batch
        e_c, e_o = target - clipped, target - values

        def lossf(e):
            if self._use_huber_loss:
                d = self.huber_delta
                return torch.where(e.abs() <= d, e ** 2 / 2, d * (e.abs() - d / 2))
            return e ** 2 / 2
        l = torch.max(lossf(e_o), lossf(e_c)) if self._use_clipped_value_loss else lossf(e_o)
        if self._use_value_active_masks:
            return (l * active_masks_batch).sum() / active_masks_batch.sum()
        return l.mean()

    def ppo_update(self, sample, update_actor=True):
        """sample: the reference's 12-tuple (shared_buffer.py:284-286), numpy or torch.  Returns
        (value_loss, critic_grad_norm, policy_loss, dist_entropy, actor_grad_norm, imp_weights) as 0-dim device
        tensors; `imp_weights` is the MEAN importance weight (the reference returns the per-sample tensor and
        only ever takes its mean, r_mappo.py:212)."""
        if self._use_recurrent_policy or self._use_naive_recurrent:
            return self._ppo_update_rnn(sample, update_actor)
        (share_obs, obs, _rs, _rsc, actions, value_preds, returns, masks, active, old_logp, adv, avail) = sample
        dev = self.device
        c = lambda x: None if x is None else check(x, dev).contiguous()
        obs, share_obs = c(obs), c(share_obs)
        B = obs.shape[0]
        self._update(obs, share_obs, 1, c(actions), c(old_logp), c(adv).reshape(-1), c(value_preds).reshape(-1),
                     c(returns).reshape(-1), c(active).reshape(-1),
                     c(avail) if self.policy.actor.discrete else None, B, update_actor=update_actor)
        i = self._info.clone()
        return i[4], i[5], i[0], i[2], i[1], i[3]

    def _ppo_update_rnn(self, sample, update_actor=True):
        """Recurrent sample of the generators: arrays [L*N, ...] step-major, rnn states [N, 1, H]."""
        (share_obs, obs, rs, rsc, actions, value_preds, returns, masks, active, old_logp, adv, avail) = sample
        dev = self.device
        c = lambda x: None if x is None else check(x, dev).to(torch.float32).contiguous()
        obs, share_obs = c(obs), c(share_obs)
        B = obs.shape[0]
        H = 64 * int(self.policy.actor._recurrent_N)         # a sample's state: [recurrent_N, 64]
        N = c(rs).reshape(-1, H).shape[0]
        L = B // N
        # rnn states exist per chunk only: place them at the rows of step 0 (the only rows the kernels read them at)
        def spread(x):
            full = torch.zeros(B, H, dtype=torch.float32, device=dev)
            full[:N] = c(x).reshape(N, H)
            return full
        v = dict(obs=obs, share=share_obs, M=1, share_is_unique=False, rnn=spread(rs), rnn_critic=spread(rsc),
                 masks=c(masks).reshape(-1), actions=c(actions), logp=c(old_logp), adv=c(adv).reshape(-1),
                 active=c(active).reshape(-1), avail=c(avail) if self.policy.actor.discrete else None,
                 value_preds=c(value_preds).reshape(-1), returns=c(returns).reshape(-1))
        idx = torch.arange(B, device=dev, dtype=torch.int64)
        self._update_rnn(v, idx, N, L, update_actor=update_actor)
        i = self._info.clone()
        return i[4], i[5], i[0], i[2], i[1], i[3]

    def train(self, buffer, update_actor=True):
        """r_mappo.py:166-219 on the device-resident buffer."""
        return self.train_finish(self.train_enqueue(buffer, update_actor))

    def train_finish(self, n_updates):
        """The only host synchronisation of train(): the six averaged scalars (r_mappo.py:207-219)."""
        info = (self._info_total / float(n_updates)).cpu().numpy()
        return {"value_loss": float(info[4]), "policy_loss": float(info[0]), "dist_entropy": float(info[2]),
                "actor_grad_norm": float(info[1]), "critic_grad_norm": float(info[5]), "ratio": float(info[3])}

    def train_enqueue(self, buffer, update_actor=True):
        """Everything of train() that runs on the stream (capturable in a CUDA graph when no host permutation is
        drawn); returns the number of ppo_update calls whose scalars were accumulated in `_info_total`."""
        vn = None if self.value_normalizer is None else self.value_normalizer.stats
        T = buffer.episode_length
        raw, s3 = ops.advantages(buffer.returns[:-1], buffer.value_preds[:-1], buffer.active_masks[:-1], vn,
                                 all_reduce=self.all_reduce)
        advantages = ops.adv_normalize(raw, s3, out=buffer._adv_scratch())
        self._info_total.zero_()
        v = buffer.flat_views()
        B, M = v["B"], v["M"]
        obs_flat, sh, acts, logp, avail = v["obs"], v["share"], v["actions"], v["logp"], v["avail"]
        vp, ret, act_m = v["value_preds"], v["returns"], v["active"]
        adv = advantages.view(B)
        rep = M if v["share_is_unique"] else 1          # samples served by one share_obs row
        n_updates = 0
        v["adv"] = adv
        # num_mini_batch == 1: every epoch's return batch is the whole buffer, so its moments (and, multi-GPU, their
        # all-reduce) are computed once per iteration; ValueNorm.update still runs every epoch
        vn_once = None
        whole_batch = not (self._use_recurrent_policy or self._use_naive_recurrent) and \
            self.num_mini_batch == 1 and not buffer.force_permutation
        if whole_batch and self.value_normalizer is not None:
            vn_once = self.value_normalizer.batch_moments(ret)
        # wide centralised critic: the share_obs rows of the whole batch become packed operand tiles once per iteration
        packed = None
        if whole_batch and self.pack_wide_inputs and self.policy.critic.cfg.in_dim > 64:
            packed = ops.critic_pack_inputs(self.policy.critic.cfg, sh)
        for _ in range(self.ppo_epoch):
            if self._use_recurrent_policy or self._use_naive_recurrent:
                if self._use_recurrent_policy:
                    batches = buffer.recurrent_rows(self.num_mini_batch, self.data_chunk_length)
                else:
                    batches = buffer.naive_recurrent_rows(self.num_mini_batch)
                for batch in batches:
                    rows_idx, n_chunks, L = batch[:3]
                    seed = batch[3] if len(batch) > 3 else None
                    vb = v if seed is None else self._seeded_states(v, rows_idx[:n_chunks], seed)
                    self._update_rnn(vb, rows_idx, n_chunks, L, update_actor=update_actor)
                    n_updates += 1
            elif self.num_mini_batch == 1 and not buffer.force_permutation:
                # the loss is a mean over the whole batch: the permutation only reorders the sum
                self._update(obs_flat, sh, rep, acts, logp, adv, vp, ret, act_m, avail, B, update_actor=update_actor,
                             vn_moments=vn_once, packed=packed)
                n_updates += 1
            else:
                mbs = B // self.num_mini_batch
                rand = torch.randperm(B)                 # CPU generator, exactly as shared_buffer.py:246
                rand_dev = rand.to(self.device, non_blocking=True)
                for k in range(self.num_mini_batch):
                    idx = rand_dev[k * mbs:(k + 1) * mbs].contiguous()
                    vn_batch = ret.index_select(0, idx) if self.value_normalizer is not None else None
                    self._update(obs_flat, sh, 1, acts, logp, adv, vp, ret, act_m, avail, mbs, index=idx,
                                 share_rows_div=rep, update_actor=update_actor, vn_batch=vn_batch)
                    n_updates += 1
        return n_updates

    def prep_training(self):
        pass

    def prep_rollout(self):
        pass
