"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE
(/root/reference/onpolicy, imported through tests/golden/ref_shims.py).

Run in the build container (the GPU box has no /root/reference):
    python tests/golden/make_golden.py [env] [gae] [gen] [ppo] [vnorm] [runner]
`popart` is the one part recorded ON THE GPU BOX (the reference's PopArt only runs on cuda, popart.py:21-27):
    MAPPO_REF_ROOT=baseline/_ref python tests/golden/make_golden.py popart      (writes gpurun_out/ppo_popart*_cuda.npz)
The reference ships no tests and no golden vectors, so these fixtures -- outputs of
the reference itself on seeded inputs -- are what pins the oracle (tests/test_oracle_*.py, CPU) and,
through it and directly, the CUDA path (tests/test_gpu_*.py).
Recorded with: python 3.12.3, torch 2.11.0 (CPU), numpy 2.3.5.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ref_shims  # noqa: E402

# MAPPO_REF_ROOT: where the unmodified reference lives -- /root/reference in the build container, baseline/_ref (the
# git-ignored copy tools/install_ref.sh makes; it travels with the snapshot) on the GPU box
REF_ROOT = os.environ.get("MAPPO_REF_ROOT", ref_shims.REFERENCE_ROOT)
ref_shims.install(REF_ROOT)


def ref_args(scenario, M, L, algo, extra=""):
    argv = ("--env_name MPE --algorithm_name %s --scenario_name %s --num_agents %d --num_landmarks %d "
            "--episode_length 25 --use_wandb --n_rollout_threads 4 --seed 1 %s" % (algo, scenario, M, L, extra)).split()
    return ref_shims.reference_args(argv)


# ------------------------------------------------------------------ env fixtures
def world_state(env):
    w = env.world
    p_pos = np.array([a.state.p_pos for a in w.agents])
    p_vel = np.array([a.state.p_vel for a in w.agents])
    lm = np.array([l.state.p_pos for l in w.landmarks])
    comm = np.array([a.state.c for a in w.agents])
    def goal_of(a):
        for name in ("goal_b", "goal_a"):      # goal_b: speaker_listener / reference; goal_a (a landmark): adversary / push
            g = getattr(a, name, None)
            if g is not None and g in w.landmarks:
                return w.landmarks.index(g)
        return -1
    goal = np.array([goal_of(a) for a in w.agents])
    if hasattr(w.agents[0], "key"):
        # simple_crypto.py:63-69: slot 0 = the goal landmark (goal_a of every agent), slot 1 = the landmark whose
        # colour is the speaker's key, slot 2 unused
        key = w.agents[2].key
        goal = np.array([goal[0], -1 if key is None else int(np.argmax(key)), -1])
    return p_pos, p_vel, lm, comm, goal


def one_hot_action(space, ids):
    name = space.__class__.__name__
    if name == "Discrete":
        return np.eye(space.n)[ids[0]]
    assert name == "MultiDiscrete"
    sizes = space.high - space.low + 1
    return np.concatenate([np.eye(int(s))[i] for s, i in zip(sizes, ids)])


def make_env_fixture(scenario, M, L, N=6, steps=56, seed=11):
    from onpolicy.envs.mpe.MPE_env import MPEEnv
    args = ref_args(scenario, M, L, "mappo")
    if scenario == "simple_tag":
        # simple_tag.py:11-12 reads two flags that train_mpe.py's parser (train_mpe.py:52-61) does not define
        args.num_adversaries, args.num_good_agents = M - 1, 1
    if scenario == "simple_world_comm":                                 # simple_world_comm.py:11-12, same two flags
        args.num_adversaries, args.num_good_agents = M - 2, 2
    rng = np.random.RandomState(seed)
    # one reference env per slot; they share the process-global np.random like DummyVecEnv
    np.random.seed(seed)
    envs = [MPEEnv(args) for _ in range(N)]
    heads = []
    for sp in envs[0].action_space:
        heads.append([sp.n] if sp.__class__.__name__ == "Discrete" else list((sp.high - sp.low + 1).astype(int)))
    H = max(len(h) for h in heads)
    obs0 = [e.reset() for e in envs]
    if scenario in ("simple_push", "simple_tag", "simple_world_comm"):
        # crowd the upper half of the envs around a landmark: agent-agent and agent-landmark contacts, tag rewards
        # (simple_world_comm: alternately the obstacle and the first forest, so the forest visibility masks switch)
        for n in range(N // 2, N):
            lm = envs[n].world.landmarks
            c = lm[len(lm) - 2 if (scenario == "simple_world_comm" and n % 2) else 0].state.p_pos
            for a in envs[n].world.agents:
                a.state.p_pos = c + rng.uniform(-0.3, 0.3, 2)
            obs0[n] = [envs[n]._get_obs(a) for a in envs[n].agents]
    if scenario == "simple_spread":
        # crowd the agents of the upper half of the envs so the soft-contact force (core.py:290-321)
        # and the collision penalty (simple_spread.py:80-83) are exercised from the first step
        for n in range(N // 2, N):
            for a in envs[n].world.agents:
                a.state.p_pos = rng.uniform(-0.3, 0.3, 2)
            obs0[n] = [envs[n]._get_obs(a) for a in envs[n].agents]
    st = [world_state(e) for e in envs]
    out = dict(M=M, L=L, N=N, steps=steps, episode_length=25,
               init_p_pos=np.array([s[0] for s in st]), init_lm_pos=np.array([s[2] for s in st]),
               init_goal=np.array([s[4] for s in st]))
    for i in range(M):
        out["init_obs_%d" % i] = np.array([o[i] for o in obs0])
    acts = np.zeros((steps, N, M, H), dtype=np.int64)
    obs_rec = [[] for _ in range(M)]
    rew_rec, done_rec, ind_rec, pp_rec, pv_rec, cm_rec = [], [], [], [], [], []
    reset_steps, r_pp, r_lm, r_goal = [], [], [], []
    for s in range(steps):
        obs_s, rew_s, done_s, ind_s = [], [], [], []
        did_reset = False
        pre_reset = []
        for n, e in enumerate(envs):
            a_ids = [[rng.randint(0, k) for k in heads[i]] for i in range(M)]
            for i in range(M):
                acts[s, n, i, :len(a_ids[i])] = a_ids[i]
            a_env = [one_hot_action(e.action_space[i], a_ids[i]) for i in range(M)]
            ob, rew, done, info = e.step(a_env)
            pre_reset.append(world_state(e))
            if np.all(done):  # env_wrappers.py:145-154
                ob = e.reset()
                did_reset = True
            obs_s.append(ob); rew_s.append(rew); done_s.append(done)
            ind_s.append([inf["individual_reward"] for inf in info])
        st = [world_state(e) for e in envs]
        if did_reset:
            reset_steps.append(s)
            r_pp.append([x[0] for x in st]); r_lm.append([x[2] for x in st]); r_goal.append([x[4] for x in st])
        for i in range(M):
            obs_rec[i].append(np.array([o[i] for o in obs_s]))
        rew_rec.append(np.array(rew_s, dtype=np.float64)); done_rec.append(np.array(done_s))
        ind_rec.append(np.array(ind_s, dtype=np.float64))
        # state right after World.step (before any auto-reset): what the physics produced
        pp_rec.append(np.array([x[0] for x in pre_reset])); pv_rec.append(np.array([x[1] for x in pre_reset]))
        cm_rec.append(np.array([x[3] for x in pre_reset]))
    out.update(actions=acts, rewards=np.array(rew_rec), dones=np.array(done_rec), indiv=np.array(ind_rec),
               post_p_pos=np.array(pp_rec), post_p_vel=np.array(pv_rec), post_comm=np.array(cm_rec),
               reset_steps=np.array(reset_steps), reset_p_pos=np.array(r_pp), reset_lm_pos=np.array(r_lm),
               reset_goal=np.array(r_goal))
    for i in range(M):
        out["obs_%d" % i] = np.array(obs_rec[i])
    return out


ENV_FIXTURES = (("simple_spread", 3, 3, "spread_m3"), ("simple_spread", 10, 3, "spread_m10"),
                ("simple_speaker_listener", 2, 3, "speaker_listener"), ("simple_reference", 2, 3, "reference"),
                ("simple_adversary", 3, 2, "adversary"), ("simple_push", 2, 2, "push"), ("simple_tag", 4, 2, "tag"),
                ("simple_crypto", 3, 2, "crypto"), ("simple_world_comm", 4, 1, "world_comm"))


def gen_env(only=None):
    for scen, M, L, tag in ENV_FIXTURES:
        if only and tag not in only:
            continue
        fx = make_env_fixture(scen, M, L, N=4 if M == 10 else 6)
        path = os.path.join(HERE, "env_%s.npz" % tag)
        np.savez_compressed(path, scenario=scen, **fx)
        print("wrote", path, os.path.getsize(path))


# ------------------------------------------------------------------ GAE / ValueNorm fixtures
def _vn_with_history(seed, n_updates=10):
    from onpolicy.utils.valuenorm import ValueNorm
    g = torch.Generator().manual_seed(seed)
    vn = ValueNorm(1)
    hist = []
    for _ in range(n_updates):
        x = torch.randn(257, 1, generator=g) * 3.0 - 4.0
        vn.update(x)
        hist.append(x.numpy().copy())
    return vn, np.stack(hist)


def vn_stats(vn):
    return np.array([vn.running_mean.item(), vn.running_mean_sq.item(), vn.debiasing_term.item()], dtype=np.float32)


def gen_vnorm():
    vn, hist = _vn_with_history(0)
    from onpolicy.utils.valuenorm import ValueNorm
    vn2 = ValueNorm(1)
    stats_seq = []
    for x in hist:
        vn2.update(torch.from_numpy(x))
        stats_seq.append(vn_stats(vn2))
    x = (torch.randn(64, 1, generator=torch.Generator().manual_seed(5)) * 5).numpy()
    out = dict(batches=hist, stats_seq=np.stack(stats_seq), x=x, normalized=vn.normalize(x).numpy(),
               denormalized=vn.denormalize(x),
               fresh_denormalized=ValueNorm(1).denormalize(x))       # first iteration: mean 0, std 0.1
    path = os.path.join(HERE, "vnorm.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path))


def _filled_buffer(args, M, obs_dim, share_dim, act_space, seed):
    from onpolicy.utils.shared_buffer import SharedReplayBuffer
    from gym import spaces
    buf = SharedReplayBuffer(args, M, spaces.Box(-np.inf, np.inf, (obs_dim,), np.float32),
                             spaces.Box(-np.inf, np.inf, (share_dim,), np.float32), act_space)
    rng = np.random.RandomState(seed)
    buf.rewards[:] = rng.normal(-5, 2, buf.rewards.shape)
    buf.value_preds[:] = rng.normal(0, 1, buf.value_preds.shape)
    buf.masks[:] = (rng.uniform(size=buf.masks.shape) > 0.1)
    buf.bad_masks[:] = (rng.uniform(size=buf.bad_masks.shape) > 0.15)
    buf.active_masks[:] = (rng.uniform(size=buf.active_masks.shape) > 0.1)
    return buf, rng


def gen_gae():
    from gym import spaces
    out = {}
    T, N, M = 25, 5, 3
    for use_gae in (1, 0):
        for proper in (0, 1):
            for norm in (1, 0):
                extra = ("" if use_gae else "--use_gae ") + ("--use_proper_time_limits " if proper else "") + \
                        ("" if norm else "--use_valuenorm ")
                args = ref_args("simple_spread", M, 3, "mappo", "--n_rollout_threads %d %s" % (N, extra))
                args.n_rollout_threads = N
                assert args.use_gae == bool(use_gae) and args.use_valuenorm == bool(norm)
                buf, rng = _filled_buffer(args, M, 18, 54, spaces.Discrete(5), seed=100 + use_gae * 4 + proper * 2 + norm)
                vn, _ = _vn_with_history(3) if norm else (None, None)
                nv = rng.normal(0, 1, (N, M, 1)).astype(np.float32)
                tag = "g%dp%dn%d" % (use_gae, proper, norm)
                out[tag + "_rewards"] = buf.rewards.copy(); out[tag + "_value_preds_in"] = buf.value_preds.copy()
                out[tag + "_masks"] = buf.masks.copy(); out[tag + "_bad_masks"] = buf.bad_masks.copy()
                out[tag + "_active_masks"] = buf.active_masks.copy(); out[tag + "_next_value"] = nv
                if norm:
                    out[tag + "_vn"] = vn_stats(vn)
                buf.compute_returns(nv, vn)
                out[tag + "_returns"] = buf.returns.copy(); out[tag + "_value_preds_out"] = buf.value_preds.copy()
                # advantage prologue, verbatim from r_mappo.py:174-182
                if norm:
                    adv = buf.returns[:-1] - vn.denormalize(buf.value_preds[:-1])
                else:
                    adv = buf.returns[:-1] - buf.value_preds[:-1]
                cp = adv.copy(); cp[buf.active_masks[:-1] == 0.0] = np.nan
                out[tag + "_adv_raw"] = adv
                out[tag + "_adv"] = (adv - np.nanmean(cp)) / (np.nanstd(cp) + 1e-5)
    path = os.path.join(HERE, "gae.npz")
    np.savez_compressed(path, T=T, N=N, M=M, gamma=0.99, gae_lambda=0.95, **out)
    print("wrote", path, os.path.getsize(path))


# ------------------------------------------------------------------ generator (minibatch indexing) fixtures
FIELDS = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "value_preds", "returns", "masks",
          "active_masks", "old_action_log_probs", "adv_targ", "available_actions"]


def _tag_rows(buf, rng):
    """Fill every buffer field with distinct random values so gathered rows identify their source."""
    for name in ("share_obs", "obs", "rnn_states", "rnn_states_critic", "value_preds", "returns", "masks",
                 "active_masks", "actions", "action_log_probs", "rewards"):
        a = getattr(buf, name)
        a[:] = rng.standard_normal(a.shape)
    if buf.available_actions is not None:
        buf.available_actions[:] = rng.standard_normal(buf.available_actions.shape)


def gen_gen():
    from gym import spaces
    from onpolicy.envs.mpe.multi_discrete import MultiDiscrete
    from onpolicy.utils.separated_buffer import SeparatedReplayBuffer
    out = {}
    T, N, M = 25, 4, 3
    args = ref_args("simple_spread", M, 3, "rmappo", "--n_rollout_threads %d --hidden_size 8" % N)
    for space_tag, space in (("disc", spaces.Discrete(5)), ("multi", MultiDiscrete([[0, 4], [0, 9]]))):
        buf, rng = _filled_buffer(args, M, 6, 7, space, seed=7)
        _tag_rows(buf, rng)
        adv = rng.standard_normal(buf.rewards.shape).astype(np.float32)
        for name in ("share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "action_log_probs",
                     "value_preds", "returns", "masks", "active_masks"):
            out["shared_%s_buf_%s" % (space_tag, name)] = getattr(buf, name).copy()
        if buf.available_actions is not None:
            out["shared_%s_buf_available_actions" % space_tag] = buf.available_actions.copy()
        out["shared_%s_buf_advantages" % space_tag] = adv
        for gname, call in (("ff", lambda: buf.feed_forward_generator(adv, 2)),
                            ("rec", lambda: buf.recurrent_generator(adv, 2, 10)),
                            ("naive", lambda: buf.naive_recurrent_generator(adv, 2))):
            torch.manual_seed(123)
            for k, sample in enumerate(call()):
                for f, v in zip(FIELDS, sample):
                    if v is not None:
                        out["shared_%s_%s_mb%d_%s" % (space_tag, gname, k, f)] = np.asarray(v)
    # separated buffer (config 3): [T+1, N, .]
    sb = SeparatedReplayBuffer(args, spaces.Box(-np.inf, np.inf, (11,), np.float32),
                               spaces.Box(-np.inf, np.inf, (14,), np.float32), spaces.Discrete(5))
    rng = np.random.RandomState(9)
    _tag_rows(sb, rng)
    adv = rng.standard_normal(sb.rewards.shape).astype(np.float32)
    for name in ("share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "action_log_probs",
                 "value_preds", "returns", "masks", "active_masks", "available_actions"):
        out["sep_buf_%s" % name] = getattr(sb, name).copy()
    out["sep_buf_advantages"] = adv
    for gname, call in (("ff", lambda: sb.feed_forward_generator(adv, 2)),
                        ("rec", lambda: sb.recurrent_generator(adv, 2, 10)),
                        ("naive", lambda: sb.naive_recurrent_generator(adv, 2))):
        torch.manual_seed(321)
        for k, sample in enumerate(call()):
            for f, v in zip(FIELDS, sample):
                if v is not None:
                    out["sep_%s_mb%d_%s" % (gname, k, f)] = np.asarray(v)
    path = os.path.join(HERE, "generators.npz")
    np.savez_compressed(path, T=T, N=N, M=M, **out)
    print("wrote", path, os.path.getsize(path))


# ------------------------------------------------------------------ PPO update fixtures
def _sd_np(module, prefix):
    return {prefix + k: v.detach().cpu().numpy().copy() for k, v in module.state_dict().items()
            if "fc_h" not in k or prefix.endswith("0.")}


def _np(t):
    return t.detach().cpu().numpy().copy()


def _popart_state(v_out):
    """PopArt's six tensors (popart.py:21-27).  On cuda they are plain attributes, not registered parameters."""
    return {k: _np(getattr(v_out, k)) for k in ("weight", "bias", "mean", "mean_sq", "debiasing_term", "stddev")}


def make_ppo_fixture(scenario, M, L, algo, extra, B, seed, zero_avail=False, n_updates=2, device="cpu"):
    """Run the reference's R_MAPPO.ppo_update `n_updates` times on one crafted minibatch and record everything:
    initial weights, the sample, the six return values, PRE-clip gradients, post-step weights and Adam state."""
    from onpolicy.envs.mpe.MPE_env import MPEEnv
    from onpolicy.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    from onpolicy.algorithms.r_mappo.r_mappo import R_MAPPO
    args = ref_args(scenario, M, L, algo, extra)
    env = MPEEnv(args)
    agent = 0 if scenario != "simple_speaker_listener" else 1
    obs_sp, sh_sp, act_sp = env.observation_space[agent], env.share_observation_space[agent], env.action_space[agent]
    torch.manual_seed(seed)
    device = torch.device(device)
    policy = R_MAPPOPolicy(args, obs_sp, sh_sp, act_sp, device=device)
    trainer = R_MAPPO(args, policy, device=device)
    popart = bool(args.use_popart)
    # make LayerNorm affines and biases non-trivial so the folded-affine chain rule is exercised
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for net in (policy.actor, policy.critic):
            for name, prm in net.named_parameters():
                if "fc_h" in name:
                    continue
                if prm.dim() == 1:
                    prm.add_((torch.randn(prm.shape, generator=g) * 0.2).to(prm.device))
                elif "action_out" in name or "v_out" in name:
                    prm.add_((torch.randn(prm.shape, generator=g) * 0.3).to(prm.device))
                elif "fc2" in name and args.layer_N > 1:
                    # the hidden blocks start as identical clones of fc_h (mlp.py:20-22): make them differ
                    prm.add_((torch.randn(prm.shape, generator=g) * 0.05).to(prm.device))
    D, S, H = obs_sp.shape[0], sh_sp.shape[0], args.hidden_size
    rng = np.random.RandomState(seed)
    recurrent = bool(args.use_recurrent_policy)
    Lc = args.data_chunk_length
    Nc = B // Lc if recurrent else B
    obs = rng.standard_normal((B, D)).astype(np.float32) * 0.7
    share_obs = rng.standard_normal((B, S)).astype(np.float32) * 0.7
    RN = args.recurrent_N
    rnn_a = (rng.standard_normal((Nc, RN, H)) * 0.5).astype(np.float32) if recurrent else np.zeros((B, 1, H), np.float32)
    rnn_c = (rng.standard_normal((Nc, RN, H)) * 0.5).astype(np.float32) if recurrent else np.zeros((B, 1, H), np.float32)
    masks = (rng.uniform(size=(B, 1)) > 0.08).astype(np.float32)
    active = (rng.uniform(size=(B, 1)) > 0.1).astype(np.float32)
    avail = None
    if act_sp.__class__.__name__ == "Discrete":
        avail = np.ones((B, act_sp.n), np.float32)
        if zero_avail:
            avail[rng.uniform(size=avail.shape) < 0.15] = 0.0
            avail[:, 0] = 1.0
    with torch.no_grad():
        policy.actor.eval(); policy.critic.eval()
        if recurrent:   # evaluate on the chunked layout to get consistent log-probs
            feats_a = policy.actor.base(torch.from_numpy(obs).to(device))
            feats_a, _ = policy.actor.rnn(feats_a, torch.from_numpy(rnn_a).to(device), torch.from_numpy(masks).to(device))
            if act_sp.__class__.__name__ == "Discrete":
                dist = [policy.actor.act.action_out(feats_a, torch.from_numpy(avail).to(device))]
            else:
                dist = [ao(feats_a) for ao in policy.actor.act.action_outs]
            actions = torch.cat([d.sample() for d in dist], -1)
            logp = torch.cat([d.log_probs(actions[:, i:i + 1]) for i, d in enumerate(dist)], -1)
            values, _ = policy.critic(share_obs, rnn_c, masks)
        else:
            values, actions, logp, _, _ = policy.get_actions(share_obs, obs, rnn_a, rnn_c, masks, avail)
    actions, logp, values = actions.cpu(), logp.cpu(), values.cpu()
    actions = actions.numpy().astype(np.float32)
    old_logp = (logp.numpy() + rng.normal(0, 0.15, logp.shape)).astype(np.float32)
    adv = rng.standard_normal((B, 1)).astype(np.float32)
    value_preds = (values.numpy() + rng.normal(0, 0.25, (B, 1))).astype(np.float32)
    if args.use_valuenorm or popart:
        returns = rng.normal(-5, 3, (B, 1)).astype(np.float32)
        if not popart:
            trainer.value_normalizer.update(torch.from_numpy(rng.normal(-4, 2, (300, 1)).astype(np.float32)))
    else:
        returns = (values.numpy() + rng.normal(0, 1.0, (B, 1))).astype(np.float32)
    sample = (share_obs, obs, rnn_a, rnn_c, actions, value_preds, returns, masks, active, old_logp, adv, avail)
    out = dict(B=B, D=D, S=S, recurrent=int(recurrent), chunk=Lc)
    out.update({"args_" + k: np.asarray(getattr(args, k)) for k in (
        "use_popart", "use_ReLU", "use_feature_normalization", "use_valuenorm", "use_huber_loss", "use_clipped_value_loss",
        "use_policy_active_masks", "use_value_active_masks", "use_max_grad_norm", "max_grad_norm", "clip_param",
        "entropy_coef", "value_loss_coef", "huber_delta", "lr", "critic_lr", "opti_eps", "weight_decay", "hidden_size",
        "layer_N", "gain")})
    out["act_heads"] = np.array([act_sp.n] if act_sp.__class__.__name__ == "Discrete"
                                else list((act_sp.high - act_sp.low + 1).astype(int)))
    for name, arr in zip(FIELDS, sample):
        if arr is not None and (recurrent or not name.startswith("rnn_states")):
            out["sample_" + name] = arr
    out.update(_sd_np(policy.actor, "actor0."))
    out.update(_sd_np(policy.critic, "critic0."))
    if args.use_valuenorm:
        out["vn0"] = vn_stats(trainer.value_normalizer)
    if popart:
        out["critic_state_dict_keys"] = np.array(sorted(policy.critic.state_dict().keys()))
        out["critic_optimizer_numel"] = sum(p.numel() for g in policy.critic_optimizer.param_groups for p in g["params"])
        out.update({"popart0." + k: v for k, v in _popart_state(policy.critic.v_out).items()})
        out["values0"] = values.numpy()
    trainer.prep_training()
    import torch.nn.utils as U
    real_clip = U.clip_grad_norm_
    for u in range(n_updates):
        pre = []

        def spy(params, max_norm, *a, **k):
            params = list(params)
            pre.append([None if p.grad is None else p.grad.detach().clone() for p in params])
            return real_clip(params, max_norm, *a, **k)
        U.clip_grad_norm_ = spy
        try:
            vl, cgn, pl, ent, agn, ratio = trainer.ppo_update(sample)
        finally:
            U.clip_grad_norm_ = real_clip
        tag = "u%d_" % u
        out[tag + "value_loss"] = vl.item(); out[tag + "critic_grad_norm"] = float(cgn)
        out[tag + "policy_loss"] = pl.item(); out[tag + "dist_entropy"] = ent.item()
        out[tag + "actor_grad_norm"] = float(agn); out[tag + "ratio_mean"] = ratio.mean().item()
        for net, nm, grads in ((policy.actor, "actor", pre[0]), (policy.critic, "critic", pre[1])):
            for (pname, _), gr in zip(net.named_parameters(), grads):
                if gr is not None:
                    out[tag + nm + "_grad." + pname] = _np(gr)
            out.update(_sd_np(net, tag + nm + "."))
        for opt, nm, net in ((policy.actor_optimizer, "actor", policy.actor), (policy.critic_optimizer, "critic", policy.critic)):
            for (pname, prm) in net.named_parameters():
                st = opt.state.get(prm, None)
                if st:
                    out[tag + nm + "_exp_avg." + pname] = _np(st["exp_avg"])
                    out[tag + nm + "_exp_avg_sq." + pname] = _np(st["exp_avg_sq"])
        if args.use_valuenorm:
            out[tag + "vn"] = vn_stats(trainer.value_normalizer)
        if popart:
            out.update({tag + "popart." + k: v for k, v in _popart_state(policy.critic.v_out).items()})
            with torch.no_grad():
                v_after, _ = policy.critic(share_obs, rnn_c, masks)
            out[tag + "values_after"] = _np(v_after)
            out[tag + "denorm_value_preds"] = np.asarray(trainer.value_normalizer.denormalize(value_preds))
    return out


PPO_CASES = {
    # C1/C2: spread, shared MLP, Tanh (the reference script passes --use_ReLU, a store_false flag), defaults
    "mlp_tanh": dict(scenario="simple_spread", M=3, L=3, algo="mappo", extra="--use_ReLU --lr 7e-4 --critic_lr 7e-4",
                     B=700, seed=3),
    # option coverage: ReLU, no ValueNorm, MSE, no active masks, tight grad clip, masked available actions
    "mlp_relu_opts": dict(scenario="simple_spread", M=3, L=3, algo="mappo",
                          extra="--use_valuenorm --use_huber_loss --use_policy_active_masks --use_value_active_masks "
                                "--max_grad_norm 0.05 --entropy_coef 0.03 --weight_decay 0.01",
                          B=300, seed=5, zero_avail=True),
    # unclipped value loss, no feature norm, listener of speaker_listener (D=11, S=14)
    "mlp_listener": dict(scenario="simple_speaker_listener", M=2, L=3, algo="mappo",
                         extra="--use_clipped_value_loss --use_feature_normalization --share_policy", B=257, seed=7),
    # recurrent_N = 2 (config.py:219; nn.GRU(num_layers=2), rnn.py:13): simple_reference, MultiDiscrete heads, chunks of 10
    "gru2_reference": dict(scenario="simple_reference", M=2, L=3, algo="rmappo",
                           extra="--recurrent_N 2 --lr 7e-4 --critic_lr 7e-4 --data_chunk_length 10", B=330, seed=19),
    # layer_N > 1 (config.py:201): two and three hidden blocks after fc1
    "mlp_layer2": dict(scenario="simple_spread", M=3, L=3, algo="mappo", extra="--use_ReLU --layer_N 2", B=500, seed=11),
    "mlp_layer3_relu": dict(scenario="simple_spread", M=3, L=3, algo="mappo",
                            extra="--layer_N 3 --use_huber_loss --max_grad_norm 0.5", B=333, seed=12),
    # C4: simple_reference, shared GRU policy (rmappo), MultiDiscrete heads (5, 10), chunks of 10 steps
    "gru_reference": dict(scenario="simple_reference", M=2, L=3, algo="rmappo",
                          extra="--use_ReLU --lr 7e-4 --critic_lr 7e-4 --data_chunk_length 10", B=370, seed=9),
}


def gen_ppo(only=None):
    for tag, kw in PPO_CASES.items():
        if only and tag not in only:
            continue
        fx = make_ppo_fixture(**kw)
        path = os.path.join(HERE, "ppo_%s.npz" % tag)
        np.savez_compressed(path, **fx)
        print("wrote", path, os.path.getsize(path), "policy_loss", fx["u0_policy_loss"], "agn", fx["u0_actor_grad_norm"],
              "cgn", fx["u0_critic_grad_norm"])


# ------------------------------------------------------------------ PopArt (cuda only)
def gen_popart():
    """--use_popart.  PopArt.__init__ (popart.py:21-27) builds its tensors as nn.Parameter(...).to(dtype, device): on
    the reference's default cuda:0 (train_mpe.py:86-96) that `.to()` returns a PLAIN tensor, so weight / bias / mean /
    mean_sq / debiasing_term / stddev are ordinary attributes: not in parameters(), not in state_dict(), never seen by
    the optimiser or by clip_grad_norm_, and only PopArt.update's rescale (popart.py:69-70) moves weight / bias.  On
    CPU `.to()` returns the registered Parameter itself and update() raises TypeError at its first assignment.  So
    this fixture can only be recorded where a GPU is: run it on the B200 box against baseline/_ref."""
    assert torch.cuda.is_available(), "the reference's PopArt only runs on cuda (see docstring)"
    out_dir = os.path.join(os.path.dirname(os.path.dirname(HERE)), "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    for tag, kw in (("popart_cuda", dict(scenario="simple_spread", M=3, L=3, algo="mappo",
                                         extra="--use_ReLU --lr 7e-4 --critic_lr 7e-4 --use_popart --use_valuenorm",
                                         B=700, seed=13, n_updates=3)),
                    ("popart_mse_cuda", dict(scenario="simple_spread", M=3, L=3, algo="mappo",
                                             extra="--use_popart --use_valuenorm --use_huber_loss --max_grad_norm 0.5",
                                             B=300, seed=15, n_updates=2))):
        fx = make_ppo_fixture(device="cuda:0", **kw)
        path = os.path.join(out_dir, "ppo_%s.npz" % tag)
        np.savez_compressed(path, torch_version=str(torch.__version__), **fx)
        print("wrote", path, os.path.getsize(path), "value_loss", fx["u0_value_loss"], "cgn", fx["u0_critic_grad_norm"],
              "state_dict keys with v_out:", [k for k in fx["critic_state_dict_keys"] if "v_out" in k],
              "optimizer numel", fx["critic_optimizer_numel"])


# ------------------------------------------------------------------ whole-runner fixtures (two consecutive iterations)
def make_runner_fixture(scenario, M, L, algo, extra, N=8, iters=2, seed=1):
    """Run the reference's OWN MPERunner.run() (runner/shared/mpe_runner.py:16-79) for `iters` iterations on a
    DummyVecEnv of N reference envs and record, by wrapping (never editing) its methods: the world state behind every
    observation (warm-up reset, every auto-reset), the sampled actions, the buffer after each rollout / compute, the
    normalised advantages train() forms (r_mappo.py:174-182, restated verbatim on the same buffer), the six train_info
    scalars, and the weights / ValueNorm state after each iteration."""
    from pathlib import Path
    import tempfile
    from onpolicy.envs.mpe.MPE_env import MPEEnv
    from onpolicy.envs.env_wrappers import DummyVecEnv
    from onpolicy.runner.shared.mpe_runner import MPERunner
    T = 25
    args = ref_args(scenario, M, L, algo, "--n_rollout_threads %d --num_env_steps %d --cuda --n_training_threads 1 "
                    "--log_interval 1000 --save_interval 1000 %s" % (N, iters * T * N, extra))
    args.n_rollout_threads = N
    torch.manual_seed(seed); np.random.seed(seed)

    def env_fn(rank):
        def init():
            e = MPEEnv(args)
            e.seed(seed + rank * 1000)
            return e
        return init
    envs = DummyVecEnv([env_fn(i) for i in range(N)])
    run_dir = Path(tempfile.mkdtemp(prefix="mappo_ref_run_"))
    runner = MPERunner({"all_args": args, "envs": envs, "eval_envs": None, "num_agents": M,
                        "device": torch.device("cpu"), "run_dir": run_dir})
    pol, tr, buf = runner.policy, runner.trainer, runner.buffer
    out = dict(N=N, M=M, L=L, T=T, iters=iters, scenario=scenario, algo=algo,
               recurrent=int(bool(args.use_recurrent_policy)), chunk=args.data_chunk_length)
    out.update({"args_" + k: np.asarray(getattr(args, k)) for k in (
        "use_ReLU", "use_feature_normalization", "use_valuenorm", "use_huber_loss", "use_clipped_value_loss",
        "use_policy_active_masks", "use_value_active_masks", "use_max_grad_norm", "max_grad_norm", "clip_param",
        "entropy_coef", "value_loss_coef", "huber_delta", "lr", "critic_lr", "opti_eps", "weight_decay", "hidden_size",
        "layer_N", "recurrent_N", "gain", "ppo_epoch", "num_mini_batch", "gamma", "gae_lambda", "use_gae",
        "use_proper_time_limits", "use_centralized_V")})
    out.update(_sd_np(pol.actor, "actor0."))
    out.update(_sd_np(pol.critic, "critic0."))
    states = lambda: [world_state(e) for e in envs.envs]
    rec = dict(step_states=[], actions_env=[], obs=[], rewards=[], dones=[])
    it = {"k": 0}
    real_reset, real_step, real_compute, real_train = envs.reset, envs.step, runner.compute, runner.train

    def reset_spy():
        o = real_reset()
        st = states()
        out["warm_p_pos"] = np.array([x[0] for x in st]); out["warm_lm_pos"] = np.array([x[2] for x in st])
        out["warm_goal"] = np.array([x[4] for x in st]); out["warm_obs"] = np.array(o)
        return o

    def step_spy(actions_env):
        r = real_step(actions_env)
        st = states()                                          # after any auto-reset: what the returned obs shows
        rec["step_states"].append((np.array([x[0] for x in st]), np.array([x[1] for x in st]),
                                   np.array([x[2] for x in st]), np.array([x[4] for x in st])))
        rec["actions_env"].append(np.array(actions_env)); rec["obs"].append(np.array(r[0]))
        rec["rewards"].append(np.array(r[1])); rec["dones"].append(np.array(r[2]))
        return r

    def compute_spy():
        k = it["k"]
        for name in ("obs", "share_obs", "masks", "rewards", "actions", "action_log_probs", "rnn_states",
                     "rnn_states_critic", "active_masks", "bad_masks"):
            out["it%d_buf_%s" % (k, name)] = getattr(buf, name).copy()
        out["it%d_value_preds_rollout" % k] = buf.value_preds.copy()
        real_compute()
        out["it%d_returns" % k] = buf.returns.copy()
        out["it%d_value_preds" % k] = buf.value_preds.copy()

    def train_spy():
        k = it["k"]
        vn = tr.value_normalizer
        if vn is not None:
            out["it%d_vn_before" % k] = vn_stats(vn)
            adv = buf.returns[:-1] - vn.denormalize(buf.value_preds[:-1])
        else:
            adv = buf.returns[:-1] - buf.value_preds[:-1]
        cp = adv.copy(); cp[buf.active_masks[:-1] == 0.0] = np.nan          # r_mappo.py:177-181, verbatim
        out["it%d_adv" % k] = (adv - np.nanmean(cp)) / (np.nanstd(cp) + 1e-5)
        info = real_train()
        for name, v in info.items():
            out["it%d_info_%s" % (k, name)] = float(v)
        out.update(_sd_np(pol.actor, "it%d_actor." % k))
        out.update(_sd_np(pol.critic, "it%d_critic." % k))
        if vn is not None:
            out["it%d_vn_after" % k] = vn_stats(vn)
        for name in ("obs", "masks", "rnn_states", "rnn_states_critic"):      # after_update: slot 0 of the next rollout
            out["it%d_after_%s0" % (k, name)] = getattr(buf, name)[0].copy()
        for key, arr in rec.items():
            if key == "step_states":
                out["it%d_p_pos" % k] = np.array([x[0] for x in arr]); out["it%d_p_vel" % k] = np.array([x[1] for x in arr])
                out["it%d_lm_pos" % k] = np.array([x[2] for x in arr]); out["it%d_goal" % k] = np.array([x[3] for x in arr])
            else:
                out["it%d_%s" % (k, key)] = np.array(arr)
            arr.clear()
        it["k"] += 1
        return info

    envs.reset, envs.step, runner.compute, runner.train = reset_spy, step_spy, compute_spy, train_spy
    runner.run()                                               # the reference's own loop
    assert it["k"] == iters
    return out


RUNNER_CASES = {
    # C1's flags (train_mpe_spread.sh:4-17 with the MLP policy), 8 envs, 2 epochs
    "spread_mappo": dict(scenario="simple_spread", M=3, L=3, algo="mappo",
                         extra="--use_ReLU --gain 0.01 --lr 7e-4 --critic_lr 7e-4 --ppo_epoch 2 --num_mini_batch 1"),
    # C4's flags (train_mpe_reference.sh): shared GRU, chunks of 10
    "reference_rmappo": dict(scenario="simple_reference", M=2, L=3, algo="rmappo",
                             extra="--use_ReLU --gain 0.01 --lr 7e-4 --critic_lr 7e-4 --ppo_epoch 2 --num_mini_batch 1 "
                                   "--data_chunk_length 10"),
}


def gen_runner():
    for tag, kw in RUNNER_CASES.items():
        fx = make_runner_fixture(**kw)
        path = os.path.join(HERE, "runner_%s.npz" % tag)
        np.savez_compressed(path, **fx)
        print("wrote", path, os.path.getsize(path), {k[9:]: round(float(v), 6) for k, v in fx.items() if k.startswith("it1_info_")})


# ------------------------------------------------------------------ checkpoint written by the reference
def gen_ckpt():
    """actor.pt / critic.pt / vnorm.pt exactly as the reference's Runner.save writes them (base_runner.py:127-135:
    torch.save of the three state_dicts), plus what those weights compute on a fixed input, for the restore test."""
    from onpolicy.envs.mpe.MPE_env import MPEEnv
    from onpolicy.algorithms.r_mappo.algorithm.rMAPPOPolicy import R_MAPPOPolicy
    from onpolicy.algorithms.r_mappo.r_mappo import R_MAPPO
    args = ref_args("simple_spread", 3, 3, "mappo", "--use_ReLU")
    env = MPEEnv(args)
    torch.manual_seed(21)
    policy = R_MAPPOPolicy(args, env.observation_space[0], env.share_observation_space[0], env.action_space[0])
    trainer = R_MAPPO(args, policy)
    g = torch.Generator().manual_seed(22)
    with torch.no_grad():
        for net in (policy.actor, policy.critic):
            for name, prm in net.named_parameters():
                if "fc_h" not in name:
                    prm.add_(torch.randn(prm.shape, generator=g) * 0.1)
    trainer.value_normalizer.update(torch.randn(100, 1, generator=g) * 2 - 3)
    d = os.path.join(HERE, "ref_ckpt")
    os.makedirs(d, exist_ok=True)
    torch.save(policy.actor.state_dict(), os.path.join(d, "actor.pt"))
    torch.save(policy.critic.state_dict(), os.path.join(d, "critic.pt"))
    torch.save(trainer.value_normalizer.state_dict(), os.path.join(d, "vnorm.pt"))
    obs = torch.randn(64, 18, generator=g).numpy()
    share = torch.randn(64, 54, generator=g).numpy()
    z, m = np.zeros((64, 1, 64), np.float32), np.ones((64, 1), np.float32)
    with torch.no_grad():
        policy.actor.eval(); policy.critic.eval()
        values, actions, logp, _, _ = policy.get_actions(share, obs, z, z, m, None, deterministic=True)
    np.savez_compressed(os.path.join(d, "io.npz"), obs=obs, share_obs=share, values=values.numpy(), actions=actions.numpy(),
                        log_probs=logp.numpy(), vn=vn_stats(trainer.value_normalizer),
                        denorm=trainer.value_normalizer.denormalize(values.numpy()))
    print("wrote", d, sorted(os.listdir(d)))


PARTS = {"ckpt": gen_ckpt, "env": gen_env, "vnorm": gen_vnorm, "gae": gen_gae, "gen": gen_gen, "ppo": gen_ppo, "runner": gen_runner,
         "popart": gen_popart}
DEFAULT_PARTS = ["env", "vnorm", "gae", "gen", "ppo", "runner", "ckpt"]

if __name__ == "__main__":
    todo = sys.argv[1:] or DEFAULT_PARTS
    for p in todo:
        PARTS[p]()
