"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE
(/root/reference/onpolicy, imported through tests/golden/ref_shims.py).

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden.py [env] [gae] [gen] [ppo] [vnorm]
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

ref_shims.install()


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
    goal = np.array([w.landmarks.index(a.goal_b) if getattr(a, "goal_b", None) is not None else -1
                     for a in w.agents])
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
    rng = np.random.RandomState(seed)
    # one reference env per slot; they share the process-global np.random like DummyVecEnv
    np.random.seed(seed)
    envs = [MPEEnv(args) for _ in range(N)]
    heads = []
    for sp in envs[0].action_space:
        heads.append([sp.n] if sp.__class__.__name__ == "Discrete" else list((sp.high - sp.low + 1).astype(int)))
    H = max(len(h) for h in heads)
    obs0 = [e.reset() for e in envs]
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


def gen_env():
    for scen, M, L, tag in (("simple_spread", 3, 3, "spread_m3"), ("simple_spread", 10, 3, "spread_m10"),
                            ("simple_speaker_listener", 2, 3, "speaker_listener"),
                            ("simple_reference", 2, 3, "reference")):
        fx = make_env_fixture(scen, M, L, N=4 if M == 10 else 6)
        path = os.path.join(HERE, "env_%s.npz" % tag)
        np.savez_compressed(path, scenario=scen, **fx)
        print("wrote", path, os.path.getsize(path))


PARTS = {"env": gen_env}

if __name__ == "__main__":
    todo = sys.argv[1:] or list(PARTS)
    for p in todo:
        PARTS[p]()
