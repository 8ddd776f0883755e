#!/usr/bin/env python
"""bench.py -- agent-steps/s of full MAPPO iterations (rollout + GAE + PPO update) on MPE simple_spread.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--envs E]

A "step" is ONE MAPPO iteration: episode_length x (policy forward, env step, insert), bootstrap value + GAE with
ValueNorm, then ppo_epoch x (actor update, critic update, clip, Adam).  One iteration produces N*T*M agent-steps.
Workload = BASELINE.json configs[1] (C2): simple_spread, 3 agents, 3 landmarks, 65,536 envs per GPU, T = 25,
shared MLP policy (Tanh), ppo_epoch = 10, num_mini_batch = 1, GAE + ValueNorm; weak scaling (envs per GPU fixed).

Prints ONE JSON line (rank 0).  `value`: device-resident runner.  `e2e`: the same iteration driven through the
reference-facing vec-env / policy / buffer calls with HOST (pinned) buffers at the boundary, copies in the timed
region.  `roofline`: the dominant kernel (fused actor forward+loss+backward) at top level plus `kernels`, one entry per
kernel of the iteration (actor / critic update, rollout forwards, env step, GAE, gradient reduce, Adam), each timed
with CUDA events inside the timed region.  `cpu_baseline`: the UNMODIFIED reference (baseline/_ref: SubprocVecEnv x 128
worker processes + MPERunner + R_MAPPO on torch CPU, the C1 flags of train_mpe_spread.sh) timed on this box's host
cores on a bounded sample; `cpu_baseline_port`: the CPU port (oracle/iteration.py) as a second figure.
`--impl reference` times the unmodified reference alone.  Under torchrun (N > 1) the line also carries `dp_check`:
replicas bit-identical after an update, and the reduced gradient of a 2-rank job equal to the single-GPU gradient on
the union of the env shards.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "agent-steps/sec (rollout+GAE+PPO update), MPE simple_spread"
UNIT = "agent-steps/s"
M_AGENTS, N_LANDMARKS, T_EP, PPO_EPOCH = 3, 3, 25, 10
D_OBS, S_OBS, N_ACT, H = 18, 54, 5, 64


SCENARIO, ALGO, SEPARATED, CONFIG = "simple_spread", "mappo", False, "C2"
ACTOR_UPDATE_DRAM_BYTES = 439284480      # 432,719,104 read + 6,565,376 written (profiles/r01_update_tc_v5.md)


def flags(n_envs):
    return ("--env_name MPE --algorithm_name %s --scenario_name %s --num_agents %d --num_landmarks %d "
            "--n_rollout_threads %d --episode_length %d --ppo_epoch %d --num_mini_batch 1 --use_ReLU --gain 0.01 "
            "--lr 7e-4 --critic_lr 7e-4 --seed 1 %s" % (ALGO, SCENARIO, M_AGENTS, N_LANDMARKS, n_envs, T_EP, PPO_EPOCH,
                                                        "--share_policy" if SEPARATED else "")).split()


def set_workload(num_agents, config=None):
    """Default M = 3 is C2 (the configuration the metric is quoted on).  --agents 10 / --config c5: the C5 shape;
    --config c3: speaker_listener with separated policies (15 epochs), c3r: the same with GRU policies; --config c4: simple_reference with the
    shared GRU policy (rmappo, chunks of 10, 15 epochs).  Only the default is the contract line; the others are
    supplementary measurements of the remaining BASELINE configs."""
    global M_AGENTS, D_OBS, S_OBS, SCENARIO, ALGO, SEPARATED, PPO_EPOCH, CONFIG
    M_AGENTS = int(num_agents)
    if config == "c5":
        M_AGENTS = 10
    if config == "c3":
        SCENARIO, SEPARATED, M_AGENTS, PPO_EPOCH = "simple_speaker_listener", True, 2, 15
    if config == "c3r":                                    # the reference's train_mpe_comm.sh as is: separated + GRU
        SCENARIO, SEPARATED, ALGO, M_AGENTS, PPO_EPOCH = "simple_speaker_listener", True, "rmappo", 2, 15
    if config == "c4":
        SCENARIO, ALGO, M_AGENTS, PPO_EPOCH = "simple_reference", "rmappo", 2, 15
    D_OBS = 4 + 2 * N_LANDMARKS + 4 * (M_AGENTS - 1)       # simple_spread.py:93-103
    S_OBS = D_OBS * M_AGENTS
    CONFIG = {"c3": "C3", "c3r": "C3 (rmappo)", "c4": "C4"}.get(config, "C2" if M_AGENTS == 3 else "C5" if M_AGENTS == 10 else "custom")


def workload_name(n_envs, n_gpus):
    kind = ("separated GRU policies, chunks of 10" if ALGO == "rmappo" else "separated MLP policies") if SEPARATED else ("shared GRU policy, chunks of 10" if ALGO == "rmappo" else "shared MLP")
    return ("%s: MPE %s M=%d L=3, %d envs/GPU x %d GPU, T=25, %s (Tanh, H=64), ppo_epoch=%d, "
            "num_mini_batch=1, GAE+ValueNorm" % (CONFIG, SCENARIO, M_AGENTS, n_envs, n_gpus, kind, PPO_EPOCH))


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------- CPU path
def cpu_iteration_rate(n_envs, steps, warmup, threads):
    """agent-steps/s of oracle/iteration.py (the CPU port of the reference path) on `threads` host threads."""
    import torch
    from mappo_b200.config import parse_args
    from mappo_b200.algorithms.r_mappo.algorithm.r_actor_critic import _reference_init
    from oracle.iteration import OracleMAPPO
    torch.set_num_threads(threads)
    a = parse_args(flags(n_envs))
    torch.manual_seed(1)
    asd = _reference_init(a, D_OBS, [N_ACT], False)
    csd = _reference_init(a, S_OBS, [1], True)
    o = OracleMAPPO(a, asd, csd)
    o.warmup()
    for _ in range(warmup):
        o.iteration()
    t0 = time.perf_counter()
    for _ in range(steps):
        o.iteration()
    dt = time.perf_counter() - t0
    return n_envs * T_EP * M_AGENTS * steps / dt, dt / steps


def reference_rate(n_envs, steps, warmup, threads):
    """agent-steps/s of the UNMODIFIED reference (baseline/_ref through baseline/ref_harness.py): its own SubprocVecEnv
    (one worker process per env) + MPERunner + R_MAPPO on torch CPU, flags of train_mpe_spread.sh with the MLP policy."""
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_harness
    if not ref_harness.available():
        return None
    fl = ref_harness.c1_flags(n_envs, threads, scenario=SCENARIO, num_agents=M_AGENTS, num_landmarks=N_LANDMARKS,
                              algo=ALGO, ppo_epoch=PPO_EPOCH)
    return ref_harness.time_reference(fl, steps, warmup)


def run_reference(ns):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores, rank 0 only.
    The reference is the unmodified `onpolicy` package from baseline/_ref when it travelled with the snapshot (kind
    "reference"); otherwise the CPU port oracle/iteration.py (kind "port")."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = len(os.sched_getaffinity(0))
    threads = cores if ns.ref_threads <= 0 else ns.ref_threads
    r = None if ns.ref_kind == "port" else reference_rate(ns.ref_envs, ns.steps, ns.warmup, threads)
    if r is not None:
        val, sec, n_envs, kind = r["value"], r["sec_per_iter"], r["n_envs"], "reference"
        sample = ("unmodified onpolicy (baseline/_ref): SubprocVecEnv x %d worker processes + MPERunner + R_MAPPO on torch "
                  "CPU (%d training threads), %d timed iterations (each a full rollout+GAE+%d-epoch update); median "
                  "iteration %.3f s = collect %.3f + env step %.3f + insert %.3f + compute %.3f + train %.3f; the reference "
                  "cannot spawn 65,536 env processes, so its own configuration size (C1) is the sample and agent-steps/s "
                  "is size-normalised" % (n_envs, threads, ns.steps, PPO_EPOCH, r["sec_median"], r["split"]["collect"],
                                          r["split"]["env_step"], r["split"]["insert"], r["split"]["compute"],
                                          r["split"]["train"]))
    else:
        n_envs, kind = ns.cpu_envs, "port"
        val, sec = cpu_iteration_rate(n_envs, ns.steps, ns.warmup, cores)
        sample = "%d envs x %d iterations (each a full rollout+GAE+%d-epoch update) of the same workload, oracle/iteration.py" % (
            n_envs, ns.steps, PPO_EPOCH)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": ns.gpus, "steps": ns.steps,
        "warmup": ns.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (policy, buffer), f64 (env physics)", "data": "synthetic",
        "config": {"workload": workload_name(ns.envs, ns.gpus), "cpu_sample_envs": n_envs,
                   "same_config_note": "same scenario, policy, T, ppo_epoch and minibatching as the GPU arm; envs per "
                                       "iteration bounded to what the CPU implementation can run (metric is per agent-step)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ---------------------------------------------------------------------------------------------- GPU path
def build_runner(n_envs, device, rank, world, host_io=False):
    import torch
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    if SEPARATED:
        from mappo_b200.runner.separated.mpe_runner import MPERunner
    else:
        from mappo_b200.runner.shared.mpe_runner import MPERunner
    a = parse_args(flags(n_envs))
    torch.manual_seed(1)                                 # same initial weights on every rank (replicated parameters)
    from mappo_b200 import parallel as _par
    envs = MPEVecEnv(a, device=device, env_id_offset=_par.env_shard(n_envs, rank), host_io=host_io)
    runner = MPERunner({"all_args": a, "envs": envs, "eval_envs": None, "device": device, "num_agents": M_AGENTS,
                        "run_dir": None})
    from mappo_b200 import parallel
    for tr in (runner.trainer if SEPARATED else [runner.trainer]):
        parallel.attach(tr, world)
    return a, envs, runner


def host_boundary_iteration(runner, envs, hb):
    """One iteration driven like the reference's MPERunner.run body (runner/shared/mpe_runner.py:24-40,95-139) with
    HOST arrays crossing every reference-facing call: policy outputs come back to the host (`_t2n` -> the library's
    drop-in `host_mirror.t2n`: pinned, one sync), the one-hot `actions_env` is built on the host and goes in through
    envs.step, obs / rewards / dones come back as host arrays, masks are built on the host, and buffer.insert receives
    host arrays.  Arrays the library itself brought out of the device (obs, rewards, policy outputs) are recognised
    by insert and served device -> device (INTEGRATION.md "pass-through"); only bytes that really cross PCIe are
    counted.  Returns (h2d_bytes, d2h_bytes, info)."""
    import numpy as np
    import torch
    from mappo_b200 import host_mirror
    b, pol = runner.buffer, runner.trainer.policy
    N, M = b.n_rollout_threads, b.num_agents
    h2d = d2h = 0

    def crossing(x):                       # bytes insert() has to send host -> device for this argument
        return 0 if host_mirror.lookup(x) is not None else x.nbytes

    for step in range(b.episode_length):
        # cent_obs = the buffer's share_obs slot [N, M, S] (its M rows per env are one stored row: R_Critic.forward
        # evaluates each env once); the reference passes np.concatenate of the same slot (mpe_runner.py:99)
        values, actions, logp, rs, rsc = pol.get_actions(b.share_obs[step], b.obs[step].view(N * M, -1), None, None, None)
        out = host_mirror.t2n_many({"values": values.view(N, M, 1), "actions": actions.view(N, M, 1),
                                    "logp": logp.view(N, M, 1)})          # mpe_runner.py:102-104: np.split(_t2n(.))
        values, actions, logp = out["values"], out["actions"], out["logp"]
        d2h += values.nbytes + actions.nbytes + logp.nbytes
        actions_env = host_mirror.onehot(actions, [N_ACT])               # mpe_runner.py:110-121 (np.eye(n)[actions]), uint8
        if host_mirror.lookup(actions_env) is not None:                  # built next to the ids on the device: came OUT as bytes
            d2h += actions_env.nbytes
        else:                                                            # built on the host: goes IN with envs.step
            h2d += actions_env.nbytes
        obs, rewards, dones, infos = envs.step(actions_env)              # host in, host out
        d2h += obs.nbytes + rewards.nbytes + dones.nbytes + N * M * 4    # + the individual rewards behind `infos`
        masks = hb["masks"].numpy()
        masks[..., 0] = ~dones                                           # mpe_runner.py:128-131 (masks[dones] = 0)
        share_obs = np.broadcast_to(obs.reshape(N, 1, -1), (N, M, obs.shape[-1] * M))   # :133-137 (a view)
        h2d += crossing(obs) + crossing(actions) + crossing(logp) + crossing(values) + crossing(rewards) + crossing(masks)
        b.insert(share_obs, obs, None, None, actions, logp, values, rewards, masks)
    runner.compute()
    info = runner.train()                                                # six scalars come back to the host
    d2h += 8 * 4
    return h2d, d2h, info


def _replica_checksum(trainers):
    """int64[4] per rank: order-sensitive checksums of the raw parameter / ValueNorm bits (bit-identical replicas give
    identical numbers)."""
    import torch
    acc = []
    for tr in trainers:
        parts = [tr.policy.actor.flat, tr.policy.critic.flat, tr.policy.actor_optimizer.exp_avg_sq,
                 tr.policy.critic_optimizer.exp_avg_sq]
        if tr.value_normalizer is not None:
            parts.append(tr.value_normalizer.stats)
        bits = torch.cat([p.reshape(-1).view(torch.int32).to(torch.int64) for p in parts])
        w = torch.arange(1, bits.numel() + 1, device=bits.device, dtype=torch.int64)
        acc.append(torch.stack([bits.sum(), (bits * w).sum(), (bits ^ w).sum(), torch.tensor(bits.numel(), device=bits.device)]))
    return torch.stack(acc).sum(0)


def _replicas_identical(trainers, dist, world):
    import torch
    mine = _replica_checksum(trainers)
    allc = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine)
    return all(bool((c == allc[0]).all()) for c in allc)


def dp_check(device, rank, world, dist, n_small=512):
    """Multi-rank parity where the driver runs: `world` ranks x n_small envs (global env ids
    rank * n_small ...) do one rollout + GAE + ONE ppo_update with the NCCL hooks; rank 0 then does the same on ONE GPU
    with the union of the envs (world x n_small, no collective).  Env resets and action sampling are Philox streams
    keyed by the global env id / row, so both see the same trajectories; the advantage moments, the ValueNorm batch
    moments and the gradient sums are global in both.  Reports the distance of the reduced gradient and of the six
    train_info scalars, and whether the replicas' parameters are bit-identical after the update."""
    import torch
    out = {"envs_per_rank": n_small, "ranks": world}

    def one_update(n, rk, wd):
        _, envs, r = build_runner(n, device, rk, wd)
        tr = r.trainer
        tr.ppo_epoch = 1
        r.warmup(); r._rollout(); r.compute()
        info = r.train()
        ws = next(iter(tr._ws.values()))
        return r, tr, ws["grad"].clone(), info

    r, tr, g_multi, info_multi = one_update(n_small, rank, world)
    out["replicas_identical"] = _replicas_identical([tr], dist, world)
    if rank == 0:
        _, tr1, g_single, info_single = one_update(n_small * world, 0, 1)
        scale = float(g_single.abs().max())
        out["grad_vs_single_gpu_max_rel"] = float((g_multi - g_single).abs().max()) / scale
        out["info_vs_single_gpu_max_rel"] = max(abs(info_multi[k] - info_single[k]) / (abs(info_single[k]) + 1e-12)
                                                for k in info_single)
        pa = float((tr.policy.actor.flat - tr1.policy.actor.flat).abs().max())
        pc = float((tr.policy.critic.flat - tr1.policy.critic.flat).abs().max())
        out["params_vs_single_gpu_max_abs"] = max(pa, pc)
        out["tolerance"] = "gradient 1e-5 relative to max|g| (summation order only)"
        out["ok"] = bool(out["replicas_identical"] and out["grad_vs_single_gpu_max_rel"] < 1e-5)
    dist.barrier()
    return out


def run_c1(ns, device):
    """BASELINE config C1 -- the reference's own configuration (128 envs, train_mpe_spread.sh) and exactly what the
    reference arm (`--impl reference`) runs -- on one GPU: the same configuration on both arms.  Timed eagerly and with
    the whole iteration captured as one CUDA graph (--use_cuda_graph); at this size the iteration is launch-bound."""
    import torch
    from mappo_b200.config import parse_args
    from mappo_b200.envs.mpe.MPE_env import MPEVecEnv
    from mappo_b200.runner.shared.mpe_runner import MPERunner
    out = {"workload": "C1: MPE simple_spread M=3 L=3, 128 envs x 1 GPU, T=25, shared MLP (Tanh, H=64), ppo_epoch=10, "
                       "num_mini_batch=1, GAE+ValueNorm (the reference arm's configuration)",
           "agent_steps_per_step": 128 * T_EP * 3, "unit": UNIT}
    for key, extra in (("eager", []), ("cuda_graph", ["--use_cuda_graph"])):
        a = parse_args(flags(128) + extra)
        torch.manual_seed(1)
        envs = MPEVecEnv(a, device=device)
        r = MPERunner({"all_args": a, "envs": envs, "eval_envs": None, "device": device, "num_agents": 3, "run_dir": None})
        r.warmup()
        for _ in range(5):
            r.run_episode()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 50
        e0.record()
        for _ in range(n):
            r.run_episode()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        out[key] = {"ms_per_step": ms, "value": 128 * T_EP * 3 / (ms * 1e-3), "steps": n, "warmup": 5}
    return out


def run_c5(ns, device, rank, world, dist, peaks):
    """The north_star target beside the contract line: BASELINE config C5 -- simple_spread with 10 agents, 16,384 envs
    per GPU (1,310,720 concurrent agents on 8 GPUs), shared MLP policy, centralised critic over 460 columns, NCCL
    gradient all-reduce -- timed like the contract line (device-resident runner, CUDA events, max over ranks), with
    its own per-kernel roofline list.  Returns the dict printed under the key `c5` (rank 0) or None."""
    import torch
    from mappo_b200 import ops
    saved = (M_AGENTS, D_OBS, S_OBS, SCENARIO, ALGO, SEPARATED, PPO_EPOCH, CONFIG)
    set_workload(10, "c5")
    n_envs, steps, warm = ns.c5_envs, max(3, min(ns.steps, 5)), 2
    try:
        _, envs, runner = build_runner(n_envs, device, rank, world)
        runner.warmup()
        for _ in range(warm):
            runner.run_episode()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ops.TIMERS.reset(); ops.TIMERS.enabled = True
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            info = runner.run_episode()
        e1.record()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        ops.TIMERS.enabled = False
        kt = ops.TIMERS.summary()
        if rank != 0:
            return None
        agent_steps = n_envs * T_EP * M_AGENTS
        tr = runner.trainer
        ws0 = next(iter(tr._ws.values()))
        kernels = kernel_rooflines(kt, n_envs, steps, peaks, ws0["nba"], ws0["nbc"], tr.policy.actor.flat.numel(),
                                   tr.policy.critic.flat.numel())
        return {"workload": workload_name(n_envs, world), "value": agent_steps * world * steps / (ms * 1e-3), "unit": UNIT,
                "ms_per_step": ms / steps, "steps": steps, "warmup": warm, "concurrent_agents": n_envs * M_AGENTS * world,
                "agent_steps_per_step": agent_steps * world, "kernels": kernels,
                "kernels_share_of_step": sum(e["ms_per_step"] for e in kernels) / (ms / steps),
                "note": "critic_update spans the wide first layer's three kernels (fc1 GEMM on packed tiles, fused 64-wide "
                        "kernel, fc1 weight-gradient GEMM; DESIGN.md 4.2); the first and last are HBM-bound streams",
                "train_info": info}
    finally:
        globals().update(dict(zip(("M_AGENTS", "D_OBS", "S_OBS", "SCENARIO", "ALGO", "SEPARATED", "PPO_EPOCH", "CONFIG"), saved)))


def run_ours(ns):
    import torch
    from mappo_b200 import _lib, ops
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path); use --impl reference for the CPU arm")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        from mappo_b200 import parallel
        parallel.init_distributed("nccl", device)
    n_envs = ns.envs
    a, envs, runner = build_runner(n_envs, device, rank, world)
    agent_steps = n_envs * T_EP * M_AGENTS

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    dp = None
    if dist is not None and not SEPARATED and not ns.no_dp_check:
        dp = dp_check(device, rank, world, dist)
    if ns.dp_check_only:
        if rank == 0:
            print(json.dumps({"dp_check": dp, "n_gpus": world}))
        if dist is not None:
            dist.destroy_process_group()
        return
    runner.warmup()
    for _ in range(ns.warmup):
        runner.run_episode()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ops.TIMERS.reset(); ops.TIMERS.enabled = True
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(ns.steps):
        info = runner.run_episode()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    launches = _lib.launch_count() - launches0
    ops.TIMERS.enabled = False
    if dp is not None:
        dp["replicas_identical_after_timed_region"] = _replicas_identical(
            runner.trainer if SEPARATED else [runner.trainer], dist, world)
    kt = ops.TIMERS.summary()
    clocks = sampler.stop() if rank == 0 else None
    value = agent_steps * world * ns.steps / (ms * 1e-3)

    if CONFIG not in ("C2", "C5", "custom"):
        # supplementary configs: device-resident figure and kernel split only
        if rank == 0:
            print(json.dumps({"metric": METRIC.replace("simple_spread", SCENARIO), "value": value, "unit": UNIT,
                              "n_gpus": world, "steps": ns.steps, "warmup": ns.warmup, "ms_per_step": ms / ns.steps,
                              "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic",
                              "dtype": "f32 (policy, buffer), f64 (env physics)",
                              "config": {"workload": workload_name(n_envs, world)}, "gpu_launches": int(launches),
                              "kernel_ms_per_step": {k: {"launches": v[0], "ms_per_step": v[1] / ns.steps}
                                                     for k, v in kt.items()},
                              "clocks": clocks, "train_info": info, "supplementary": True}))
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- e2e: host buffers at the reference-facing boundary
    envs.set_host_io(True)
    pin = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt).pin_memory()
    import numpy as _np
    hb = {"onehot": pin(n_envs, M_AGENTS, N_ACT, dt=torch.uint8), "masks": pin(n_envs, M_AGENTS, 1),
          "row5": _np.arange(n_envs * M_AGENTS, dtype=_np.int64) * N_ACT}
    host_boundary_iteration(runner, envs, hb)               # warm-up of the host path
    barrier()
    e_steps = max(1, min(ns.steps, ns.e2e_steps))
    t0 = time.perf_counter()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(e_steps):
        h2d, d2h, _ = host_boundary_iteration(runner, envs, hb)
    g1.record()
    barrier()
    e_ms = torch.tensor([max(g0.elapsed_time(g1), 0.0)], dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    e2e_value = agent_steps * world * e_steps / (float(e_ms.item()) * 1e-3)
    e2e_wall_s = time.perf_counter() - t0                      # the e2e leg alone (the CPU legs run later)
    envs.set_host_io(False)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    c5 = None
    if CONFIG == "C2" and (ns.c5 or world >= 8):
        c5 = run_c5(ns, device, rank, world, dist, peaks)
    c1 = None
    if CONFIG == "C2" and world == 1 and not ns.no_c1 and M_AGENTS == 3:
        c1 = run_c1(ns, device)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (see DESIGN.md "Kernels" for the per-row work)
    rows = agent_steps                                        # one actor_update launch covers the whole batch
    flop_row = actor_update_flops_per_row()
    n, tot = kt.get("actor_update", (0, 0.0))
    per_launch_ms = tot / max(n, 1)
    achieved = rows * flop_row / (per_launch_ms * 1e-3) / 1e12 if n else None
    peak = peaks.get("bf16_tflops_sustained", 1400.0)
    roofline = {"kernel": "mlp_tc_kernel<MODE=2> (actor forward + PPO loss + backward, tcgen05 fp16-pair GEMMs)", "bound": "tensor",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if achieved else None,
                "peak_source": "measured bf16 sustained (MEASURED_PEAKS.json)" if peaks else "fallback",
                "traffic": ACTOR_UPDATE_DRAM_BYTES if (CONFIG == "C2" and n_envs == 65536) else None,
                "traffic_source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch "
                                  "(profiles/r01_update_tc_v5.md); algorithmic 88 B x 4,915,200 rows = 432.5 MB",
                "launches": n, "ms_per_launch": per_launch_ms,
                "flop_per_row": flop_row, "rows_per_launch": rows,
                "note": "algorithmic fp32-equivalent FLOP; the kernel issues 3-4 fp16 MMAs per product and is bound by "
                        "the fp32 epilogues between the GEMMs (see DESIGN.md 4.1, profiles/)"}
    share = {k: {"launches": v[0], "ms_per_step": v[1] / ns.steps} for k, v in kt.items()}
    tr0 = runner.trainer[0] if SEPARATED else runner.trainer
    ws0 = next(iter(tr0._ws.values()))
    roofline["kernels"] = kernel_rooflines(kt, n_envs, ns.steps, peaks, ws0["nba"], ws0["nbc"],
                                           tr0.policy.actor.flat.numel(), tr0.policy.critic.flat.numel())
    timed = sum(e["ms_per_step"] for e in roofline["kernels"])
    roofline["kernels_ms_per_step"] = timed
    roofline["kernels_share_of_step"] = timed / (ms / ns.steps)

    # ---- CPU baseline beside it (N = 1 only): the unmodified reference (child process: its 128 env workers are
    # forked from a process that never initialised CUDA), and the CPU port as a second figure
    cpu = cpu_port = None
    if world == 1 and not ns.no_cpu:
        cores = len(os.sched_getaffinity(0))
        try:
            cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", str(ns.ref_iters),
                   "--warmup", "1", "--ref-envs", str(ns.ref_envs), "--ref-threads", str(ns.ref_threads),
                   "--agents", str(M_AGENTS)]
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=900)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
            cpu = json.loads(line)["cpu_baseline"]
        except Exception as ex:                                  # noqa: BLE001 -- the figure is reported, never fatal
            cpu = {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": "failed: %r" % (ex,)}
        if cpu.get("kind") == "reference" and not ns.no_port:
            v, sec = cpu_iteration_rate(ns.cpu_envs, ns.cpu_iters, 1, cores)
            cpu_port = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": "%d envs x %d full iterations after 1 warm-up (%.2f s each) of the same workload, "
                                  "oracle/iteration.py (numpy env batched over envs + torch CPU autograd), torch %d threads"
                                  % (ns.cpu_envs, ns.cpu_iters, sec, cores)}

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": ns.steps, "warmup": ns.warmup,
           "ms_per_step": ms / ns.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32 (policy, buffer), f64 (env physics)", "data": "synthetic",
           "config": {"workload": workload_name(n_envs, world), "agent_steps_per_step": agent_steps * world,
                      "l2": "inputs larger than L2 (rollout buffer 0.4 GB per GPU)"},
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "steps": e_steps, "wall_s": e2e_wall_s},
           "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "cpu_baseline_port": cpu_port,
           "kernel_ms_per_step": share,
           "clocks": clocks, "train_info": info}
    if c1 is not None:
        out["c1"] = c1
    if c5 is not None:
        out["c5"] = c5
    if dp is not None:
        out["dp_check"] = dp
        out["collective"] = ("NCCL all-reduce (torch.distributed): ONE float64 buffer [actor grad sums | critic grad sums | "
                             "8 loss sums] per ppo_update; advantage moments and ValueNorm batch moments (3 doubles each) "
                             "once per iteration")
    print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def kernel_rooflines(kt, n_envs, steps, peaks, nblocks_a, nblocks_c, P_a, P_c):
    """Per-kernel roofline entries for the bench line (north_star: "each kernel at a stated fraction of its HBM or
    tensor roofline").  Times: CUDA events around every launch inside the timed region (ops.TIMERS, on the launching
    stream).  Algorithmic work: SURVEY.md section 8(d) formulas (fp32 storage; FLOP = 2 x MAC, training pass = 3 x
    forward minus the first-layer data gradient).  Denominators: MEASURED_PEAKS.json (HBM copy bandwidth; bf16 tensor
    sustained, since every kernel is timed inside a long step).  `traffic`: dram__bytes_read.sum + dram__bytes_write.sum
    of one launch from the committed `ncu --set full` summaries (profiles/r02_traffic.json), C2 shape only."""
    M, L, D, S, A, T = M_AGENTS, N_LANDMARKS, D_OBS, S_OBS, N_ACT, T_EP
    hbm = peaks.get("hbm_gbs", 6650.0)
    tens = peaks.get("bf16_tflops_sustained", 1400.0)
    src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"
    rows_a, rows_c = n_envs * M, n_envs                      # per rollout step: actor rows, unique share_obs rows
    B = n_envs * T * M
    fwd_a = 2 * (D * H + H * H + H * A)
    fwd_c = 2 * (S * H + H * H + H)
    env_bytes = 4 * (4 * M + 2 * L + M + 1) + 4 * (4 * M + M * D + M) + M      # per env, SURVEY 8(d): 367 B at C2
    gae_bytes_traj = (T + 2 * (T + 1)) * 4 + 2 * T * 4 + 4                    # per (env, agent) trajectory: 512 B
    traffic = {}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if CONFIG == "C2" and n_envs == tj.get("n_envs"):
            traffic = tj.get("kernels", {})
    except Exception:
        pass
    spec = [
        ("actor_update", "mlp_tc_kernel<MODE 2> actor forward + PPO loss + backward", "tensor",
         B * (3 * fwd_a - 2 * D * H), "%d rows x (3 x %d - 2 D H) FLOP" % (B, fwd_a)),
        ("critic_update", "mlp_tc_kernel<MODE 3> critic forward + value loss + backward (one row per env-step, group = M)",
         "tensor", n_envs * T * (3 * fwd_c - 2 * S * H), "%d rows x (3 x %d - 2 S H) FLOP" % (n_envs * T, fwd_c)),
        ("actor_forward", "mlp_tc_kernel<MODE 0> rollout actor forward + sampling", "tensor", rows_a * fwd_a,
         "%d rows x %d FLOP" % (rows_a, fwd_a)),
        ("critic_forward", "mlp_tc_kernel<MODE 1> rollout critic forward", "tensor", rows_c * fwd_c,
         "%d rows x %d FLOP" % (rows_c, fwd_c)),
        ("env_step", "env_kernel<simple_spread> physics + reward + obs + auto-reset", "hbm", n_envs * env_bytes,
         "%d envs x %d B" % (n_envs, env_bytes)),
        ("gae_returns", "gae_kernel backward scan + ValueNorm denormalise + raw advantages", "hbm",
         n_envs * M * gae_bytes_traj, "%d trajectories x %d B" % (n_envs * M, gae_bytes_traj)),
        ("grad_reduce", "grad_reduce_kernel (fixed-order sum of the per-CTA partial rows)", "hbm", None, None),
        ("adam_step", "adam_step_kernel (affine chain rule + global-norm clip + Adam)", "hbm", None, None),
    ]
    out = []
    for key, name, bound, work, formula in spec:
        n, tot = kt.get(key, (0, 0.0))
        if n == 0:
            continue
        ms = tot / n
        if key == "grad_reduce":      # two launches per update: actor and critic partials, read once, sums written
            work = ((nblocks_a + 1) * P_a + (nblocks_c + 1) * P_c) * 4 // 2
            formula = "mean of (%d + 1) x %d and (%d + 1) x %d floats" % (nblocks_a, P_a, nblocks_c, P_c)
        if key == "adam_step":        # params + m + v read and written, sums read, gradient written
            work = (P_a + P_c) * 4 * 8 // 2
            formula = "mean of 8 x 4 B x %d and x %d parameters" % (P_a, P_c)
        e = {"key": key, "kernel": name, "bound": bound, "launches_per_step": n / steps, "ms_per_launch": ms,
             "ms_per_step": tot / steps, "work_per_launch": work, "work_formula": formula,
             "traffic": traffic.get(key)}
        if bound == "tensor":
            ach = work / (ms * 1e-3) / 1e12
            e.update(achieved=ach, peak=tens, unit="TFLOP/s", frac=ach / tens,
                     frac_of_fp32_fma_peak=ach / 74.4)
        else:
            ach = work / (ms * 1e-3) / 1e9
            e.update(achieved=ach, peak=hbm, unit="GB/s", frac=ach / hbm)
        e["peak_source"] = src
        out.append(e)
    return out


def actor_update_flops_per_row():
    """Algorithmic FLOP of forward + dgrad + wgrad per sample for the actor (MACs x 2 x 3 minus the
    first-layer dgrad)."""
    macs = D_OBS * H + H * H + H * N_ACT
    return 2 * macs * 3 - 2 * D_OBS * H


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU")
    ap.add_argument("--agents", type=int, default=3, help="agents per env (3 = C2, the headline; 10 = C5 shape)")
    ap.add_argument("--config", default=None, choices=["c3", "c3r", "c4", "c5"], help="supplementary BASELINE configs")
    ap.add_argument("--cpu-envs", dest="cpu_envs", type=int, default=1024,
                    help="envs of the bounded CPU sample (the port's throughput peaks near 1,024 envs per iteration: "
                         "58 k agent-steps/s on 8 cores against 40 k at 8,192)")
    ap.add_argument("--cpu-iters", dest="cpu_iters", type=int, default=10,
                    help="timed iterations of the port's leg (about 5 s on 16 cores)")
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=5)
    ap.add_argument("--no-cpu", dest="no_cpu", action="store_true")
    ap.add_argument("--c5", action="store_true", help="also measure BASELINE config C5 (10 agents) and print it under the "
                                                      "key `c5`; on by default at 8 GPUs (the north_star target)")
    ap.add_argument("--c5-envs", dest="c5_envs", type=int, default=16384, help="envs per GPU of the C5 measurement")
    ap.add_argument("--no-c1", dest="no_c1", action="store_true", help="skip the C1 (128 envs) figures of the `c1` key")
    ap.add_argument("--no-dp-check", dest="no_dp_check", action="store_true")
    ap.add_argument("--dp-check-only", dest="dp_check_only", action="store_true",
                    help="under torchrun: run the multi-rank parity check, print it and exit (tests/test_gpu_dist.py)")
    ap.add_argument("--no-port", dest="no_port", action="store_true", help="skip the second (port) CPU figure")
    ap.add_argument("--ref-envs", dest="ref_envs", type=int, default=128,
                    help="SubprocVecEnv workers of the unmodified reference (C1: 128, train_mpe_spread.sh)")
    ap.add_argument("--ref-threads", dest="ref_threads", type=int, default=1,
                    help="n_training_threads of the reference arm: 1 = the reference's own default (config.py:168, "
                         "train_mpe_spread.sh) and the faster setting on the B200 box's 16 cores, where the 128 env worker "
                         "processes already occupy every core (measured: 23.0 k agent-steps/s with 1, 15.6 k with 16); "
                         "0 = all host cores")
    ap.add_argument("--ref-kind", dest="ref_kind", default="auto", choices=["auto", "port"],
                    help="auto: the unmodified reference from baseline/_ref when present, else the port")
    ap.add_argument("--ref-iters", dest="ref_iters", type=int, default=5,
                    help="timed iterations of the unmodified reference in the cpu_baseline leg of the GPU arm")
    ns = ap.parse_args()
    set_workload(ns.agents, ns.config)
    if ns.config in ("c3", "c3r", "c4") and ns.envs == 65536:
        ns.envs = 32768
    ns.warmup = max(ns.warmup, 0)
    if ns.impl == "reference":
        run_reference(ns)
    else:
        run_ours(ns)


if __name__ == "__main__":
    main()
