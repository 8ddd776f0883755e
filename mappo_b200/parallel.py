"""Data-parallel plumbing for 1..8 GPUs of one box: one process per GPU, `torch.distributed` (NCCL over NVLink on
the GPUs; gloo in the CPU tests).

The reference has no counterpart (single device).  What shards and what is exchanged:
  * envs: rank r owns the global env ids [r*N, (r+1)*N); reset draws are Philox streams keyed by the GLOBAL env id, so
    a rollout does not depend on how the envs are sharded; no exchange during rollout, insert or GAE;
  * update: parameters and Adam state are replicated; per `ppo_update` the ranks sum (all-reduce) ONE flat float64
    buffer [actor grads | critic grads | 8 loss sums / active counts] of UNNORMALISED sums (written by
    mappo_grad_reduce_f64), then every rank divides by the global count, clips and steps identically
    (mappo_adam_step_f64);
  * statistics: the advantage moments (double[3]: sum, sum of squares, count) once per iteration and the ValueNorm
    batch moments (double[3]) once per `ppo_update`.
"""
import os

import torch
import torch.distributed as dist


def env_shard(n_envs_per_rank, rank):
    """Global id of this rank's env 0 (mappo_env_config.env_id_offset)."""
    return int(rank) * int(n_envs_per_rank)


def init_distributed(backend=None, device=None):
    """Join the job described by RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT (torchrun).  Returns (rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl" and device is not None:
            kw["device_id"] = device
        dist.init_process_group(backend, **kw)
    return rank, world


def sum_all_reduce():
    """callable(tensor): in-place sum over ranks; what R_MAPPO.set_all_reduce / ValueNorm.all_reduce expect."""
    def fn(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t
    return fn


def attach(trainer, world):
    """Make `trainer` (R_MAPPO) data-parallel over the initialised process group."""
    if world > 1:
        trainer.set_all_reduce(sum_all_reduce(), world)
    return trainer


def global_moments(stats3):
    """(mean, population std) from an all-reduced double[3] = (sum, sum of squares, count): np.nanmean / np.nanstd of
    the union of the ranks' samples (r_mappo.py:180-182)."""
    s, ss, n = float(stats3[0]), float(stats3[1]), float(stats3[2])
    mean = s / n
    var = max(ss / n - mean * mean, 0.0)
    return mean, var ** 0.5
