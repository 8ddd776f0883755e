"""CPU, world_size 2 over gloo: the host side of the multi-GPU path (mappo_b200/parallel.py) against the oracle.
  * env sharding: rank r steps envs [r*N, (r+1)*N) with Philox streams keyed by the global env id -> the
    concatenation of the shards' rollouts equals the single-process rollout bit for bit;
  * advantage moments: all-reduced (sum, sumsq, count) reproduce np.nanmean / np.nanstd of the whole batch;
  * gradients: all-reduced UNNORMALISED gradient sums divided by the global count equal the full-batch gradient
    (what R_MAPPO._update does with mappo_grad_reduce -> all_reduce -> mappo_adam_step)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from mappo_b200 import parallel
    from oracle.mpe import OracleMPE
    from oracle import policy as op
    r, w = parallel.init_distributed("gloo")
    assert (r, w) == (rank, world)
    allred = parallel.sum_all_reduce()
    N, M, L, T = 6, 3, 3, 5
    env = OracleMPE("simple_spread", N, M, L, T, seed=7, env_id_offset=parallel.env_shard(N, rank))
    obs = [np.stack(env.reset(), 1)]
    rng = np.random.RandomState(0)
    acts_all = rng.randint(0, 5, (T + 2, world * N, M, 1))
    rews = []
    for t in range(T + 2):                     # crosses one auto-reset
        o, rw, d, _ = env.step(acts_all[t, rank * N:(rank + 1) * N])
        obs.append(np.stack(o, 1)); rews.append(rw)
    # moments of a per-rank "advantage" array
    adv = np.concatenate(rews, 0).astype(np.float32).reshape(-1)
    s3 = torch.tensor([adv.astype(np.float64).sum(), (adv.astype(np.float64) ** 2).sum(), adv.size], dtype=torch.float64)
    allred(s3)
    # gradient of a mean loss over the union of the shards, from per-rank unnormalised sums
    torch.manual_seed(0)
    W = torch.randn(4, 18, requires_grad=True)
    x = torch.as_tensor(np.concatenate(obs, 0).reshape(-1, 18), dtype=torch.float32)
    loss_sum = (torch.tanh(x @ W.t()) ** 2).sum()
    g = torch.autograd.grad(loss_sum, W)[0].reshape(-1)
    cnt = torch.tensor([float(x.shape[0])], dtype=torch.float64)
    allred(g); allred(cnt)
    torch.save(dict(obs=np.stack(obs), rews=np.stack(rews), s3=s3, g=g / cnt.item().__float__(), W=W.detach()),
               os.path.join(out, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_reductions(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sys.path.insert(0, ROOT)
    from mappo_b200 import parallel
    from oracle.mpe import OracleMPE
    res = [torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r), weights_only=False) for r in range(world)]
    N, M, L, T = 6, 3, 3, 5
    env = OracleMPE("simple_spread", world * N, M, L, T, seed=7)
    obs = [np.stack(env.reset(), 1)]
    rng = np.random.RandomState(0)
    acts_all = rng.randint(0, 5, (T + 2, world * N, M, 1))
    rews = []
    for t in range(T + 2):
        o, rw, d, _ = env.step(acts_all[t])
        obs.append(np.stack(o, 1)); rews.append(rw)
    obs, rews = np.stack(obs), np.stack(rews)
    np.testing.assert_array_equal(np.concatenate([r["obs"] for r in res], axis=1), obs)       # bit-exact sharding
    np.testing.assert_array_equal(np.concatenate([r["rews"] for r in res], axis=1), rews)
    adv = rews.astype(np.float32).reshape(-1)
    mean, std = parallel.global_moments(res[0]["s3"])
    np.testing.assert_allclose(mean, np.nanmean(adv), rtol=1e-6)
    np.testing.assert_allclose(std, np.nanstd(adv), rtol=1e-5)
    assert torch.equal(res[0]["s3"], res[1]["s3"])
    W = res[0]["W"].clone().requires_grad_(True)
    x = torch.as_tensor(np.concatenate([r["obs"].reshape(-1, 18) for r in res], 0), dtype=torch.float32)
    full = torch.autograd.grad((torch.tanh(x @ W.t()) ** 2).mean(0).sum(), W)[0].reshape(-1)
    np.testing.assert_allclose(res[0]["g"].numpy(), full.numpy(), rtol=1e-5, atol=1e-7)
    assert torch.equal(res[0]["g"], res[1]["g"])                                              # replicas stay identical
