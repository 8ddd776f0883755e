"""ORACLE (test infrastructure, not product code): Philox4x32-10 counter-based RNG in numpy.

The reference draws reset states from the process-global MT19937 `np.random`
(onpolicy/envs/mpe/scenarios/simple_spread.py:38-45, seeded in scripts/train/train_mpe.py:25),
which a device-resident batched env cannot reproduce stream-for-stream.
The build therefore defines its own reset stream -- Philox4x32-10 (Salmon et al., SC'11; the
published algorithm, same constants as Random123/cuRAND) keyed by (seed, stream) with
counter = (global env id, episode index, draw block, 0) -- and this file is the CPU statement of it
that the CUDA kernel (mappo_b200/csrc/philox.cuh) is checked against bit-for-bit.
Known-answer vectors from the Random123 distribution (kat_vectors) are pinned in tests/test_oracle_philox.py.
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1, rounds=10):
    """Vectorised Philox4x32-R. All inputs broadcastable uint32 arrays / ints. Returns 4 uint32 arrays."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & _MASK for c in np.broadcast_arrays(c0, c1, c2, c3)]
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(rounds):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def u32_to_unit(x):
    """uint32 -> float64 in the open interval (0,1): (x + 0.5) * 2**-32 (exact in f64)."""
    return (np.asarray(x, dtype=np.float64) + 0.5) * (1.0 / 4294967296.0)


def u32_to_sym(x):
    """uint32 -> float64 uniform in (-1, 1): 2u - 1."""
    return 2.0 * u32_to_unit(x) - 1.0


def draws(env_id, episode, n_draws, seed, stream):
    """`n_draws` uint32 per env: draw i comes from block i//4, lane i%4 of
    philox(counter=(env_id, episode, block, 0), key=(seed, stream)).  Returns [n_env, n_draws] uint32."""
    env_id = np.asarray(env_id, dtype=np.uint64).reshape(-1)
    episode = np.broadcast_to(np.asarray(episode, dtype=np.uint64).reshape(-1), env_id.shape)
    out = np.empty((env_id.shape[0], n_draws), dtype=np.uint32)
    nblk = (n_draws + 3) // 4
    for b in range(nblk):
        r = philox4x32(env_id, episode, np.uint64(b), np.uint64(0), seed, stream)
        for lane in range(4):
            i = 4 * b + lane
            if i < n_draws:
                out[:, i] = r[lane]
    return out
