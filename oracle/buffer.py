"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's replay-buffer maths.

  ValueNorm                      onpolicy/utils/valuenorm.py:8-78
  SharedReplayBuffer.compute_returns   onpolicy/utils/shared_buffer.py:168-224   (all four branches)
  advantage prologue             onpolicy/algorithms/r_mappo/r_mappo.py:174-182
  feed_forward_generator         onpolicy/utils/shared_buffer.py:226-286
  recurrent_generator            onpolicy/utils/shared_buffer.py:385-494
  naive_recurrent_generator      onpolicy/utils/shared_buffer.py:288-383
  SeparatedReplayBuffer generators   onpolicy/utils/separated_buffer.py:162-393

float32 numpy with the reference's operation order.  Pinned against fixtures recorded from the
reference (tests/golden/gae_*.npz, gen_*.npz) by tests/test_oracle_buffer.py.
"""
import numpy as np

f32 = np.float32


class OracleValueNorm:
    """State = (running_mean, running_mean_sq, debiasing_term), float32 scalars (valuenorm.py:19-21)."""

    def __init__(self, beta=0.99999, epsilon=1e-5):
        self.beta, self.epsilon = beta, epsilon
        self.running_mean = f32(0.0)
        self.running_mean_sq = f32(0.0)
        self.debiasing_term = f32(0.0)

    def stats(self):
        return np.array([self.running_mean, self.running_mean_sq, self.debiasing_term], dtype=f32)

    def load(self, stats):
        self.running_mean, self.running_mean_sq, self.debiasing_term = [f32(x) for x in stats]

    def running_mean_var(self):                                  # valuenorm.py:31-35
        db = max(self.debiasing_term, f32(self.epsilon))
        mean = f32(self.running_mean / db)
        mean_sq = f32(self.running_mean_sq / db)
        var = max(f32(mean_sq - f32(mean * mean)), f32(1e-2))
        return mean, var

    def update(self, x):                                         # valuenorm.py:38-54 (per_element_update=False)
        x = np.asarray(x, dtype=f32)
        batch_mean = f32(np.mean(x, dtype=np.float64))
        batch_sq_mean = f32(np.mean((x * x).astype(f32), dtype=np.float64))
        w, omw = f32(self.beta), f32(1.0 - self.beta)
        self.running_mean = f32(f32(self.running_mean * w) + f32(batch_mean * omw))
        self.running_mean_sq = f32(f32(self.running_mean_sq * w) + f32(batch_sq_mean * omw))
        self.debiasing_term = f32(f32(self.debiasing_term * w) + omw)

    def normalize(self, x):                                      # valuenorm.py:56-65
        mean, var = self.running_mean_var()
        return ((np.asarray(x, dtype=f32) - mean) / np.sqrt(var, dtype=f32)).astype(f32)

    def denormalize(self, x):                                    # valuenorm.py:67-78
        mean, var = self.running_mean_var()
        return (np.asarray(x, dtype=f32) * np.sqrt(var, dtype=f32) + mean).astype(f32)


def compute_returns(rewards, value_preds, masks, next_value, gamma, gae_lambda, use_gae=True,
                    use_proper_time_limits=False, bad_masks=None, normalizer=None):
    """Arrays are [T(+1), ...] float32 with any trailing shape.  Returns (value_preds, returns) copies.
    shared_buffer.py:168-224, branch for branch."""
    rewards = np.asarray(rewards, dtype=f32)
    value_preds = np.array(value_preds, dtype=f32)
    masks = np.asarray(masks, dtype=f32)
    returns = np.zeros_like(value_preds)
    T = rewards.shape[0]
    if bad_masks is None:
        bad_masks = np.ones_like(masks)
    den = (lambda v: normalizer.denormalize(v)) if normalizer is not None else (lambda v: v)
    if use_gae:
        value_preds[-1] = next_value
        gae = 0
        for step in reversed(range(T)):
            delta = rewards[step] + gamma * den(value_preds[step + 1]) * masks[step + 1] - den(value_preds[step])
            if use_proper_time_limits:
                gae = delta + gamma * gae_lambda * gae * masks[step + 1]
                gae = gae * bad_masks[step + 1]
            else:
                gae = delta + gamma * gae_lambda * masks[step + 1] * gae
            returns[step] = gae + den(value_preds[step])
    else:
        returns[-1] = next_value
        for step in reversed(range(T)):
            if use_proper_time_limits:
                returns[step] = (returns[step + 1] * gamma * masks[step + 1] + rewards[step]) * bad_masks[step + 1] \
                    + (1 - bad_masks[step + 1]) * den(value_preds[step])
            else:
                returns[step] = returns[step + 1] * gamma * masks[step + 1] + rewards[step]
    return value_preds, returns


def advantages(returns, value_preds, active_masks=None, normalizer=None):
    """r_mappo.py:174-182: raw and normalised advantages."""
    den = normalizer.denormalize(value_preds[:-1]) if normalizer is not None else value_preds[:-1]
    adv = (returns[:-1] - den).astype(f32)
    cp = adv.copy()
    if active_masks is not None:
        cp[active_masks[:-1] == 0.0] = np.nan
    mean, std = np.nanmean(cp), np.nanstd(cp)
    return adv, ((adv - mean) / (std + 1e-5)).astype(f32)


# ------------------------------------------------------------------ minibatch index logic
def feed_forward_rows(T, N, M, perm, num_mini_batch):
    """shared_buffer.py:226-247: flat row i of the [T,N,M] arrays is (t*N + n)*M + m; minibatch k is
    perm[k*mbs:(k+1)*mbs] with mbs = B // num_mini_batch (tail dropped).  Returns list of int64 arrays."""
    B = T * N * M
    mbs = B // num_mini_batch
    perm = np.asarray(perm, dtype=np.int64)
    return [perm[k * mbs:(k + 1) * mbs] for k in range(num_mini_batch)]


def cast_rows(T, N, M):
    """Row map of `_cast` (shared_buffer.py:10-11): position p in the [N,M,T]-ordered stream -> flat
    (t*N + n)*M + m row of the time-major arrays."""
    n, m, t = np.meshgrid(np.arange(N), np.arange(M), np.arange(T), indexing="ij")
    return ((t * N + n) * M + m).reshape(-1).astype(np.int64)


def recurrent_rows(T, N, M, perm, num_mini_batch, L):
    """shared_buffer.py:385-494: chunk c covers stream positions [cL, cL+L) of the [N,M,T]-ordered data
    (straddling trajectories when L does not divide T); a minibatch stacks its chunks on axis 1 and
    flattens time-major.  Returns per minibatch (rows [L*Nc] int64 into the time-major flat arrays,
    rnn_rows [Nc] int64 = the row whose rnn state seeds each chunk)."""
    B = T * N * M
    chunks = B // L
    mbs = chunks // num_mini_batch
    stream = cast_rows(T, N, M)
    perm = np.asarray(perm, dtype=np.int64)
    out = []
    for k in range(num_mini_batch):
        idx = perm[k * mbs:(k + 1) * mbs]
        pos = idx[None, :] * L + np.arange(L)[:, None]          # [L, Nc] stream positions, time-major
        out.append((stream[pos].reshape(-1), stream[idx * L]))
    return out


def separated_recurrent_rows(T, N, perm, num_mini_batch, L):
    """separated_buffer.py:289-393: data is _cast to [N,T] order (transpose(1,0,2)), chunks as above, but the
    minibatch is stacked as (Nc, L) on axis 0 and then flattened AS IF it were (L, Nc) (:361-391) --
    reproduced as is.  Returns per minibatch (rows [L*Nc] into the [T,N] time-major flat arrays, rnn_rows)."""
    B = T * N
    chunks = B // L
    mbs = chunks // num_mini_batch
    n, t = np.meshgrid(np.arange(N), np.arange(T), indexing="ij")
    stream = (t * N + n).reshape(-1).astype(np.int64)
    perm = np.asarray(perm, dtype=np.int64)
    out = []
    for k in range(num_mini_batch):
        idx = perm[k * mbs:(k + 1) * mbs]
        pos = idx[:, None] * L + np.arange(L)[None, :]          # [Nc, L]: np.stack(axis=0) of L-row chunks
        out.append((stream[pos].reshape(-1), stream[idx * L]))   # _flatten(L, N, x) is a plain reshape
    return out


def naive_recurrent_rows(T, R, perm, num_mini_batch):
    """shared_buffer.py:288-383 (R = N*M) and separated_buffer.py:215-287 (R = N): whole trajectories,
    minibatch k = perm[k*Nb:(k+1)*Nb]; rows are time-major (T, Nb) into the flat [T, R] arrays; the rnn
    state comes from t = 0 of each trajectory.  Returns per minibatch (rows [T*Nb], rnn_rows [Nb])."""
    nb = R // num_mini_batch
    perm = np.asarray(perm, dtype=np.int64)
    out = []
    for start in range(0, R, nb):
        idx = perm[start:start + nb]
        if len(idx) < nb:
            break
        rows = (np.arange(T)[:, None] * R + idx[None, :]).reshape(-1)
        out.append((rows, idx.copy()))
    return out
