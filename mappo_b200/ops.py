"""Thin Python wrappers over the C ABI (include/mappo_b200.h): argument checking and pointer plumbing only.
Every function launches CUDA kernels on torch's current stream; none synchronises."""
import ctypes as C

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


class GaeConfig(C.Structure):
    _fields_ = [("T", C.c_int32), ("R", C.c_int32), ("gamma", C.c_float), ("gamma_lambda", C.c_float),
                ("use_gae", C.c_int32), ("use_proper_time_limits", C.c_int32), ("use_value_norm", C.c_int32),
                ("reserved", C.c_int32)]


class KernelTimers:
    """Per-kernel CUDA-event timing on the launching stream (bench.py's roofline figures).  Disabled by default:
    `span` then costs one attribute test."""

    def __init__(self):
        self.enabled = False
        self._ev = {}

    class _Span:
        def __init__(self, owner, name):
            self.o, self.n = owner, name

        def __enter__(self):
            if self.o.enabled:
                self.a = torch.cuda.Event(enable_timing=True)
                self.b = torch.cuda.Event(enable_timing=True)
                self.a.record()
            return self

        def __exit__(self, *exc):
            if self.o.enabled:
                self.b.record()
                self.o._ev.setdefault(self.n, []).append((self.a, self.b))
            return False

    def span(self, name):
        return KernelTimers._Span(self, name)

    def reset(self):
        self._ev = {}

    def summary(self):
        """{name: (launches, total_ms)}; synchronises."""
        torch.cuda.synchronize()
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in self._ev.items()}


TIMERS = KernelTimers()

_P = C.c_void_p
_lib._EXTRA_SIGS.update({
    "mappo_gae_returns": [C.POINTER(GaeConfig), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "mappo_adv_stats_reduce": [_P, C.c_int, _P, _P],
    "mappo_adv_normalize": [_P, _P, C.c_longlong, _P, _P],
    "mappo_moments": [_P, C.c_longlong, _P, _P],
    "mappo_valuenorm_apply": [_P, _P, C.c_double, _P],
    "mappo_valuenorm_transform": [_P, _P, _P, C.c_longlong, C.c_int, _P],
    "mappo_popart_rescale": [_P, _P, _P, _P, C.c_int, C.c_int, _P],
})


def _f32c(t, name):
    assert t.dtype == torch.float32 and t.is_cuda and t.is_contiguous(), "%s must be contiguous float32 CUDA" % name
    return t


def gae_returns(rewards, value_preds, returns, masks, next_value, gamma, gae_lambda, use_gae=True,
                use_proper_time_limits=False, bad_masks=None, active_masks=None, vn_stats=None,
                advantages_raw=None, stats3=None, all_reduce=None):
    """compute_returns + raw advantages + advantage moments.  Arrays [T(+1), ...] with any trailing shape.
    Returns (advantages_raw [T,...], stats3 double[3] = {sum, sumsq, count})."""
    L = _lib.lib()
    T = rewards.shape[0]
    R = rewards[0].numel()
    assert value_preds.shape[0] == T + 1 and returns.shape[0] == T + 1 and masks.shape[0] == T + 1
    dev = rewards.device
    if advantages_raw is None:
        advantages_raw = torch.empty_like(rewards)
    nparts = L.mappo_gae_num_partials(R)
    partials = torch.empty(nparts, 3, dtype=torch.float64, device=dev)
    if stats3 is None:
        stats3 = torch.empty(3, dtype=torch.float64, device=dev)
    cfg = GaeConfig(T, R, float(gamma), float(gamma * gae_lambda), int(use_gae), int(use_proper_time_limits),
                    int(vn_stats is not None), 0)
    with TIMERS.span("gae_returns"):
        check(L.mappo_gae_returns(C.byref(cfg), ptr(_f32c(rewards, "rewards")), ptr(_f32c(value_preds, "value_preds")),
                                  ptr(_f32c(returns, "returns")), ptr(_f32c(masks, "masks")),
                                  ptr(bad_masks), ptr(active_masks), ptr(next_value), ptr(vn_stats),
                                  ptr(advantages_raw), ptr(partials), stream_ptr()), "gae_returns")
    check(L.mappo_adv_stats_reduce(ptr(partials), nparts, ptr(stats3), stream_ptr()), "adv_stats_reduce")
    if all_reduce is not None:
        all_reduce(stats3)
    return advantages_raw, stats3


def adv_normalize(advantages_raw, stats3, out=None):
    if out is None:
        out = torch.empty_like(advantages_raw)
    check(_lib.lib().mappo_adv_normalize(ptr(advantages_raw), ptr(out), advantages_raw.numel(), ptr(stats3),
                                         stream_ptr()), "adv_normalize")
    return out


def moments(x, stats3=None, all_reduce=None):
    """double[3] = {sum x, sum x^2 (x^2 rounded to f32 first), n} over a contiguous float32 tensor."""
    L = _lib.lib()
    n = x.numel()
    nparts = L.mappo_moments_num_partials(n)
    partials = torch.empty(nparts, 3, dtype=torch.float64, device=x.device)
    if stats3 is None:
        stats3 = torch.empty(3, dtype=torch.float64, device=x.device)
    check(L.mappo_moments(ptr(_f32c(x, "x")), n, ptr(partials), stream_ptr()), "moments")
    check(L.mappo_adv_stats_reduce(ptr(partials), nparts, ptr(stats3), stream_ptr()), "stats_reduce")
    if all_reduce is not None:
        all_reduce(stats3)
    return stats3


def valuenorm_apply(vn_stats, stats3, beta):
    check(_lib.lib().mappo_valuenorm_apply(ptr(vn_stats), ptr(stats3), float(beta), stream_ptr()), "valuenorm_apply")


def popart_rescale(old_stats, new_stats, head_w, head_b):
    """PopArt.update's rescale of the value head (popart.py:69-70); head_w [n_out, n_in], head_b [n_out] in place."""
    check(_lib.lib().mappo_popart_rescale(ptr(old_stats), ptr(new_stats), ptr(head_w), ptr(head_b),
                                          int(head_w.shape[1]), int(head_w.shape[0]), stream_ptr()), "popart_rescale")


def valuenorm_transform(vn_stats, x, denormalize, out=None):
    if out is None:
        out = torch.empty_like(x)
    check(_lib.lib().mappo_valuenorm_transform(ptr(vn_stats), ptr(_f32c(x, "x")), ptr(out), x.numel(),
                                               int(denormalize), stream_ptr()), "valuenorm_transform")
    return out


# ---------------------------------------------------------------------------------------------------
# networks
class NetConfig(C.Structure):
    _fields_ = [("in_dim", C.c_int32), ("hidden", C.c_int32), ("layer_N", C.c_int32), ("use_relu", C.c_int32),
                ("use_feature_norm", C.c_int32), ("recurrent", C.c_int32), ("n_heads", C.c_int32),
                ("head_dim", C.c_int32 * 4), ("is_critic", C.c_int32)]


class LossConfig(C.Structure):
    _fields_ = [("clip_param", C.c_float), ("entropy_coef", C.c_float), ("value_loss_coef", C.c_float),
                ("huber_delta", C.c_float), ("use_huber_loss", C.c_int32), ("use_clipped_value_loss", C.c_int32),
                ("use_value_norm", C.c_int32), ("use_policy_active_masks", C.c_int32),
                ("use_value_active_masks", C.c_int32), ("reserved", C.c_int32)]


class AdamConfig(C.Structure):
    _fields_ = [("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("weight_decay", C.c_float), ("max_grad_norm", C.c_float), ("use_max_grad_norm", C.c_int32),
                ("bias_correction1", C.c_float), ("bias_correction2_sqrt", C.c_float), ("one_minus_beta1", C.c_float),
                ("one_minus_beta2", C.c_float), ("beta1_exact", C.c_double), ("beta2_exact", C.c_double),
                ("frozen_head", C.c_int32), ("reserved2", C.c_int32)]


_NC, _LC, _AC = C.POINTER(NetConfig), C.POINTER(LossConfig), C.POINTER(AdamConfig)
_lib._EXTRA_SIGS.update({
    "mappo_actor_forward": [_NC, _P, _P, C.c_int, C.c_longlong, _P, C.c_int, C.c_uint64, C.c_uint64, C.c_longlong,
                            _P, _P, _P, _P, _P, _P, _P],
    "mappo_advantages": [_P, _P, _P, _P, C.c_longlong, _P, _P, _P],
    "mappo_critic_forward": [_NC, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, _P, _P],
    "mappo_actor_update": [_NC, _LC, _P, _P, C.c_int, C.c_longlong, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "mappo_critic_update": [_NC, _LC, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P,
                            _P],
    "mappo_critic_pack_inputs": [_NC, _P, C.c_int, C.c_longlong, _P, _P, _P],
    "mappo_grad_reduce": [_P, _P, C.c_int, C.c_longlong, _P, _P, _P],
    "mappo_grad_reduce_f64": [_P, _P, C.c_int, C.c_longlong, _P, _P, _P],
    "mappo_adam_step": [_NC, _AC, _P, _P, _P, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, _P, _P, _P],
    "mappo_adam_step_f64": [_NC, _AC, _P, _P, _P, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, _P, _P, _P],
})


def net_config(in_dim, hidden, layer_N, use_relu, use_feature_norm, recurrent, head_dims, is_critic):
    hd = (C.c_int32 * 4)(*(list(head_dims) + [0] * (4 - len(head_dims))))
    # `recurrent` carries recurrent_N (0 = feed-forward; True counts as 1)
    return NetConfig(int(in_dim), int(hidden), int(layer_N), int(bool(use_relu)), int(bool(use_feature_norm)),
                     int(recurrent), len(head_dims), hd, int(bool(is_critic)))


def net_param_count(cfg):
    L = _lib.lib()
    L.mappo_net_param_count.restype = C.c_longlong
    L.mappo_net_param_count.argtypes = [_NC]
    n = L.mappo_net_param_count(C.byref(cfg))
    assert n > 0
    return int(n)


def _rows_view(x):
    """2-D float32 CUDA view [rows, width] with unit inner stride -> (tensor, row stride in floats)."""
    assert x.dtype == torch.float32 and x.is_cuda and x.dim() == 2 and x.stride(1) == 1, \
        "expected a float32 CUDA matrix with contiguous rows"
    return x, int(x.stride(0))


def _vp(t):
    """pointer of a possibly strided view (base pointer of its first element); None -> NULL"""
    return None if t is None else C.c_void_p(t.data_ptr())


def actor_forward(cfg, params, obs, actions_out, logp_out, available_actions=None, deterministic=False, seed=0,
                  step=0, row_offset=0, given_actions=None, entropy_out=None, step_offset=None):
    """step_offset: int64 device scalar added to `step` in the kernel (CUDA-graph replays), or None."""
    obs, ld = _rows_view(obs)
    ws = mlp_workspace(cfg, obs.shape[0], obs.device)
    with TIMERS.span("actor_forward"):
        check(_lib.lib().mappo_actor_forward(C.byref(cfg), ptr(params), _vp(obs), ld, obs.shape[0],
                                             ptr(available_actions), int(deterministic), int(seed) & (2 ** 64 - 1),
                                             int(step), int(row_offset), ptr(given_actions), ptr(actions_out),
                                             ptr(logp_out), ptr(entropy_out), ptr(step_offset), ptr(ws), stream_ptr()),
              "actor_forward")


def advantages(returns, value_preds, active_masks, vn_stats, out=None, stats3=None, all_reduce=None):
    """returns/value_preds: the [T, ...] slices (contiguous).  -> (advantages_raw, stats3)"""
    L = _lib.lib()
    n = returns.numel()
    if out is None:
        out = torch.empty_like(returns)
    nparts = L.mappo_moments_num_partials(n)
    partials = torch.empty(nparts, 3, dtype=torch.float64, device=returns.device)
    if stats3 is None:
        stats3 = torch.empty(3, dtype=torch.float64, device=returns.device)
    check(L.mappo_advantages(ptr(_f32c(returns, "returns")), ptr(_f32c(value_preds, "value_preds")), ptr(active_masks),
                             ptr(vn_stats), n, ptr(out), ptr(partials), stream_ptr()), "advantages")
    check(L.mappo_adv_stats_reduce(ptr(partials), nparts, ptr(stats3), stream_ptr()), "adv_stats_reduce")
    if all_reduce is not None:
        all_reduce(stats3)
    return out, stats3


_WORKSPACES = {}
# A captured CUDA graph holds the raw pointers of the scratch buffers that existed at capture time.  Once a runner has
# captured one (pin_workspaces()), a buffer that has to grow is REPLACED for eager callers but the old allocation is
# kept alive here, so later replays never touch freed memory (eval / render / a public ppo_update() may ask for more).
_RETIRED = []
GRAPH_PINNED = False


def pin_workspaces():
    global GRAPH_PINNED
    GRAPH_PINNED = True


def _store_workspace(key, new):
    old = _WORKSPACES.get(key)
    if old is not None and GRAPH_PINNED:
        _RETIRED.append(old)
    _WORKSPACES[key] = new
    return new


def mlp_workspace(cfg, rows, device):
    """Scratch for a wide (in_dim > 64) actor or critic call, grown on demand and kept per device; None when not needed."""
    L = _lib.lib()
    L.mappo_mlp_workspace_bytes.argtypes = [_NC, C.c_longlong]
    L.mappo_mlp_workspace_bytes.restype = C.c_longlong
    need = int(L.mappo_mlp_workspace_bytes(C.byref(cfg), int(rows)))
    if need <= 0:
        return None
    ws = _WORKSPACES.get(device)
    if ws is None or ws.numel() * 4 < need:
        ws = _store_workspace(device, torch.empty((need + 3) // 4, dtype=torch.float32, device=device))
    return ws


def critic_forward(cfg, params, share_obs, values_out, group=1):
    x, ld = _rows_view(share_obs)
    assert values_out.numel() == x.shape[0] * group
    ws = mlp_workspace(cfg, x.shape[0], x.device)
    with TIMERS.span("critic_forward"):
        check(_lib.lib().mappo_critic_forward(C.byref(cfg), ptr(params), _vp(x), ld, x.shape[0], int(group),
                                              ptr(values_out), ptr(ws), stream_ptr()), "critic_forward")


def rnn_workspace(cfg, positions, update, device):
    L = _lib.lib()
    L.mappo_rnn_workspace_bytes.argtypes = [_NC, C.c_longlong, C.c_int]
    L.mappo_rnn_workspace_bytes.restype = C.c_longlong
    need = int(L.mappo_rnn_workspace_bytes(C.byref(cfg), int(positions), int(update)))
    key = (device, "rnn")
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() * 4 < need:
        ws = _store_workspace(key, torch.empty((need + 3) // 4, dtype=torch.float32, device=device))
    return ws


def _h2(x, rows):
    """rnn states [rows, recurrent_N, H] (or [rows, H]) -> contiguous float32 [rows, recurrent_N * H] view"""
    x = x.reshape(rows, -1)
    assert x.dtype == torch.float32 and x.is_cuda and x.is_contiguous() and x.shape[1] % 64 == 0, \
        "rnn states must be contiguous [rows, recurrent_N * 64]"
    return x


def actor_forward_rnn(cfg, params, obs, rnn_in, masks, rnn_out, actions_out, logp_out, available_actions=None,
                      deterministic=False, seed=0, step=0, row_offset=0, given_actions=None, entropy_out=None,
                      step_offset=None):
    obs, ld = _rows_view(obs)
    rows = obs.shape[0]
    ws = rnn_workspace(cfg, rows, 0, obs.device)
    L = _lib.lib()
    L.mappo_actor_forward_rnn.argtypes = [_NC, _P, _P, C.c_int, C.c_longlong, _P, _P, _P, C.c_int, C.c_uint64, C.c_uint64,
                                          C.c_longlong, _P, _P, _P, _P, _P, _P, _P, _P]
    with TIMERS.span("actor_forward"):
        check(L.mappo_actor_forward_rnn(C.byref(cfg), ptr(params), _vp(obs), ld, rows, ptr(_h2(rnn_in, rows)),
                                        ptr(_f32c(masks.reshape(rows), "masks")), ptr(available_actions),
                                        int(deterministic), int(seed) & (2 ** 64 - 1), int(step), int(row_offset),
                                        ptr(given_actions), ptr(_h2(rnn_out, rows)), ptr(actions_out), ptr(logp_out),
                                        ptr(entropy_out), ptr(step_offset), ptr(ws), stream_ptr()), "actor_forward_rnn")


def critic_forward_rnn(cfg, params, share_obs, rnn_in, masks, rnn_out, values_out, share_rows_div=1):
    x, ld = _rows_view(share_obs)
    rows = values_out.numel()
    ws = rnn_workspace(cfg, rows, 0, x.device)
    L = _lib.lib()
    L.mappo_critic_forward_rnn.argtypes = [_NC, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, _P, _P, _P, _P, _P]
    with TIMERS.span("critic_forward"):
        check(L.mappo_critic_forward_rnn(C.byref(cfg), ptr(params), _vp(x), ld, rows, int(share_rows_div),
                                         ptr(_h2(rnn_in, rows)), ptr(_f32c(masks.reshape(rows), "masks")),
                                         ptr(_h2(rnn_out, rows)), ptr(values_out), ptr(ws), stream_ptr()),
              "critic_forward_rnn")


def actor_update_rnn(cfg, loss, params, obs, n_chunks, chunk_len, index, rnn_states, masks, actions, old_logp, adv,
                     active_masks, available_actions, grad_partials, loss_partials):
    obs, ld = _rows_view(obs)
    ws = rnn_workspace(cfg, n_chunks * chunk_len, 1, obs.device)
    L = _lib.lib()
    L.mappo_actor_update_rnn.argtypes = [_NC, _LC, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, _P, _P, _P, _P, _P, _P,
                                         _P, _P, C.c_int, _P, _P, _P]
    assert index.dtype == torch.int64 and index.numel() == n_chunks * chunk_len
    with TIMERS.span("actor_update"):
        check(L.mappo_actor_update_rnn(C.byref(cfg), C.byref(loss), ptr(params), _vp(obs), ld, int(n_chunks),
                                       int(chunk_len), ptr(index), ptr(rnn_states), ptr(masks), ptr(actions),
                                       ptr(old_logp), ptr(adv), ptr(active_masks), ptr(available_actions),
                                       ptr(grad_partials), int(grad_partials.shape[0]), ptr(loss_partials), ptr(ws),
                                       stream_ptr()), "actor_update_rnn")


def critic_update_rnn(cfg, loss, params, share_obs, n_chunks, chunk_len, index, rnn_states, masks, value_preds, returns,
                      active_masks, vn_stats, grad_partials, loss_partials, share_rows_div=1):
    x, ld = _rows_view(share_obs)
    ws = rnn_workspace(cfg, n_chunks * chunk_len, 1, x.device)
    L = _lib.lib()
    L.mappo_critic_update_rnn.argtypes = [_NC, _LC, _P, _P, C.c_int, C.c_longlong, C.c_int, _P, C.c_int, _P, _P, _P, _P,
                                          _P, _P, _P, C.c_int, _P, _P, _P]
    with TIMERS.span("critic_update"):
        check(L.mappo_critic_update_rnn(C.byref(cfg), C.byref(loss), ptr(params), _vp(x), ld, int(n_chunks),
                                        int(chunk_len), ptr(index), int(share_rows_div), ptr(rnn_states), ptr(masks),
                                        ptr(value_preds), ptr(returns), ptr(active_masks), ptr(vn_stats),
                                        ptr(grad_partials), int(grad_partials.shape[0]), ptr(loss_partials), ptr(ws),
                                        stream_ptr()), "critic_update_rnn")


def update_num_blocks(cfg, rows):
    L = _lib.lib()
    L.mappo_update_num_blocks.argtypes = [_NC, C.c_longlong]
    L.mappo_update_num_blocks.restype = C.c_int
    return int(L.mappo_update_num_blocks(C.byref(cfg), int(rows)))


def kernel_path(cfg):
    """0 fp32 tile kernels, 1 fused tcgen05 kernels, 2 the same behind a wide first layer, 3 recurrent chain (host-side
    decision of the library, include/mappo_b200.h: mappo_mlp_kernel_path)"""
    L = _lib.lib()
    L.mappo_mlp_kernel_path.argtypes = [_NC]
    L.mappo_mlp_kernel_path.restype = C.c_int
    return int(L.mappo_mlp_kernel_path(C.byref(cfg)))


def actor_update(cfg, loss, params, obs, actions, old_logp, adv, active_masks, available_actions, grad_partials,
                 loss_partials, index=None):
    obs, ld = _rows_view(obs)
    rows = obs.shape[0] if index is None else index.numel()
    ws = mlp_workspace(cfg, rows, obs.device)
    with TIMERS.span("actor_update"):
        check(_lib.lib().mappo_actor_update(C.byref(cfg), C.byref(loss), ptr(params), _vp(obs), ld, rows, ptr(index),
                                            ptr(actions), ptr(old_logp), ptr(adv), ptr(active_masks),
                                            ptr(available_actions), ptr(grad_partials), ptr(loss_partials), ptr(ws),
                                            stream_ptr()), "actor_update")


def critic_pack_inputs(cfg, share_obs):
    """Wide first layer: convert the share_obs rows once into the packed fp16-pair operand tiles every epoch's forward
    pass then streams (mappo_critic_pack_inputs).  Returns the packed buffer (kept per device), or None when this
    network does not pack.  Pass it to critic_update(packed=...) with the SAME rows."""
    x, ld = _rows_view(share_obs)
    L = _lib.lib()
    L.mappo_critic_packed_bytes.argtypes = [_NC, C.c_longlong]
    L.mappo_critic_packed_bytes.restype = C.c_longlong
    need = int(L.mappo_critic_packed_bytes(C.byref(cfg), int(x.shape[0])))
    if need <= 0:
        return None
    key = (x.device, "packed")
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < need:
        buf = _store_workspace(key, torch.empty(need, dtype=torch.uint8, device=x.device))
    ws = mlp_workspace(cfg, x.shape[0], x.device)
    check(L.mappo_critic_pack_inputs(C.byref(cfg), _vp(x), ld, x.shape[0], ptr(buf), ptr(ws), stream_ptr()),
          "critic_pack_inputs")
    return buf


def critic_update(cfg, loss, params, share_obs, value_preds, returns, active_masks, vn_stats, grad_partials,
                  loss_partials, group=1, index=None, share_rows_div=1, packed=None):
    x, ld = _rows_view(share_obs)
    rows = x.shape[0] if index is None else index.numel()
    ws = mlp_workspace(cfg, rows, x.device)
    with TIMERS.span("critic_update"):
        check(_lib.lib().mappo_critic_update(C.byref(cfg), C.byref(loss), ptr(params), _vp(x), ld, rows, int(group),
                                             ptr(index), int(share_rows_div), ptr(value_preds), ptr(returns),
                                             ptr(active_masks), ptr(vn_stats), ptr(grad_partials), ptr(loss_partials),
                                             ptr(ws), ptr(packed), stream_ptr()), "critic_update")


def grad_reduce(grad_partials, loss_partials, blocks, grad_sum, loss_sum):
    fn = _lib.lib().mappo_grad_reduce_f64 if grad_sum.dtype == torch.float64 else _lib.lib().mappo_grad_reduce
    with TIMERS.span("grad_reduce"):
        check(fn(ptr(grad_partials), ptr(loss_partials), int(blocks), grad_sum.numel(),
                                           ptr(grad_sum), ptr(loss_sum), stream_ptr()), "grad_reduce")


def adam_step(cfg, adam, params, exp_avg, exp_avg_sq, grad_sum, loss_sum, denominator_slot, rows_total, grad_out,
              info_out=None, apply_update=True, step_dev=None):
    """step_dev: int64 device scalar holding the optimiser step count (read for the bias corrections and incremented
    by the kernel), or None for the host-computed corrections in `adam`."""
    fn = _lib.lib().mappo_adam_step_f64 if grad_sum.dtype == torch.float64 else _lib.lib().mappo_adam_step
    with TIMERS.span("adam_step"):
        check(fn(C.byref(cfg), C.byref(adam), ptr(params), ptr(exp_avg), ptr(exp_avg_sq),
                                         ptr(grad_sum), ptr(loss_sum), int(denominator_slot), int(rows_total),
                                         int(apply_update), ptr(grad_out), ptr(info_out), ptr(step_dev), stream_ptr()),
              "adam_step")
