"""CPU: the oracle's buffer maths (oracle/buffer.py) against fixtures recorded from the unmodified reference."""
import os

import numpy as np
import pytest
import torch

from oracle import buffer as ob

FIELDS = ["share_obs", "obs", "rnn_states", "rnn_states_critic", "actions", "value_preds", "returns", "masks",
          "active_masks", "old_action_log_probs", "adv_targ", "available_actions"]
SRC = {"old_action_log_probs": "action_log_probs", "adv_targ": "advantages"}


def test_valuenorm_matches_reference(golden_dir):
    fx = np.load(os.path.join(golden_dir, "vnorm.npz"))
    vn = ob.OracleValueNorm()
    np.testing.assert_array_equal(vn.denormalize(fx["x"]), fx["fresh_denormalized"])
    for x, want in zip(fx["batches"], fx["stats_seq"]):
        vn.update(x)
        np.testing.assert_allclose(vn.stats(), want, rtol=2e-7, atol=0)
    vn.load(fx["stats_seq"][-1])
    np.testing.assert_array_equal(vn.normalize(fx["x"]), fx["normalized"])
    np.testing.assert_array_equal(vn.denormalize(fx["x"]), fx["denormalized"])


@pytest.mark.parametrize("tag", ["g%dp%dn%d" % (g, p, n) for g in (1, 0) for p in (0, 1) for n in (1, 0)])
def test_compute_returns_bit_exact(tag, golden_dir):
    fx = np.load(os.path.join(golden_dir, "gae.npz"))
    use_gae, proper, norm = int(tag[1]), int(tag[3]), int(tag[5])
    vn = None
    if norm:
        vn = ob.OracleValueNorm(); vn.load(fx[tag + "_vn"])
    vp, ret = ob.compute_returns(fx[tag + "_rewards"], fx[tag + "_value_preds_in"], fx[tag + "_masks"],
                                 fx[tag + "_next_value"], float(fx["gamma"]), float(fx["gae_lambda"]),
                                 use_gae=bool(use_gae), use_proper_time_limits=bool(proper),
                                 bad_masks=fx[tag + "_bad_masks"], normalizer=vn)
    np.testing.assert_array_equal(ret[:-1], fx[tag + "_returns"][:-1])
    if use_gae:
        np.testing.assert_array_equal(vp, fx[tag + "_value_preds_out"])
    raw, adv = ob.advantages(ret, vp, fx[tag + "_active_masks"], vn)
    np.testing.assert_array_equal(raw, fx[tag + "_adv_raw"])
    np.testing.assert_allclose(adv, fx[tag + "_adv"], rtol=1e-6, atol=1e-6)


def _check(fx, prefix, bufprefix, rows, rnn_rows, k, T_rows_of):
    for f in FIELDS:
        key = "%s_mb%d_%s" % (prefix, k, f)
        if key not in fx.files:
            continue
        src = fx["%s_%s" % (bufprefix, SRC.get(f, f))]
        if f in ("rnn_states", "rnn_states_critic"):
            flat = src[:-1].reshape(-1, *src.shape[-2:])
            want = flat[rnn_rows] if rnn_rows is not None else flat[rows]
        else:
            has_tp1 = f not in ("actions", "old_action_log_probs", "adv_targ")
            flat = (src[:-1] if has_tp1 else src).reshape(-1, src.shape[-1])
            want = flat[rows]
        np.testing.assert_array_equal(fx[key], want, err_msg=key)


@pytest.mark.parametrize("space", ["disc", "multi"])
def test_shared_generators_bit_exact(space, golden_dir):
    fx = np.load(os.path.join(golden_dir, "generators.npz"))
    T, N, M = int(fx["T"]), int(fx["N"]), int(fx["M"])
    bp = "shared_%s_buf" % space
    torch.manual_seed(123); perm = torch.randperm(T * N * M).numpy()
    for k, rows in enumerate(ob.feed_forward_rows(T, N, M, perm, 2)):
        _check(fx, "shared_%s_ff" % space, bp, rows, None, k, None)
    torch.manual_seed(123); perm = torch.randperm(T * N * M // 10).numpy()
    for k, (rows, rnn) in enumerate(ob.recurrent_rows(T, N, M, perm, 2, 10)):
        _check(fx, "shared_%s_rec" % space, bp, rows, rnn, k, None)
    torch.manual_seed(123); perm = torch.randperm(N * M).numpy()
    for k, (rows, rnn) in enumerate(ob.naive_recurrent_rows(T, N * M, perm, 2)):
        _check(fx, "shared_%s_naive" % space, bp, rows, rnn, k, None)


def test_separated_generators_bit_exact(golden_dir):
    fx = np.load(os.path.join(golden_dir, "generators.npz"))
    T, N = int(fx["T"]), int(fx["N"])
    torch.manual_seed(321); perm = torch.randperm(T * N).numpy()
    for k, rows in enumerate(ob.feed_forward_rows(T, N, 1, perm, 2)):
        _check(fx, "sep_ff", "sep_buf", rows, None, k, None)
    torch.manual_seed(321); perm = torch.randperm(T * N // 10).numpy()
    for k, (rows, rnn) in enumerate(ob.separated_recurrent_rows(T, N, perm, 2, 10)):
        _check(fx, "sep_rec", "sep_buf", rows, rnn, k, None)
    torch.manual_seed(321); perm = torch.randperm(N).numpy()
    for k, (rows, rnn) in enumerate(ob.naive_recurrent_rows(T, N, perm, 2)):
        _check(fx, "sep_naive", "sep_buf", rows, rnn, k, None)
