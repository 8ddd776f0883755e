#!/usr/bin/env python
"""Determinism probe: the same rollout forward launched repeatedly must give bit-identical outputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from test_gpu_policy import build, sample_of
fx = np.load(os.path.join(ROOT, "tests", "golden", "ppo_mlp_tanh.npz"))
args, pol, tr = build(fx)
s = sample_of(fx)
obs = torch.as_tensor(s[1]).cuda()
share = torch.as_tensor(s[0]).cuda()
B = obs.shape[0]
acts = torch.empty(B, 1, device="cuda"); lps = torch.empty(B, 1, device="cuda"); vals = torch.empty(B, 1, device="cuda")
ref_lp = ref_v = None
bad = 0
for it in range(400):
    pol.critic.forward_into(share, vals)
    pol.actor.forward_into(obs, acts, lps, deterministic=True)
    lp = lps.clone(); v = vals.clone()
    if ref_lp is None:
        ref_lp, ref_v = lp, v
        continue
    d = (lp != ref_lp).nonzero()[:, 0]
    dv = (v != ref_v).nonzero()[:, 0]
    if d.numel() or dv.numel():
        bad += 1
        if bad <= 6:
            print("iter", it, "lp rows", d.tolist()[:20], "max", (lp - ref_lp).abs().max().item(), "| v rows", dv.tolist()[:20],
                  "max", (v - ref_v).abs().max().item() if dv.numel() else 0)
print("nondeterministic launches:", bad, "of 399")
from oracle import policy as op
from ppo_common import sd_from
asd = sd_from(fx, "actor0.")
feats, _ = op.actor_features(asd, s[1], s[2], s[7], args.use_ReLU, args.use_feature_normalization, False)
a_ref, lp_ref, _ = op.act_forward(op.actor_logits(asd, feats, s[11]), True)
same = (acts.cpu().numpy() == a_ref.numpy())
err = np.abs(ref_lp.cpu().numpy() - lp_ref.numpy())[same]
import hashlib
print("hash gpu", hashlib.md5(ref_lp.cpu().numpy().tobytes()).hexdigest()[:10], "hash oracle", hashlib.md5(lp_ref.numpy().tobytes()).hexdigest()[:10], "threads", torch.get_num_threads())
print("vs oracle: same actions %.4f, max |dlp| %.3e, n > 6e-5: %d" % (same.mean(), err.max(), (err > 6e-5).sum()))

if (err > 6e-5).any():
    full = np.abs(ref_lp.cpu().numpy() - lp_ref.numpy())[:, 0]
    rows = np.nonzero(full > 6e-5)[0]
    print("rows", rows.tolist(), "tile", (rows // 128).tolist(), "r", (rows % 128).tolist())
    print("gpu", ref_lp.cpu().numpy()[rows, 0], "oracle", lp_ref.numpy()[rows, 0])
    # second launch through a fresh policy object with the same weights
    args2, pol2, tr2 = build(fx)
    pol2.actor.forward_into(obs, acts, lps, deterministic=True)
    print("fresh policy same as first:", bool((lps == ref_lp).all()), "max", (lps - ref_lp).abs().max().item())
    print("flat equal:", bool((pol2.actor.flat == pol.actor.flat).all()))
    # FMA path
    os.environ["MAPPO_MLP_IMPL"] = "fma"
