"""ValueNorm mirror (onpolicy/utils/valuenorm.py:8-78): same constructor, attributes and methods; the three
statistics live in one float32[3] device tensor read directly by the GAE and value-loss kernels."""
import numpy as np
import torch
import torch.nn as nn

from .. import ops


class ValueNorm(nn.Module):
    def __init__(self, input_shape=1, norm_axes=1, beta=0.99999, per_element_update=False, epsilon=1e-5,
                 device="cuda"):
        super().__init__()
        if input_shape not in (1, (1,), [1]) or per_element_update or norm_axes != 1:
            raise NotImplementedError("ValueNorm: only the scalar, per-batch update the MPE path uses")
        assert abs(epsilon - 1e-5) < 1e-12, "epsilon is fixed at 1e-5 in the kernels"
        self.input_shape, self.norm_axes, self.beta, self.epsilon = input_shape, norm_axes, beta, epsilon
        self.per_element_update = per_element_update
        # [running_mean, running_mean_sq, debiasing_term]
        self.register_buffer("stats", torch.zeros(3, dtype=torch.float32, device=device), persistent=False)
        self.all_reduce = None   # set by the trainer in multi-GPU jobs: sums a double[3] over ranks

    # the reference exposes three nn.Parameters; keep them readable/writable under the same names
    running_mean = property(lambda s: s.stats[0:1])
    running_mean_sq = property(lambda s: s.stats[1:2])
    debiasing_term = property(lambda s: s.stats[2])

    def reset_parameters(self):
        self.stats.zero_()

    def state_dict(self, *a, **k):   # same keys/shapes as the reference's vnorm.pt
        return {"running_mean": self.stats[0:1].clone(), "running_mean_sq": self.stats[1:2].clone(),
                "debiasing_term": self.stats[2].clone()}

    def load_state_dict(self, sd, strict=True):
        self.stats[0] = sd["running_mean"].reshape(()).to(self.stats)
        self.stats[1] = sd["running_mean_sq"].reshape(()).to(self.stats)
        self.stats[2] = sd["debiasing_term"].reshape(()).to(self.stats)

    def running_mean_var(self):
        db = self.stats[2].clamp(min=self.epsilon)
        mean = self.stats[0:1] / db
        var = (self.stats[1:2] / db - mean ** 2).clamp(min=1e-2)
        return mean, var

    def _dev(self, x):
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        return x.to(self.stats.device, torch.float32).contiguous()

    @torch.no_grad()
    def update(self, input_vector, moments=None):
        """`moments`: the double[3] batch moments of `input_vector` if the caller already has them (the trainer
        computes them once per iteration when every epoch sees the same return batch)."""
        s3 = moments if moments is not None else self.batch_moments(input_vector)
        ops.valuenorm_apply(self.stats, s3, self.beta)

    @torch.no_grad()
    def batch_moments(self, input_vector):
        return ops.moments(self._dev(input_vector), all_reduce=self.all_reduce)

    def normalize(self, input_vector):
        return ops.valuenorm_transform(self.stats, self._dev(input_vector), False)

    def denormalize(self, input_vector):
        """Returns a device tensor (the reference returns numpy, forcing a D2H copy: valuenorm.py:76)."""
        return ops.valuenorm_transform(self.stats, self._dev(input_vector), True)
