"""PopArt mirror (onpolicy/algorithms/utils/popart.py:7-99) -- the behaviour the reference has ON CUDA, its default
device (scripts/train/train_mpe.py:86-96).

`PopArt.__init__` builds every tensor as `nn.Parameter(...).to(dtype, device)` (popart.py:21-27).  When that `.to()`
changes the device it returns a PLAIN tensor, so on cuda `weight`, `bias`, `stddev`, `mean`, `mean_sq` and
`debiasing_term` are ordinary attributes: absent from `critic.parameters()` and `critic.state_dict()`, never seen by
`clip_grad_norm_` or Adam (rMAPPOPolicy.py:35-37); the value head only ever moves through the rescale at the end of
`update` (popart.py:69-70).  (On the CPU `.to()` returns the registered Parameter itself and `update` raises TypeError
at `self.stddev = ...`: the flag cannot run there at all.)  Pinned by tests/golden/ppo_popart*_cuda.npz, recorded from
the unmodified reference on the B200 box (tests/golden/make_golden.py::gen_popart).

The running statistics are the same three numbers, updated and debiased the same way, as ValueNorm's
(popart.py:52-66,72-76 vs valuenorm.py:31-54), so this class IS a ValueNorm (`stats` float32[3] read by the GAE and
value-loss kernels) plus the head rescale (csrc/gae.cu: popart_rescale_kernel)."""
import numpy as np
import torch

from ... import ops
from ...utils.valuenorm import ValueNorm


class PopArt(ValueNorm):
    def __init__(self, input_shape, output_shape, norm_axes=1, beta=0.99999, epsilon=1e-5, device="cuda",
                 weight=None, bias=None):
        super().__init__(1, norm_axes=norm_axes, beta=beta, epsilon=epsilon, device=device)
        assert output_shape == 1, "the value head is scalar"
        self.input_shape, self.output_shape = input_shape, output_shape
        dev = self.stats.device
        # views into the critic's flat parameter buffer when the critic owns the head (R_Critic), else own storage
        self.weight = weight if weight is not None else torch.zeros(output_shape, input_shape, device=dev)
        self.bias = bias if bias is not None else torch.zeros(output_shape, device=dev)
        self._old = torch.zeros(3, dtype=torch.float32, device=dev)

    # popart.py:24-27 under the reference's names
    mean = property(lambda s: s.stats[0:1])
    mean_sq = property(lambda s: s.stats[1:2])

    @property
    def stddev(self):
        """popart.py:64 (not debiased, clamp 1e-4; 1 before the first update)"""
        if float(self.stats[2]) == 0.0:
            return torch.ones(1, device=self.stats.device)
        return (self.stats[1:2] - self.stats[0:1] ** 2).sqrt().clamp(min=1e-4)

    def debiased_mean_var(self):
        return self.running_mean_var()

    def forward(self, input_vector):
        """popart.py:41-47.  API helper only: the training and rollout paths evaluate the head inside the fused
        critic kernels."""
        return torch.nn.functional.linear(self._dev(input_vector), self.weight, self.bias)

    @torch.no_grad()
    def update(self, input_vector, moments=None, defer_rescale=False):
        """popart.py:49-70.  defer_rescale: only move the statistics and return the statistics before the step; the
        caller rescales the head (`rescale`) after the critic kernel has run -- the reference's value loss uses
        `values` computed BEFORE update() with the old head and targets normalised with the NEW statistics
        (r_mappo.py:114-122 then :64-67)."""
        self._old.copy_(self.stats)
        ValueNorm.update(self, input_vector, moments=moments)
        if defer_rescale:
            return self._old
        self.rescale(self._old)
        return None

    @torch.no_grad()
    def rescale(self, old_stats):
        ops.popart_rescale(old_stats, self.stats, self.weight, self.bias)

    def state_dict(self, *a, **k):
        return {}                      # on cuda nothing of PopArt is registered: critic.pt / vnorm.pt do not hold it

    def load_state_dict(self, sd, strict=True):
        pass
