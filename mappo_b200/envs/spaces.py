"""Minimal gym-style spaces (gym is not a dependency).  The CLASS NAMES matter: the reference
dispatches on `space.__class__.__name__` (onpolicy/utils/shared_buffer.py:58, onpolicy/utils/util.py:31-51,
onpolicy/algorithms/utils/act.py:18-27), and so does the mirror."""
import numpy as np


class Box:
    def __init__(self, low, high, shape, dtype=np.float32):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype

    def __repr__(self):
        return "Box%s" % (self.shape,)


class Discrete:
    def __init__(self, n):
        self.n = int(n)
        self.shape = ()

    def __repr__(self):
        return "Discrete(%d)" % self.n


class MultiDiscrete:
    """onpolicy/envs/mpe/multi_discrete.py: array of [min, max] per head; `shape` is the number of heads."""

    def __init__(self, array_of_param_array):
        a = np.array(array_of_param_array)
        self.low = a[:, 0]
        self.high = a[:, 1]
        self.num_discrete_space = self.low.shape[0]
        self.n = int(np.sum(self.high) + 2)

    @property
    def shape(self):
        return self.num_discrete_space

    def __repr__(self):
        return "MultiDiscrete" + str(self.num_discrete_space)


def head_sizes(space):
    """Sizes of the categorical heads of an MPE action space."""
    name = space.__class__.__name__
    if name == "Discrete":
        return [space.n]
    if name == "MultiDiscrete":
        return [int(h - l + 1) for l, h in zip(space.low, space.high)]
    raise NotImplementedError(name)
