"""Import shims that let the UNMODIFIED reference (`/root/reference/onpolicy`) import in this
container (Python 3.12, numpy 2.3, torch 2.11), as listed in SURVEY.md §8(c).

Used only by `tests/golden/make_golden.py` (fixture generation, run in the build container where
`/root/reference` exists).  Nothing in the product, the `-m gpu` tests, `smoke()` or `bench.py`
imports this module: the GPU box has no `/root/reference`.
"""
import importlib.util
import sys
import types

import numpy as np

REFERENCE_ROOT = "/root/reference"


# The shim classes live at MODULE level and carry `__module__ = "gym..."`: SubprocVecEnv's workers send the spaces back
# through a Pipe (env_wrappers.py:171-172, 252-253), i.e. they are pickled by reference as `gym.spaces.Box` and must be
# importable under that name in both processes (function-local classes cannot be pickled: the worker dies and the
# parent blocks in recv() forever).
class Env:
    metadata = {}

    def close(self):
        pass


class Space:
    def __init__(self, shape=None, dtype=None):
        self.shape = None if shape is None else tuple(shape)
        self.dtype = None if dtype is None else np.dtype(dtype)


class Discrete(Space):
    def __init__(self, n):
        self.n = int(n)
        super().__init__((), np.int64)

    def __repr__(self):
        return "Discrete(%d)" % self.n


class Box(Space):
    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.shape(low)
        super().__init__(shape, dtype)
        self.low = np.full(self.shape, low, dtype=dtype)
        self.high = np.full(self.shape, high, dtype=dtype)

    def __repr__(self):
        return "Box%s" % (self.shape,)


class Tuple(Space):
    def __init__(self, spaces):
        self.spaces = tuple(spaces)
        super().__init__(None, None)

    def __getitem__(self, i):
        return self.spaces[i]


class EnvSpec:
    def __init__(self, *a, **k):
        pass


Env.__module__ = "gym"
for _c in (Space, Discrete, Box, Tuple):
    _c.__module__ = "gym.spaces"
EnvSpec.__module__ = "gym.envs.registration"


def _gym_shim():
    gym = types.ModuleType("gym")
    spaces = types.ModuleType("gym.spaces")
    spaces.Space, spaces.Discrete, spaces.Box, spaces.Tuple = Space, Discrete, Box, Tuple
    gym.Env, gym.Space, gym.spaces = Env, Space, spaces
    envs = types.ModuleType("gym.envs")
    registration = types.ModuleType("gym.envs.registration")
    registration.EnvSpec = EnvSpec
    envs.registration = registration
    gym.envs = envs
    return {"gym": gym, "gym.spaces": spaces, "gym.envs": envs, "gym.envs.registration": registration}


def _imp_shim():
    imp = types.ModuleType("imp")

    def load_source(name, pathname):
        spec = importlib.util.spec_from_file_location(name or "_mpe_scenario", pathname)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod

    imp.load_source = load_source
    return imp


def install(reference_root=REFERENCE_ROOT):
    """Install the shims and put the reference on sys.path. Idempotent."""
    for name, mod in _gym_shim().items():
        sys.modules.setdefault(name, mod)
    sys.modules.setdefault("seaborn", types.ModuleType("seaborn"))
    sys.modules.setdefault("imageio", types.ModuleType("imageio"))
    if "imp" not in sys.modules:
        sys.modules["imp"] = _imp_shim()
    if "tensorboardX" not in sys.modules:
        tbx = types.ModuleType("tensorboardX")

        class SummaryWriter:
            def __init__(self, *a, **k):
                pass

            def add_scalars(self, *a, **k):
                pass

        tbx.SummaryWriter = SummaryWriter
        sys.modules["tensorboardX"] = tbx
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    global REFERENCE_ROOT
    REFERENCE_ROOT = reference_root


def reference_args(argv):
    """Build the reference's `all_args` namespace exactly as train_mpe.py does
    (onpolicy/config.py:4, scripts/train/train_mpe.py:52-80)."""
    install(REFERENCE_ROOT)
    from onpolicy.config import get_config

    parser = get_config()
    parser.add_argument("--scenario_name", type=str, default="simple_spread")
    parser.add_argument("--num_landmarks", type=int, default=3)
    parser.add_argument("--num_agents", type=int, default=2)
    a = parser.parse_known_args(argv)[0]
    if a.algorithm_name == "rmappo":
        a.use_recurrent_policy, a.use_naive_recurrent_policy = True, False
    elif a.algorithm_name == "mappo":
        a.use_recurrent_policy, a.use_naive_recurrent_policy = False, False
    elif a.algorithm_name == "ippo":
        a.use_centralized_V = False
    return a
