"""TEST INFRASTRUCTURE -- minimal ``gymnasium`` stand-in (Env, Wrapper, spaces.Box, register)."""
from . import spaces  # noqa: F401


class Env:
    _np_random = None

    @property
    def np_random(self):
        import numpy as np
        if self._np_random is None:
            self._np_random = np.random.default_rng()
        return self._np_random

    @np_random.setter
    def np_random(self, rng):
        self._np_random = rng

    def reset(self, seed=None, options=None):
        pass


class Wrapper(Env):
    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):                     # gymnasium.Wrapper forwards unknown attributes to the wrapped env
        if name.startswith("_") or name == "env":
            raise AttributeError(name)
        return getattr(self.env, name)
