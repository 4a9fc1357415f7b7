"""Throw-away stand-in for the `gym` package (absent in this image, no network).

TEST INFRASTRUCTURE ONLY. It lets the *unmodified* reference package under
/root/reference import and run so that oracle/refharness can export golden
vectors from it. It implements only the API surface
the reference touches: Env, the wrapper base classes, spaces with seed/sample,
and the registration registry used by gym.make.
"""
import importlib

from . import spaces  # noqa: F401
from .envs import registration  # noqa: F401
from .envs.registration import make, register  # noqa: F401

__version__ = "0.0-stub"


class Env:
    metadata = {}
    action_space = None
    observation_space = None

    def reset(self, *a, **k):
        raise NotImplementedError

    def step(self, action):
        raise NotImplementedError

    def render(self, mode="human"):
        return None

    def close(self):
        return None

    def seed(self, seed=None):
        return None

    @property
    def unwrapped(self):
        return self


class Wrapper(Env):
    def __init__(self, env):
        self.env = env
        self.action_space = env.action_space
        self.observation_space = env.observation_space

    def __getattr__(self, name):
        if name == "env":
            raise AttributeError(name)
        return getattr(self.env, name)

    @property
    def unwrapped(self):
        return self.env.unwrapped

    def reset(self, **kwargs):
        return self.env.reset(**kwargs)

    def step(self, action):
        return self.env.step(action)


class ObservationWrapper(Wrapper):
    def reset(self, **kwargs):
        return self.observation(self.env.reset(**kwargs))

    def step(self, action):
        out = self.env.step(action)
        return (self.observation(out[0]),) + tuple(out[1:])


class ActionWrapper(Wrapper):
    def step(self, action):
        return self.env.step(self.action(action))


class RewardWrapper(Wrapper):
    def step(self, action):
        out = self.env.step(action)
        return (out[0], self.reward(out[1])) + tuple(out[2:])
