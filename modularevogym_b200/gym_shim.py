"""Import-time shims that let the reference's Python layer (``evogym``, ``evogym.envs``) run on this engine unchanged
(SURVEY 8f-3).  Nothing here computes anything.

``install()`` does up to three things, each only if needed:

* ``evogym.simulator_cpp`` -> this package's pybind11 module (``modularevogym_b200/simulator_cpp*.so``), registered in
  ``sys.modules`` so that ``from evogym.simulator_cpp import Sim`` (``evogym/sim.py:14``) and
  ``from evogym.simulator_cpp import Viewer, Camera`` (``evogym/viewer.py:6``) resolve to the CUDA engine without copying
  files into the reference's package (the alternative in INTEGRATION.md section 2);
* ``numpy.float`` -> ``float``: the alias numpy 1.24 removed, used by every environment's space declaration
  (``evogym/envs/walk.py:29-30``);
* a minimal ``gym`` (``Env``, ``spaces.Box`` / ``Dict``, ``envs.registration.register`` + ``make``, ``Wrapper``,
  ``utils.seeding``) when neither ``gym`` is installed: the reference's environments use gym for declarations, registration
  and the step-limit wrapper only.
"""
from __future__ import annotations

import importlib
import sys
import types
from typing import Optional

import numpy as np


def _minimal_gym() -> types.ModuleType:
    gym = types.ModuleType("gym")
    gym.__doc__ = "minimal stand-in installed by modularevogym_b200.gym_shim (gym is not installed)"

    class Env:
        metadata: dict = {}
        reward_range = (-float("inf"), float("inf"))
        spec = None
        action_space = None
        observation_space = None

        def seed(self, seed=None):
            return [seed]

        def close(self):
            pass

        @property
        def unwrapped(self):
            return self

    class Space:
        def __init__(self, shape=None, dtype=None):
            self.shape, self.dtype = (None if shape is None else tuple(shape)), dtype

    class Box(Space):
        def __init__(self, low, high, shape=None, dtype=np.float32):
            shape = tuple(shape) if shape is not None else np.asarray(low).shape
            super().__init__(shape, dtype)
            self.low = np.broadcast_to(np.asarray(low, dtype=np.float64), shape)
            self.high = np.broadcast_to(np.asarray(high, dtype=np.float64), shape)

        def sample(self):
            return np.random.uniform(self.low, self.high).astype(self.dtype if self.dtype is not float else np.float64)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    class Dict_(Space):
        def __init__(self, spaces=None, **kw):
            super().__init__(None, None)
            self.spaces = dict(spaces or {}, **kw)

        def __getitem__(self, k):
            return self.spaces[k]

        def keys(self):
            return self.spaces.keys()

        def items(self):
            return self.spaces.items()

    class Wrapper(Env):
        def __init__(self, env):
            self.env = env
            self.action_space, self.observation_space = env.action_space, env.observation_space

        def __getattr__(self, name):
            if name.startswith("_"):
                raise AttributeError(name)
            return getattr(self.env, name)

        def step(self, action):
            return self.env.step(action)

        def reset(self, **kw):
            return self.env.reset(**kw)

        @property
        def unwrapped(self):
            return self.env.unwrapped

    class TimeLimit(Wrapper):
        """gym.wrappers.TimeLimit: done after max_episode_steps, info['TimeLimit.truncated'] = not done before"""

        def __init__(self, env, max_episode_steps=None):
            super().__init__(env)
            self._max_episode_steps, self._elapsed_steps = max_episode_steps, None

        def step(self, action):
            obs, rew, done, info = self.env.step(action)
            self._elapsed_steps += 1
            if self._max_episode_steps is not None and self._elapsed_steps >= self._max_episode_steps:
                info["TimeLimit.truncated"] = not done
                done = True
            return obs, rew, done, info

        def reset(self, **kw):
            self._elapsed_steps = 0
            return self.env.reset(**kw)

    registry = {}

    def register(id, entry_point=None, max_episode_steps=None, **kw):
        registry[id] = (entry_point, max_episode_steps, kw)

    def make(id, **kwargs):
        if id not in registry:
            raise KeyError(f"No registered env with id: {id}")
        entry_point, limit, kw = registry[id]
        mod_name, cls_name = entry_point.split(":")
        cls = getattr(importlib.import_module(mod_name), cls_name)
        env = cls(**dict(kw.get("kwargs", {}), **kwargs))
        return TimeLimit(env, limit) if limit else env

    def np_random(seed=None):
        return np.random.RandomState(seed), seed

    def mk(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    spaces = mk("gym.spaces", Box=Box, Dict=Dict_, Space=Space)
    mk("gym.spaces.box", Box=Box)
    error = mk("gym.error", Error=type("Error", (Exception,), {}))
    seeding = mk("gym.utils.seeding", np_random=np_random)
    utils = mk("gym.utils", seeding=seeding)
    registration = mk("gym.envs.registration", register=register, make=make, registry=registry)
    envs = mk("gym.envs", registration=registration)
    wrappers = mk("gym.wrappers", TimeLimit=TimeLimit)
    gym.Env, gym.Wrapper, gym.ObservationWrapper = Env, Wrapper, Wrapper
    gym.spaces, gym.error, gym.utils, gym.envs, gym.wrappers = spaces, error, utils, envs, wrappers
    gym.make, gym.register = make, register
    sys.modules["gym"] = gym
    return gym


def install(simulator: bool = True, simulator_module: Optional[types.ModuleType] = None, force_gym: bool = False) -> dict:
    """Returns what was installed: ``{'simulator': bool, 'np_float': bool, 'gym': bool}``.  ``simulator_module`` overrides the
    module registered as ``evogym.simulator_cpp`` (the fixture generator registers the compiled reference engine)."""
    done = {"simulator": False, "np_float": False, "gym": False}
    if simulator or simulator_module is not None:
        if simulator_module is None:
            from . import simulator_cpp as simulator_module      # the pybind11 module over libmevo_b200.so
        sys.modules["evogym.simulator_cpp"] = simulator_module
        done["simulator"] = True
    if not hasattr(np, "float"):
        np.float = float                                          # noqa: NPY001
        done["np_float"] = True
    have_gym = False
    if not force_gym:
        try:
            import gym  # noqa: F401
            have_gym = True
        except ImportError:
            pass
    if not have_gym:
        _minimal_gym()
        done["gym"] = True
    return done
