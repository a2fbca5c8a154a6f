"""`gymnasium` when it is installed, otherwise a minimal stand-in with the pieces Pulselib uses.

Pulselib's envs subclass `gym.Env`, declare `spaces.Discrete/Box`, are registered with
`gymnasium.envs.registration.register` and created with `gym.make(id, **kwargs)`; callers then use
`env.unwrapped`, `env.close()` and read attributes through the wrapper
(/root/reference/environments/__init__.py:1-31, scripts/Poker/trainGPU.py:177-188).  The image this
engine is developed on has no gymnasium, so the same surface is provided here.
"""
from __future__ import annotations

import importlib
import sys
import types

import numpy as np

try:  # pragma: no cover - depends on the image
    import gymnasium as gym  # type: ignore
    from gymnasium import spaces  # type: ignore
    from gymnasium.envs.registration import register  # type: ignore

    HAVE_GYMNASIUM = True
except Exception:  # ModuleNotFoundError in this image
    HAVE_GYMNASIUM = False

    class Space:
        def __init__(self, shape=None, dtype=None):
            self.shape = shape
            self.dtype = dtype
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

    class Discrete(Space):
        def __init__(self, n, start=0):
            super().__init__((), np.int64)
            self.n = int(n)
            self.start = int(start)

        def sample(self):
            return int(self._rng.integers(self.start, self.start + self.n))

        def contains(self, x):
            return self.start <= int(x) < self.start + self.n

    class Box(Space):
        def __init__(self, low, high, shape=None, dtype=np.float32):
            if shape is None:
                shape = np.shape(low)
            super().__init__(tuple(shape), dtype)
            self.low = np.broadcast_to(self._bound(low, dtype), self.shape)
            self.high = np.broadcast_to(self._bound(high, dtype), self.shape)

        @staticmethod
        def _bound(v, dtype):
            v = np.asarray(v, dtype=np.float64)
            if np.issubdtype(np.dtype(dtype), np.integer):  # gymnasium clips infinite bounds of integer boxes
                info = np.iinfo(dtype)
                v = np.clip(v, info.min, info.max)
            return v.astype(dtype)

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1e6)
            hi = np.where(np.isfinite(self.high), self.high, 1e6)
            return self._rng.uniform(lo, hi).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    class Env:
        metadata: dict = {}
        action_space = None
        observation_space = None
        spec = None
        render_mode = None

        def reset(self, *, seed=None, options=None):
            if seed is not None:
                self.np_random = np.random.default_rng(seed)
            return None

        def step(self, action):
            raise NotImplementedError

        def render(self):
            return None

        def close(self):
            return None

        @property
        def unwrapped(self):
            return self

    class _Spec:
        def __init__(self, id, entry_point, max_episode_steps=None, kwargs=None):
            self.id = id
            self.entry_point = entry_point
            self.max_episode_steps = max_episode_steps
            self.kwargs = dict(kwargs or {})

    registry: dict = {}

    def register(id, entry_point=None, max_episode_steps=None, kwargs=None, **_):
        registry[id] = _Spec(id, entry_point, max_episode_steps, kwargs)

    class _Wrapper:
        """Attribute-transparent wrapper, like gymnasium's wrapper chain (reads fall through)."""

        def __init__(self, env, spec):
            self.__dict__["env"] = env
            self.__dict__["spec"] = spec

        @property
        def unwrapped(self):
            return self.__dict__["env"]

        def __getattr__(self, name):
            return getattr(self.__dict__["env"], name)

        def reset(self, **kw):
            return self.__dict__["env"].reset(**kw)

        def step(self, action):
            return self.__dict__["env"].step(action)

        def close(self):
            return self.__dict__["env"].close()

    def make(id, **kwargs):
        if id not in registry:
            raise KeyError(f"No registered env with id: {id}")
        spec = registry[id]
        ep = spec.entry_point
        if isinstance(ep, str):
            mod, _, attr = ep.partition(":")
            ep = getattr(importlib.import_module(mod), attr)
        kw = dict(spec.kwargs)
        kw.update(kwargs)
        return _Wrapper(ep(**kw), spec)

    spaces = types.ModuleType("gymnasium.spaces")
    spaces.Space, spaces.Discrete, spaces.Box = Space, Discrete, Box
    gym = types.ModuleType("gymnasium")
    gym.Env, gym.spaces, gym.make, gym.register = Env, spaces, make, register
    gym.registry = registry
    _envs = types.ModuleType("gymnasium.envs")
    _reg = types.ModuleType("gymnasium.envs.registration")
    _reg.register, _reg.registry = register, registry
    _envs.registration = _reg
    gym.envs = _envs


def install_shim() -> None:
    """Makes `import gymnasium` resolve to the stand-in (no-op when the real package exists)."""
    if HAVE_GYMNASIUM:
        return
    sys.modules.setdefault("gymnasium", gym)
    sys.modules.setdefault("gymnasium.spaces", gym.spaces)
    sys.modules.setdefault("gymnasium.envs", gym.envs)
    sys.modules.setdefault("gymnasium.envs.registration", gym.envs.registration)
