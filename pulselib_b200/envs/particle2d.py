"""Particle2D behind the reference's gym.Env surface
(/root/reference/environments/Particle2D/Particle2D.py:6-30), one fused CUDA launch per step."""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .._lib import c_f, c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces


class Particle2D(gym.Env):
    metadata = {"render.modes": ["human"]}

    def __init__(self, device, batch_size, dt=0.1, max_steps=200, seed=0, env_base=0):
        super().__init__()
        self.device = _lib.require_cuda(device)
        self._lib = _lib.load()
        self.batch_size, self.dt, self.max_steps = int(batch_size), dt, max_steps
        self.action_space = spaces.Box(-1, 1, (2,), np.float32)
        self.observation_space = spaces.Box(-np.inf, np.inf, (4,), np.float32)
        self.seed_value, self.env_base, self._episode = int(seed), int(env_base), 0
        n, dev = self.batch_size, self.device
        self.state = torch.zeros((n, 4), dtype=torch.float32, device=dev)
        self.steps = torch.zeros(n, dtype=torch.int32, device=dev)
        self.terminated = torch.zeros(n, dtype=torch.bool, device=dev)
        self._truncated = torch.zeros(n, dtype=torch.bool, device=dev)

    def reset(self, seed=None, options=None):
        if seed is not None:
            self.seed_value, self._episode = int(seed), 0
        n, dev = self.batch_size, self.device
        # fresh tensors each reset, like the reference (Particle2D.py:16-19)
        self.state = torch.empty((n, 4), dtype=torch.float32, device=dev)
        self.steps = torch.empty(n, dtype=torch.int32, device=dev)
        self.terminated = torch.empty(n, dtype=torch.bool, device=dev)
        obs = torch.empty((n, 4), dtype=torch.float32, device=dev)
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_particle2d_reset(
                ptr(self.state), ptr(self.steps), ptr(self.terminated), ptr(obs),
                c_u64(self.seed_value), c_u64(self.env_base), c_u32(self._episode), c_i(n),
                _lib.cuda_stream(dev)), "pulse_particle2d_reset")
        self._episode += 1
        return obs, {}

    def step(self, action):
        n, dev = self.batch_size, self.device
        action = torch.as_tensor(action, device=dev)
        if action.dtype != torch.float32 or not action.is_contiguous():
            action = action.to(torch.float32).contiguous()
        if tuple(action.shape) != (n, 2):
            raise ValueError(f"action must have shape {(n, 2)}, got {tuple(action.shape)}")
        obs = torch.empty((n, 4), dtype=torch.float32, device=dev)
        rewards = torch.empty(n, dtype=torch.float32, device=dev)
        # the reference rebinds self.terminated to a new tensor every step (Particle2D.py:29)
        self.terminated = torch.empty(n, dtype=torch.bool, device=dev)
        truncated = torch.empty(n, dtype=torch.bool, device=dev)
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_particle2d_step(
                ptr(self.state), ptr(action), ptr(self.steps), ptr(obs), ptr(rewards),
                ptr(self.terminated), ptr(truncated), c_f(self.dt), c_i(self.max_steps), c_i(n),
                _lib.cuda_stream(dev)), "pulse_particle2d_step")
        return obs, rewards, self.terminated, truncated, {}
