"""Batched 10x20 Tetris (placement + line clear), one fused CUDA launch per step.

The reference ships only empty shells for Tetris (/root/reference/environments/Tetris/Tetris.py:1-13,
TetrisNumba.py:1-16: `__init__`, `reset`, `step`, `render` are `pass`), so there is no reference
behaviour to keep.  The rules are this repo's spec (oracle/oracle_tetris.c):
action = rotation * 10 + column, hard drop, lock, clear full rows, reward = rows cleared, game over
when a piece locks above row 19.  The class keeps the gym.Env shape of the other Pulselib envs.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .._lib import c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces


class TetrisBatched(gym.Env):
    metadata = {"render_modes": ["human"]}
    WIDTH, HEIGHT, NUM_ACTIONS, NUM_PIECES = 10, 20, 40, 7

    def __init__(self, device, n_envs, seed=0, env_base=0, auto_reset=True, device_counter=False):
        super().__init__()
        self.device = _lib.require_cuda(device)
        self._lib = _lib.load()
        self.n_envs = int(n_envs)
        self.seed_value, self.env_base, self.auto_reset = int(seed), int(env_base), bool(auto_reset)
        self.action_space = spaces.Discrete(self.NUM_ACTIONS)
        self.observation_space = spaces.Box(low=0, high=1023, shape=(self.HEIGHT,), dtype=np.int16)
        n, dev = self.n_envs, self.device
        # rows[i, y] bit x = cell (x, y) of env i, row 0 at the bottom; this tensor is the observation
        self.rows = torch.zeros((n, self.HEIGHT), dtype=torch.int16, device=dev)
        self.piece = torch.zeros(n, dtype=torch.uint8, device=dev)
        self.lines = torch.zeros(n, dtype=torch.int32, device=dev)
        self.done = torch.zeros(n, dtype=torch.bool, device=dev)
        self.rewards = torch.zeros(n, dtype=torch.int32, device=dev)
        self._t = 0
        self._started = False
        # device-resident tape counter, advanced by each step launch (CUDA-graph replay; see envs/tfe.py)
        self._ctr = torch.zeros(2, dtype=torch.int32, device=dev) if device_counter else None

    def _pieces(self, p):
        if p is None:
            return None
        p = torch.as_tensor(p, device=self.device)
        if tuple(p.shape) != (self.n_envs,):
            raise ValueError(f"pieces must have shape {(self.n_envs,)}")
        return p.to(torch.uint8).contiguous()

    def reset(self, seed=None, options=None):
        """options["first_piece"]: optional uint8 [N] piece ids 0..6 (I O T S Z J L) instead of the tape."""
        if seed is not None:
            self.seed_value, self._t = int(seed), 0
        if self._ctr is not None and self._started:     # fold the device counter back into the host one
            self._t += int(self._ctr[0].item())
            self._ctr.zero_()
        first = self._pieces((options or {}).get("first_piece"))
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tetris_reset(
                ptr(self.rows), ptr(self.piece), ptr(self.lines), ptr(self.done), ptr(first), c_u64(self.seed_value),
                c_u64(self.env_base), c_u32(self._t), c_i(self.n_envs), _lib.cuda_stream(self.device)),
                "pulse_tetris_reset")
        self._t += 1
        self._started = True
        return self.rows, self.get_info()

    def step(self, actions, next_piece=None, reset_piece=None):
        """actions uint8 or int64 [N] in 0..39.  Returns (rows int16 [N,20] (the state tensor itself),
        reward int32 [N] = rows cleared, done bool [N], False, {'piece', 'lines'})."""
        if not self._started:
            raise _lib.PulseError("TetrisBatched.step called before reset")
        actions = torch.as_tensor(actions, device=self.device)
        if tuple(actions.shape) != (self.n_envs,):
            raise ValueError(f"actions must have shape {(self.n_envs,)}, got {tuple(actions.shape)}")
        if actions.dtype not in (torch.uint8, torch.int64) or not actions.is_contiguous():
            actions = actions.to(torch.int64).contiguous()
        nx, rp = self._pieces(next_piece), self._pieces(reset_piece)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tetris_step(
                ptr(self.rows), ptr(self.piece), ptr(self.lines), ptr(self.done), ptr(actions),
                c_i(actions.element_size()), ptr(nx), ptr(rp), c_u64(self.seed_value), c_u64(self.env_base),
                c_u32(self._t), ptr(self._ctr), c_i(self.auto_reset), ptr(self.rewards), c_i(self.n_envs),
                _lib.cuda_stream(self.device)), "pulse_tetris_step")
        if self._ctr is None:
            self._t += 1
        return self.rows, self.rewards, self.done, False, self.get_info()

    def get_obs(self):
        return self.rows

    def get_info(self):
        return {"piece": self.piece, "lines": self.lines}

    def cells(self) -> torch.Tensor:
        """bool [N, 20, 10] occupancy grid (row 0 at the bottom), for agents that want a dense board."""
        bits = torch.arange(self.WIDTH, device=self.device, dtype=torch.int32)
        return ((self.rows.to(torch.int32).unsqueeze(-1) >> bits) & 1).bool()

    def close(self):
        pass
