"""Device ring buffer fed straight from the env step.

Replaces the hand-off in /root/reference/scripts/Poker/trainGPU.py:79-94 (five boolean-mask gathers and a clone per
step) and /root/reference/utils/ReplayBuffer.py::ReplayBufferTorch (`add` at arange(position, position + batch) %
capacity, `sample` = randint gathers): `append` is three small launches that select the learning seat's rows of the
step and write them in table order into the ring; the running total stays on the device, so the step loop never
waits for the host.

    ring = TransitionRing(device, capacity=1 << 22, obs_width=env.obs_size)
    ring.append(state, actions, rewards, next_state, dones, seat_idx_before, terminated_before, q_seat)
    s, a, r, s2, d = ring.sample(4096)
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import c_i, ptr


class TransitionRing:
    def __init__(self, device, capacity, obs_width):
        self.device = dev = _lib.require_cuda(device)
        self._lib = _lib.load()
        self._lib.pulse_replay_scratch_bytes.restype = C.c_size_t
        self.capacity, self.obs_width = int(capacity), int(obs_width)
        if self.capacity <= 0 or self.obs_width <= 0:
            raise ValueError("capacity and obs_width must be positive")
        self.state_buffer = torch.zeros((self.capacity, self.obs_width), dtype=torch.float32, device=dev)
        self.next_state_buffer = torch.zeros((self.capacity, self.obs_width), dtype=torch.float32, device=dev)
        self.action_buffer = torch.zeros(self.capacity, dtype=torch.int64, device=dev)
        self.reward_buffer = torch.zeros(self.capacity, dtype=torch.float32, device=dev)
        self.terminated_buffer = torch.zeros(self.capacity, dtype=torch.bool, device=dev)
        self.total = torch.zeros(1, dtype=torch.int64, device=dev)      # transitions appended so far (device side)
        self.appended = torch.zeros(1, dtype=torch.int64, device=dev)   # rows added by the last append
        self._scratch = None

    # position / size as ReplayBufferTorch keeps them; reading them synchronises with the stream
    @property
    def position(self):
        return int(self.total.item()) % self.capacity

    def __len__(self):
        return min(int(self.total.item()), self.capacity)

    def append(self, state, actions, rewards, next_state, dones, seat_idx_before, terminated_before=None, q_seat=0):
        n = state.shape[0]
        if n > self.capacity:
            raise ValueError("one step's batch must fit the ring (n_games <= capacity)")

        def as_(t, dtype):
            return t if t.dtype == dtype and t.is_contiguous() else t.to(dtype).contiguous()

        state, next_state = as_(state, torch.float32), as_(next_state, torch.float32)
        if tuple(state.shape) != (n, self.obs_width) or tuple(next_state.shape) != (n, self.obs_width):
            raise ValueError(f"state and next_state must have shape {(n, self.obs_width)}")
        actions, rewards = as_(actions, torch.int64), as_(rewards, torch.float32)
        dones, seat = as_(dones, torch.bool), as_(seat_idx_before, torch.int32)
        term = None if terminated_before is None else as_(terminated_before, torch.bool)
        need = int(self._lib.pulse_replay_scratch_bytes(C.c_longlong(n)))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_replay_append(
                ptr(state), ptr(next_state), c_i(self.obs_width), ptr(actions), ptr(rewards), ptr(dones), ptr(seat),
                ptr(term), c_i(int(q_seat)), ptr(self.state_buffer), ptr(self.next_state_buffer), ptr(self.action_buffer),
                ptr(self.reward_buffer), ptr(self.terminated_buffer), C.c_longlong(self.capacity), ptr(self.total),
                ptr(self.appended), ptr(self._scratch), C.c_longlong(n), _lib.cuda_stream(self.device)),
                "pulse_replay_append")

    def sample(self, batch_size):
        """ReplayBufferTorch.sample (utils/ReplayBuffer.py:62-71)."""
        idx = torch.randint(0, len(self), (batch_size,), device=self.device)
        return (self.state_buffer[idx], self.action_buffer[idx], self.reward_buffer[idx], self.next_state_buffer[idx],
                self.terminated_buffer[idx])
