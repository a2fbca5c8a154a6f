"""CUDA-graph capture of env step loops.

A step of every env here is a single kernel launch of a few tens of microseconds at the config sizes
(1M 2048 boards: ~20 us), so a Python loop that pays ~15 us of interpreter + ctypes time per step is
launch-bound.  `StepGraph` captures K consecutive `env.step(actions[k])` calls once and replays them as
one graph launch.  The env must not allocate inside `step` (the classes in pulselib_b200.envs do not when
handed contiguous tensors of the right dtype) and, if it draws chance from the tape, must keep its tape
counter on the device (`device_counter=True`) so every replay sees fresh draws.
"""
from __future__ import annotations

import torch


class StepGraph:
    def __init__(self, env, actions: torch.Tensor, warmup: int = 1, step=None):
        """actions: [K, N, ...] device tensor; the graph reads it in place on every replay, so refill it
        (e.g. `actions.random_(...)`, a policy writing into it) between replays."""
        self.env, self.actions = env, actions
        self.k = int(actions.shape[0])
        self._step = step or (lambda a: env.step(a))
        dev = actions.device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                      # warm-up off the capture stream (lazy init, allocs)
            for _ in range(warmup):
                self.out = self._step(actions[0])
        torch.cuda.current_stream(dev).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            for k in range(self.k):
                self.out = self._step(actions[k])
        self.launches_per_replay = self.k

    def replay(self):
        """Runs the K captured steps; returns the last step's outputs (the env's persistent tensors)."""
        self.graph.replay()
        return self.out
