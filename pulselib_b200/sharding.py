"""Env-sharded data parallelism: one process per GPU, a contiguous range of env ids each, no data-path
collective (no table, board or particle reads another's state).  The chance tape is keyed
by the GLOBAL env id (`env_base + i`), so a run's results do not depend on how many GPUs it was split over.
The only collective is a sum all-reduce of the run-statistics vector (NCCL over NVLink/NVSwitch; `gloo` in
the CPU tests), once per reporting interval, outside the step loop.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch

from . import _lib

STAT_NAMES = ("env_steps", "episodes_done", "reward_sum", "reward_sq_sum")


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """(first global env id, number of envs) of `rank`: contiguous, sizes differ by at most one."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(int(n_total), int(world))
    return rank * base + min(rank, extra), base + (1 if rank < extra else 0)


@dataclass
class Shard:
    rank: int
    world: int
    env_base: int
    n_envs: int


def init_from_env(n_total: int | None = None, n_per_gpu: int | None = None, backend: str | None = None) -> Shard:
    """Reads RANK / WORLD_SIZE / LOCAL_RANK (torchrun), initialises torch.distributed when world > 1 and
    returns this rank's shard: strong scaling when `n_total` is given, weak when `n_per_gpu` is."""
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not torch.distributed.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        torch.distributed.init_process_group(backend, **kw)
    if n_total is not None:
        start, count = shard_range(n_total, rank, world)
    else:
        start, count = rank * int(n_per_gpu), int(n_per_gpu)
    return Shard(rank, world, start, count)


class RunStats:
    """float64 [4] accumulator on `device` (see STAT_NAMES), fed by one small kernel per step and summed
    across ranks on demand."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.vec = torch.zeros(len(STAT_NAMES), dtype=torch.float64, device=self.device)

    def update(self, reward: torch.Tensor, done: torch.Tensor | None = None) -> None:
        if self.device.type != "cuda":
            raise _lib.PulseError("RunStats.update runs on the GPU (pulse_stats_accumulate)")
        if reward.dtype not in (torch.float32, torch.int32):
            raise TypeError("reward must be float32 or int32")
        import ctypes as C

        lib = _lib.load()
        with _lib.device_guard(self.device):
            _lib.check(lib.pulse_stats_accumulate(_lib.ptr(reward), C.c_int(reward.dtype == torch.float32), _lib.ptr(done),
                                                  _lib.ptr(self.vec), C.c_longlong(reward.numel()),
                                                  _lib.cuda_stream(self.device)), "pulse_stats_accumulate")

    def all_reduce(self) -> torch.Tensor:
        """Sum over ranks (no-op for a single process); returns the reduced copy, leaves `vec` local."""
        out = self.vec.clone()
        if torch.distributed.is_available() and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            torch.distributed.all_reduce(out, op=torch.distributed.ReduceOp.SUM)
        return out

    def summary(self, reduced: torch.Tensor | None = None) -> dict:
        v = (self.all_reduce() if reduced is None else reduced).tolist()
        d = dict(zip(STAT_NAMES, v))
        n = max(d["env_steps"], 1.0)
        d["reward_mean"] = d["reward_sum"] / n
        d["reward_var"] = max(d["reward_sq_sum"] / n - d["reward_mean"] ** 2, 0.0)
        return d
