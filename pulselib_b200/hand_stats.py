"""Hand-level statistics accumulated on the device.

Replaces the per-step host bookkeeping of /root/reference/scripts/Poker/trainGPU_performance.py:186-206
(`newly_done = dones & ~terminated_before`, masked gathers of the Q seat's stacks, stages, positions, lists of
tensors) and the reductions of /root/reference/utils/performance.py (:55-59 positions, :104-121 BB/100 and its
95% lower bound, :135-167 episode summary, :170-174 stage buckets) with one small kernel per step that folds
the hands finished on that step into a 67-double vector (`pulse_poker_hand_stats`, csrc/stats.cu).  The same
vector is what env-sharded ranks sum with one all-reduce (NCCL on GPUs, gloo on CPU tensors).

    stats = HandStats(env, q_seat=0)
    obs, info = env.reset(...);  stats.begin_hand()
    ...  obs, rew, done, _, info = env.step(actions);  stats.update(done)
    stats.summary()   # mean_bb_delta, hand_win_rate, hands_completed, field_bb_per_100, lcb95_bb_per_100, ...
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import c_i, ptr

LEN = 67
CONFIDENCE_Z_95 = 1.959963984540054  # utils/performance.py:15
STAGE_BUCKETS = ("preflop", "flop", "turn", "river", "showdown")


def summarize(vec, active_players=None):
    """67-vector (tensor or list) -> the reference's summary numbers as Python floats."""
    v = [float(x) for x in (vec.tolist() if hasattr(vec, "tolist") else vec)]
    n, s, s2, wins = v[0], v[1], v[2], v[3]
    mean = s / n if n else 0.0
    var = max(s2 / n - mean * mean, 0.0) if n else 0.0          # std(unbiased=False)
    lcb = (mean - CONFIDENCE_Z_95 * math.sqrt(var) / math.sqrt(n)) * 100.0 if n > 1 else mean * 100.0
    ap = active_players or 16
    stage = {name: {"hands": int(v[4 + i]), "wins": int(v[14 + i]),
                    "bb_per_100": (v[9 + i] / v[4 + i] * 100.0) if v[4 + i] else 0.0,
                    "win_percentage": (v[14 + i] / n) if n else 0.0}          # calculate_street_win_percentages
             for i, name in enumerate(STAGE_BUCKETS)}
    pos = {f"position_{p}": {"hands": int(v[19 + p]), "wins": int(v[51 + p]),  # calculate_position_win_rates
                             "win_rate": (v[51 + p] / v[19 + p]) if v[19 + p] else 0.0,
                             "bb_per_100": (v[35 + p] / v[19 + p] * 100.0) if v[19 + p] else 0.0}
           for p in range(ap) if v[19 + p]}
    seat_vals = [m["bb_per_100"] for m in pos.values()]
    out = {
        "hands_completed": int(n), "mean_bb_delta": mean, "hand_win_rate": wins / n if n else 0.0,
        "field_bb_per_100": mean * 100.0, "lcb95_bb_per_100": lcb if n else 0.0,
        "seat_balanced_bb_per_100": sum(seat_vals) / len(seat_vals) if seat_vals else 0.0,
        "by_stage": stage, "by_position": pos,
    }
    return out


class HandStats:
    """Accumulates finished hands of one env (reference-shaped PokerGPU or PokerGPUPacked) for one Q seat."""

    def __init__(self, env, q_seat=0):
        self.env, self.q_seat = env, int(q_seat)
        self.device = _lib.require_cuda(env.device)
        self._lib = _lib.load()
        self.vec = torch.zeros(LEN, dtype=torch.float64, device=self.device)
        self.terminated = torch.zeros(env.n_games, dtype=torch.bool, device=self.device)
        self.initial = torch.zeros(env.n_games, dtype=torch.int32, device=self.device)

    def _engine(self):
        """The packed engine behind `env` while it is authoritative (PokerGPUPacked itself, or the drop-in PokerGPU
        in packed mode), else None: the reference-shaped tensors are read."""
        env = self.env
        if hasattr(env, "packed_engine"):
            return env.packed_engine()
        return env if hasattr(env, "_packed") else None

    def _q_stacks(self):
        eng = self._engine()
        if eng is not None:
            return eng._stack[self.q_seat]
        return self.env.stacks[:, self.q_seat]

    def begin_hand(self):
        """Call after env.reset(): remembers the Q seat's starting stacks, clears `terminated`
        (initial_q_stacks / terminated of trainGPU_performance.py:166, :171)."""
        self.initial.copy_(self._q_stacks())
        self.terminated.zero_()

    def update(self, done, rearm=None):
        """Fold the hands that finished on this step (done & ~terminated) into the vector; terminated |= done.
        rearm (default: the env's auto_reset): every done flag is a finished hand and the table's starting stack is
        re-armed for the hand the engine deals next."""
        if rearm is None:
            rearm = bool(getattr(self.env, "auto_reset", False))
        env, dev = self.env, self.device
        done = done if done.dtype == torch.bool and done.is_contiguous() else done.to(torch.bool).contiguous()
        eng = self._engine()
        with _lib.device_guard(dev):
            if eng is not None:
                cfg = eng._config()
                rc = self._lib.pulse_poker_packed_hand_stats(C.byref(eng._packed), C.byref(cfg), ptr(done),
                                                             ptr(self.terminated), ptr(self.initial), c_i(self.q_seat),
                                                             c_i(1 if rearm else 0), ptr(self.vec), _lib.cuda_stream(dev))
            else:
                stacks = env._tensor("stacks", torch.int32)
                stages = env._tensor("stages", torch.int32)
                button = env._tensor("button", torch.int32)
                rc = self._lib.pulse_poker_hand_stats(ptr(stacks), c_i(env.n_players), ptr(stages), ptr(button), ptr(done),
                                                      ptr(self.terminated), ptr(self.initial), c_i(self.q_seat),
                                                      c_i(int(env.active_players)), ptr(self.vec),
                                                      C.c_longlong(env.n_games), _lib.cuda_stream(dev))
        _lib.check(rc, "pulse_poker_hand_stats")

    def allreduce(self):
        """Sum the vector over all ranks (one collective of 67 doubles; outside the step loop)."""
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.vec, op=dist.ReduceOp.SUM)
        return self.vec

    def summary(self):
        return summarize(self.vec, int(self.env.active_players))
