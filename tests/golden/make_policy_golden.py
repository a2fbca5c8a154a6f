"""Generates tests/golden/poker_policy.npz from the reference's OWN scripted agents and build_actions
(build container only: needs /root/reference).

/root/reference/environments/Poker/Player.py (HeuristicHandsPlayerGPU, TightAggressivePlayerGPU,
LoosePassivePlayerGPU, SmallBallPlayerGPU) and /root/reference/environments/Poker/utils.py::build_actions
are imported in place and run on CPU.  Their random draws are injected through the seam they already have
-- the torch.randint / torch.rand calls -- from a per-table chance pair (w0, w1):

    torch.randint(lo, hi, (n,)) -> lo + mulhi32(w0, hi - lo)      torch.rand(n) -> (w1 >> 8) * 2^-24

Two kinds of cases are stored: each agent class called directly on a batch of observations, and whole
build_actions calls with mixed seat types (including a stand-in Q agent whose actions must be left alone).
The masks inside build_actions compact the batch (`state[mask]`), so the patched draws are compacted with
the same mask -- each table still sees its own pair.

    python tests/golden/make_policy_golden.py
"""
from __future__ import annotations

import sys
import types
from pathlib import Path
from unittest import mock

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "poker_policy.npz"
KIND = {"external": 0, "random": 1, "heuristic_hands": 2, "tight_aggressive": 3, "loose_passive": 4, "small_ball": 5}


def load_reference():
    from pulselib_b200 import gym_compat

    gym_compat.install_shim()
    ev = types.ModuleType("eval7")
    ev.Card = lambda s: None
    ev.evaluate = lambda cards: 0
    ev.HandRange = lambda s: None
    ev.Deck = lambda: None
    sys.modules["eval7"] = ev
    for name, rel in (("environments", "environments"), ("environments.Poker", "environments/Poker")):
        m = types.ModuleType(name)
        m.__path__ = [str(REF / rel)]
        sys.modules[name] = m
    sys.path.insert(0, str(REF))
    import environments.Poker.Player as player
    import environments.Poker.utils as utils

    return player, utils


class Draws:
    """Stands in for torch.randint / torch.rand inside the reference's agent code."""

    def __init__(self, chance2):
        self.c = chance2.astype(np.uint64)
        self.mask = None  # the rows the next call sees (None = all)

    def rows(self, n):
        c = self.c if self.mask is None else self.c[self.mask]
        assert c.shape[0] == n, (c.shape, n)
        return c

    def randint(self, lo, hi, size, device=None, **kw):
        c = self.rows(size[0])
        return torch.from_numpy((lo + ((c[:, 0] * np.uint64(hi - lo)) >> np.uint64(32))).astype(np.int64))

    def rand(self, n, device=None, **kw):
        c = self.rows(n)
        return torch.from_numpy(((c[:, 1] >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)))


def random_obs(rng, n, width, n_players):
    obs = np.zeros((n, width), np.float32)
    obs[:, 0:5] = rng.integers(-1, 53, (n, 5))
    obs[:, 5:7] = rng.integers(1, 53, (n, 2))
    obs[: n // 50, 5:7] = -1                      # sit-out rows carry -1 hole cards
    obs[:, 7] = rng.integers(0, 5, n)
    obs[:, 8] = rng.integers(0, n_players, n)
    obs[:, 9] = rng.integers(0, 200, n)           # pot: both SmallBall thresholds (30, 80) are crossed
    obs[:, 10:13] = rng.integers(0, 100, (n, 3))
    return obs


def main():
    player, utils = load_reference()
    T = utils.PokerAgentType
    dev = torch.device("cpu")
    classes = {"heuristic_hands": player.HeuristicHandsPlayerGPU, "tight_aggressive": player.TightAggressivePlayerGPU,
               "loose_passive": player.LoosePassivePlayerGPU, "small_ball": player.SmallBallPlayerGPU}
    rng = np.random.default_rng(20260401)
    rec = {}
    # 1. each agent class on its own
    n = 20000
    obs = random_obs(rng, n, 28, 6)
    chance = rng.integers(0, 2 ** 32, (n, 2), dtype=np.uint64).astype(np.uint32)
    chance[:200, 1] = np.uint32(0xFFFFFFFF)       # LoosePassive: probs just below 1
    chance[200:400, 1] = np.uint32(int(0.9 * 2 ** 32))
    rec["direct/obs"], rec["direct/chance2"] = obs, chance
    for name, cls in classes.items():
        d = Draws(chance)
        with mock.patch.object(torch, "randint", d.randint), mock.patch.object(torch, "rand", d.rand):
            acts = cls(100, 0, dev).action(torch.from_numpy(obs))
        rec[f"direct/{name}"] = acts.numpy().astype(np.int64)
        print(name, np.bincount(rec[f"direct/{name}"], minlength=13))
    # 2. build_actions with mixed tables
    tables = [
        ("mixed6", 6, ["qlearning", "random", "heuristic_hands", "tight_aggressive", "loose_passive", "small_ball"]),
        ("hu", 2, ["qlearning", "tight_aggressive"]),
        ("scripted4", 4, ["small_ball", "small_ball", "random", "loose_passive"]),
    ]

    class FakeQ:                                   # the Q agent: some deterministic function of the row
        def get_actions(self, states):
            return (states[:, 9].long() + states[:, 5].long()) % 13

    for tag, P, names in tables:
        n = 12000
        W = 13 + 3 * (P - 1)
        obs = random_obs(rng, n, W, P)
        chance = rng.integers(0, 2 ** 32, (n, 2), dtype=np.uint64).astype(np.uint32)
        seat = rng.integers(0, P, n).astype(np.int32)
        types_ = [T(x) for x in names]
        agents = [FakeQ() if x == "qlearning" else (classes[x](100, i, dev) if x in classes else None)
                  for i, x in enumerate(names)]
        actions = torch.zeros(n, dtype=torch.long)
        d = Draws(chance)
        real_zeros_like = torch.zeros_like
        state_t, seat_t = torch.from_numpy(obs), torch.from_numpy(seat)

        # build_actions calls agent.action(state[mask]): tell the draw stand-in which rows that is
        orig_actions = {}
        for i, a in enumerate(agents):
            if a is not None and hasattr(a, "action"):
                def wrap(states, _a=a, _i=i):
                    kinds = [k for k, t in enumerate(types_) if t == types_[_i]]
                    d.mask = np.isin(seat, kinds)
                    try:
                        return type(_a).action(_a, states)
                    finally:
                        d.mask = None
                a.action = wrap
        with mock.patch.object(torch, "randint", d.randint), mock.patch.object(torch, "rand", d.rand):
            utils.build_actions(state_t, actions, seat_t, agents, types_, dev)
        rec[f"{tag}/obs"], rec[f"{tag}/chance2"], rec[f"{tag}/seat_idx"] = obs, chance, seat
        rec[f"{tag}/kinds"] = np.array([KIND.get(x, 0) if x != "qlearning" else 0 for x in names], np.uint8)
        rec[f"{tag}/q_actions"] = FakeQ().get_actions(state_t).numpy().astype(np.int64)
        rec[f"{tag}/actions"] = actions.numpy().astype(np.int64)
        print(tag, np.bincount(rec[f"{tag}/actions"], minlength=13))
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, OUT.stat().st_size)


if __name__ == "__main__":
    main()
