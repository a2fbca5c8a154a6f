"""CPU suite: the oracle replays every env call the reference's OWN unit tests make.

tests/golden/poker_reftests.npz holds (state before, call, state after, return values) for each top-level
PokerGPU call in /root/reference/tests/poker/test_poker_gpu_*.py, recorded while those 136 tests PASSED on
the reference itself (tests/golden/record_reference_tests.py).  That makes them the reference's known-answer
tests for this path: pot-fraction bets, side pots, heads-up blinds, street openers, rotation, no-actor
rewards ...  The oracle must reproduce each one."""
import ctypes as C

import numpy as np

from oracle import pyoracle
from tests import poker_util as pu
from tests.hostemu.emu import EmuPoker, lib as emu_lib, p as emu_p


class OracleKat:
    def __init__(self, hr, ctor):
        self.o = pyoracle.OraclePoker(hr, **ctor)
        self.P, self.N = ctor["n_players"], ctor["n_games"]

    def set_state(self, before, active_players, rec):
        o = self.o
        o.active_players = active_players
        for f in pu.KAT_N + pu.KAT_NP + ("decks", "board", "hands", "is_done", "equity_dirty"):
            if f in before:
                dst = getattr(o, f)
                dst[...] = np.asarray(before[f]).reshape(dst.shape)
        if "equities" in before:
            # a test may shrink env.active_players after a reset at a larger table size: the reference keeps
            # its [N, AP_at_reset] tensor and only reads columns < active_players
            e = np.ascontiguousarray(before["equities"][:, :active_players])
            o._eq[: e.size] = e.reshape(-1)

    def get_state(self, name):
        if name == "active_players":
            return self.o.active_players
        return self.o.equities if name == "equities" else getattr(self.o, name)

    def reset(self, rec):
        o, a = self.o, rec.args
        # chance and table size are injected through the seams the reference has: the deck the reference
        # drew (or was given) and the table size it ended up with
        decks, ap = rec.after["decks"], int(rec.after["active_players"])
        rotation = int(a["opt_rotation"]) if int(a["opt_rotation"]) >= 0 else int(a["arg_rotation"])
        if not rec.first and not rec.had_button:
            o.button[...] = -1          # stacks were preset but no reset ran yet: button starts at 0 (:121)
        o._first = rec.first
        obs = o.reset(decks, active_players=ap, rotation=rotation)
        return {"obs": obs}

    def call(self, rec):
        o, m, a = self.o, rec.method, rec.args
        if m == "step":
            stub = "stub_equities" in a   # the test replaced calculate_equities on the instance: inject its result
            if stub:
                e = np.ascontiguousarray(a["stub_equities"][:, : o.active_players])
                o._eq[: e.size] = e.reshape(-1)
                o.equity_dirty[...] = a["stub_equity_dirty"]
            obs, rew, done = o.step(a["a0"], use_cached_equities=stub)
            return {"obs": obs, "reward": rew, "done": done.astype(np.uint8)}
        if m == "get_obs":
            o.call("get_obs", pyoracle.p(o.obs))
            return {"value": o.obs}
        if m == "execute_actions":
            acts = np.ascontiguousarray(a["a0"], dtype=np.int64)
            o.call("execute_actions", pyoracle.p(acts))
            return {}
        if m in ("calculate_equities", "resolve_fold_winners", "resolve_terminated_games"):
            o.call(m)
            return {}
        if m == "poker_reward_gpu":
            acts = np.ascontiguousarray(a["a0"], dtype=np.int64)
            actor = np.ascontiguousarray(a["a1"], dtype=np.int32)
            out = np.zeros(self.N, np.float32)
            o.call("reward", pyoracle.p(acts), pyoracle.p(actor), pyoracle.p(out))
            return {"value": out}
        if m == "post_blinds":            # PokerGPU.py:188-199, big blind only
            g, bb = np.arange(self.N), o.bb
            o.stacks[g, bb] -= 1; o.current_round_bet[g, bb] += 1; o.total_invested[g, bb] += 1; o.pots += 1
            o.status[g, bb] = np.where(o.stacks[g, bb] == 0, 2, 0)
            return {}
        if m == "deal_players_cards":     # :201-206
            n = int(a["a0"])
            pos = o.deck_positions[:, None] + np.arange(n)[None]
            cards = np.take_along_axis(o.decks, pos, 1)
            o.deck_positions += n
            return {"value": cards}
        if m == "deal_cards":             # :208-214
            g, n = np.asarray(a["a0"]).astype(np.int64), int(a["a1"])
            pos = o.deck_positions[g][:, None] + np.arange(n)[None]
            cards = o.decks[g[:, None], pos]
            o.deck_positions[g] += n
            return {"value": cards}
        raise AssertionError(f"unhandled reference method {m}")


class EmuKat(OracleKat):
    """The same replay through the CUDA kernel bodies compiled for the host (tests/hostemu): the code that
    runs on the GPU, one phase bit per reference method (pulselib_b200/envs/poker.py)."""

    def __init__(self, hr, ctor):
        self.o = EmuPoker(hr, **ctor)
        self.P, self.N = ctor["n_players"], ctor["n_games"]

    def reset(self, rec):
        o, a = self.o, rec.args
        o.active_players = int(rec.after["active_players"])
        rotation = int(a["opt_rotation"]) if int(a["opt_rotation"]) >= 0 else int(a["arg_rotation"])
        s = o._state()
        d = np.ascontiguousarray(rec.after["decks"], dtype=np.int32)
        flags = (1 if rec.first else 0) | (0 if rec.had_button else 2)
        emu_lib().emu_poker_reset(C.byref(s), C.byref(o.cfg), emu_p(d), C.c_uint64(0), C.c_uint64(0), C.c_uint32(0),
                                  C.c_int(0 if rec.first else rotation), C.c_int(flags), emu_p(o.obs))
        return {"obs": o.obs}

    def call(self, rec):
        o, m, a = self.o, rec.method, rec.args
        zeros = np.zeros(self.N, np.int64)
        if m == "step":
            stub = "stub_equities" in a
            if stub:
                e = np.ascontiguousarray(a["stub_equities"][:, : o.active_players])
                o._eq[: e.size] = e.reshape(-1)
                o.equity_dirty[...] = a["stub_equity_dirty"]
            obs, rew, done = o.step(a["a0"], phases=254 if stub else 255)
            return {"obs": obs, "reward": rew, "done": done.astype(np.uint8)}
        if m == "get_obs":
            o.step(zeros, phases=128)
            return {"value": o.obs}
        if m == "execute_actions":
            o.step(a["a0"], phases=2)
            return {}
        if m in ("calculate_equities", "resolve_fold_winners", "resolve_terminated_games"):
            o.step(zeros, phases={"calculate_equities": 1, "resolve_fold_winners": 8 | 256,
                                  "resolve_terminated_games": 16 | 256}[m])
            return {}
        if m == "poker_reward_gpu":
            actor = np.ascontiguousarray(a["a1"], dtype=np.int32)
            o.step(a["a0"], phases=64 | 512, actor_idx=actor)
            return {"value": o.rewards}
        return super().call(rec)   # post_blinds / deal_*: plain tensor indexing in the env class, not kernels


def test_recording_covers_the_reference_suite():
    recs = pu.load_reference_kat()
    z = np.load(pu.GOLDEN / "poker_reftests.npz")
    assert len(z["not_passed"]) == 0, "every reference test passed on the reference when recorded"
    assert len(recs) >= 280 and len({r.test for r in recs}) >= 120
    assert {r.method for r in recs} >= {"reset", "step", "execute_actions", "resolve_terminated_games",
                                        "resolve_fold_winners", "calculate_equities", "poker_reward_gpu", "get_obs"}


def test_oracle_reproduces_reference_unit_tests(hr_table):
    n = pu.replay_reference_kat(lambda ctor: OracleKat(hr_table, ctor))
    assert n >= 280


def test_kernel_bodies_hostbuild_reproduce_reference_unit_tests(hr_table):
    n = pu.replay_reference_kat(lambda ctor: EmuKat(hr_table, ctor))
    assert n >= 280


def test_dropin_class_reproduces_reference_unit_tests_on_host_bodies():
    """The registered drop-in class (packed engine, mirror on access) replays all records too -- here on the
    host-compiled kernel bodies, in tests/test_gpu_reference_kat.py on the B200."""
    import torch

    from tests.hostemu.shim import hostemu_backend
    from tests.kat_dropin import DropinKat

    with hostemu_backend():
        n = pu.replay_reference_kat(lambda ctor: DropinKat(torch.device("cpu"), ctor))
    assert n >= 280
