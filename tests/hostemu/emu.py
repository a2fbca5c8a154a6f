"""TEST INFRASTRUCTURE ONLY: numpy-backed driver of the host-compiled kernel bodies (libhostemu.so)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import build as _build

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(str(_build.build()))
    return _LIB


def p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class StateStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "decks", "deck_positions", "board", "pots", "stages", "stacks", "hands", "current_round_bet",
        "total_invested", "status", "button", "sb", "bb", "idx", "highest", "agg", "acted", "last_raise_size",
        "prev_invested", "is_done", "equity_dirty", "equities", "hand_ranks")] + [("n_hand_ranks", C.c_int64)]


class ConfigStruct(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_games", "n_players", "active_players", "max_players", "starting_bbs",
                                         "max_bbs", "K", "alpha")] + [("w1", C.c_float), ("w2", C.c_float)]


N_FIELDS = ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg", "acted",
            "last_raise_size", "prev_invested")
NP_FIELDS = ("stacks", "current_round_bet", "total_invested", "status")


class EmuPoker:
    def __init__(self, hand_ranks, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000,
                 w1=.5, w2=.5, K=20, alpha=300):
        N, P = n_games, n_players
        self.n_games, self.n_players, self.max_players = N, P, max_players
        self.obs_size = 13 + (max_players - 1) * 3
        self.hand_ranks = np.ascontiguousarray(hand_ranks, dtype=np.int32)
        self.decks = np.zeros((N, 52), np.int32)
        self.board = np.full((N, 5), -1, np.int32)
        self.hands = np.full((N, P, 2), -1, np.int32)
        for f in N_FIELDS:
            setattr(self, f, np.zeros(N, np.int32))
        for f in NP_FIELDS:
            setattr(self, f, np.zeros((N, P), np.int32))
        self.is_done = np.zeros(N, np.uint8)
        self.equity_dirty = np.zeros(N, np.uint8)
        self._eq = np.zeros(N * P, np.float32)
        self.active_players = P
        self._first = True
        self.cfg = ConfigStruct(N, P, P, max_players, starting_bbs, max_bbs, K, alpha, w1, w2)
        self.obs = np.zeros((N, self.obs_size), np.float32)
        self.rewards = np.zeros(N, np.float32)

    @property
    def equities(self):
        return self._eq[: self.n_games * self.active_players].reshape(self.n_games, self.active_players)

    def _state(self):
        s = StateStruct()
        for n in ("decks", "board", "hands", "is_done", "equity_dirty") + N_FIELDS + NP_FIELDS:
            setattr(s, n, getattr(self, n).ctypes.data)
        s.equities = self._eq.ctypes.data
        s.hand_ranks = self.hand_ranks.ctypes.data
        s.n_hand_ranks = self.hand_ranks.size
        self.cfg.active_players = self.active_players
        return s

    def reset(self, decks=None, active_players=None, rotation=0, seed=0, env_base=0, hand=0):
        self.active_players = active_players or self.n_players
        s = self._state()
        d = None if decks is None else np.ascontiguousarray(decks, dtype=np.int32)
        lib().emu_poker_reset(C.byref(s), C.byref(self.cfg), p(d), C.c_uint64(seed), C.c_uint64(env_base),
                              C.c_uint32(hand), C.c_int(rotation), C.c_int(3 if self._first else 0), p(self.obs))
        self._first = False
        return self.obs

    def step(self, actions, phases=255, actor_idx=None):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        s = self._state()
        lib().emu_poker_step(C.byref(s), C.byref(self.cfg), p(a), p(actor_idx), C.c_uint32(phases), p(self.obs),
                             p(self.rewards), p(self.is_done))
        return self.obs, self.rewards, self.is_done.astype(bool)


class PackedStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("tbl", "tbl2", "stack", "crb", "inv", "hole", "ev", "hand_ctr", "hand_ranks")] + \
               [("n_hand_ranks", C.c_int64)]


class EmuPokerPacked(EmuPoker):
    """Packed-layout kernel bodies on the host; the reference-shaped arrays are refreshed by unpack
    after every call so the same comparison helpers apply."""

    def __init__(self, hand_ranks, auto_reset=False, seed=0, env_base=0, **kw):
        super().__init__(hand_ranks, **kw)
        N, P = self.n_games, self.n_players
        self.tbl = np.zeros((N, 4), np.uint32)
        self.tbl2 = np.zeros((N, 2), np.uint32)
        self.pstack = np.zeros((P, N), np.int32)
        self.pcrb = np.zeros((P, N), np.int32)
        self.pinv = np.zeros((P, N), np.int32)
        self.phole = np.zeros((P, N), np.uint16)
        self.hand_ctr = np.zeros(N, np.uint32)
        self.pev = np.zeros((P, N), np.uint64)
        self.auto_reset, self.seed, self.env_base = auto_reset, seed, env_base

    def _packed(self):
        s = PackedStruct()
        s.tbl, s.tbl2 = self.tbl.ctypes.data, self.tbl2.ctypes.data
        s.stack, s.crb, s.inv, s.hole = (a.ctypes.data for a in (self.pstack, self.pcrb, self.pinv, self.phole))
        s.hand_ctr = self.hand_ctr.ctypes.data
        s.ev = self.pev.ctypes.data
        s.hand_ranks, s.n_hand_ranks = self.hand_ranks.ctypes.data, self.hand_ranks.size
        return s

    def _unpack(self):
        s, c = self._packed(), self._state()
        lib().emu_poker_unpack(C.byref(s), C.byref(c), C.byref(self.cfg))

    def reset(self, decks=None, active_players=None, rotation=0, seed=None, env_base=None, hand=0):
        self.active_players = active_players or self.n_players
        self.cfg.active_players = self.active_players
        s = self._packed()
        d = None if decks is None else np.ascontiguousarray(decks, dtype=np.int32)
        lib().emu_poker_packed_reset(C.byref(s), C.byref(self.cfg), p(d),
                                     C.c_uint64(self.seed if seed is None else seed),
                                     C.c_uint64(self.env_base if env_base is None else env_base), C.c_uint32(hand),
                                     C.c_int(rotation), C.c_int(3 if self._first else 0), p(self.obs))
        self._first = False
        self._unpack()
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        s = self._packed()
        lib().emu_poker_packed_step(C.byref(s), C.byref(self.cfg), p(a), p(self.obs), p(self.rewards), p(self.is_done),
                                    C.c_int(self.auto_reset), C.c_uint64(self.seed), C.c_uint64(self.env_base))
        self._unpack()
        return self.obs, self.rewards, self.is_done.astype(bool)
