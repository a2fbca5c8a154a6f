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
               [("n_hand_ranks", C.c_int64), ("idx", C.c_void_p), ("prev_invested", C.c_void_p)]


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
        self.live_idx = np.zeros(N, np.int32)
        self.live_prev_inv = np.zeros(N, np.int32)
        self._prefixed, self._stepped = None, False

    def _packed(self):
        s = PackedStruct()
        s.tbl, s.tbl2 = self.tbl.ctypes.data, self.tbl2.ctypes.data
        s.stack, s.crb, s.inv, s.hole = (a.ctypes.data for a in (self.pstack, self.pcrb, self.pinv, self.phole))
        s.hand_ctr = self.hand_ctr.ctypes.data
        s.ev = self.pev.ctypes.data
        s.hand_ranks, s.n_hand_ranks = self.hand_ranks.ctypes.data, self.hand_ranks.size
        s.idx, s.prev_invested = self.live_idx.ctypes.data, self.live_prev_inv.ctypes.data
        return s

    def _unpack(self):
        """the full mirror (pulse_poker_unpack_full): every reference attribute, decks included"""
        s, c = self._packed(), self._state()
        lib().emu_poker_unpack_full(C.byref(s), C.byref(c), C.byref(self.cfg), p(self._prefixed), C.c_uint64(self.seed),
                                    C.c_uint64(self.env_base), C.c_int(0 if self._stepped else 1))

    def reset(self, decks=None, active_players=None, rotation=0, seed=None, env_base=None, hand=0):
        self.active_players = active_players or self.n_players
        self.cfg.active_players = self.active_players
        s = self._packed()
        d = None if decks is None else np.ascontiguousarray(decks, dtype=np.int32)
        if seed is not None:
            self.seed = seed
        if env_base is not None:
            self.env_base = env_base
        lib().emu_poker_packed_reset(C.byref(s), C.byref(self.cfg), p(d), C.c_uint64(self.seed),
                                     C.c_uint64(self.env_base), C.c_uint32(hand),
                                     C.c_int(rotation), C.c_int(7 if self._first else 0), p(self.obs))
        self._first = False
        self._prefixed, self._stepped = d, False
        self._unpack()
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        s = self._packed()
        lib().emu_poker_packed_step(C.byref(s), C.byref(self.cfg), p(a), p(self.obs), p(self.rewards), p(self.is_done),
                                    C.c_int(self.auto_reset), C.c_uint64(self.seed), C.c_uint64(self.env_base))
        self._stepped = True
        self._unpack()
        return self.obs, self.rewards, self.is_done.astype(bool)


    def pack_from_mirror(self):
        """csrc/poker_packed_body.cuh::packed_pack_one on the host: the reference-shaped arrays of this object -> the
        packed arrays; returns the number of tables the packed layout could not express."""
        s, c = self._packed(), self._state()
        return int(lib().emu_poker_pack(C.byref(c), C.byref(s), C.byref(self.cfg)))

    def showdown_equity(self, samples=256, counter=0, lanes=1):
        """csrc/poker_packed_body.cuh::packed_showdown_equity_part on the host, boards split over `lanes` workers"""
        out = np.zeros((self.n_games, self.active_players), np.float32)
        s = self._packed()
        lib().emu_showdown_equity(C.byref(s), C.byref(self.cfg), p(out), C.c_int(samples), C.c_uint64(self.seed),
                                  C.c_uint64(self.env_base), C.c_uint32(counter), C.c_int(lanes))
        return out

    def wire_rows(self):
        """the compact step record (csrc/poker_wire.cuh) of every table, reward = self.rewards"""
        words = 3 + (13 + 3 * (self.max_players - 1) - 8 + 1) // 2
        rows = np.zeros((self.n_games, words), np.uint32)
        s = self._packed()
        rc = lib().emu_poker_wire_rows(C.byref(s), C.byref(self.cfg), p(self.rewards), p(rows))
        assert rc == 0
        return rows

    def step_policy(self, kinds, chance2=None, counter=0, actions=None):
        """packed step with the scripted seats evaluated inside; returns the played actions too"""
        k = np.ascontiguousarray(kinds, dtype=np.uint8)
        a = None if actions is None else np.ascontiguousarray(actions, dtype=np.int64)
        c = None if chance2 is None else np.ascontiguousarray(chance2, dtype=np.uint32)
        chosen = np.zeros(self.n_games, np.int64)
        s = self._packed()
        lib().emu_poker_packed_step_policy(C.byref(s), C.byref(self.cfg), p(a), p(k), p(c), C.c_uint32(counter),
                                           p(chosen), p(self.obs), p(self.rewards), p(self.is_done),
                                           C.c_int(self.auto_reset), C.c_uint64(self.seed), C.c_uint64(self.env_base))
        self._unpack()
        return self.obs, self.rewards, self.is_done.astype(bool), chosen


def emu_policy(obs, seat_idx, kinds, actions, chance2=None, seed=0, env_base=0, counter=0):
    """pulse_policy_action over observation rows (the body of pulse_poker_policy), in place on `actions`."""
    obs = np.ascontiguousarray(obs, dtype=np.float32)
    seat = np.ascontiguousarray(seat_idx, dtype=np.int32)
    k = np.ascontiguousarray(kinds, dtype=np.uint8)
    c = None if chance2 is None else np.ascontiguousarray(chance2, dtype=np.uint32)
    assert actions.dtype == np.int64 and actions.flags.c_contiguous
    lib().emu_poker_policy(p(obs), C.c_int(obs.shape[1]), p(seat), p(k), C.c_int(len(k)), p(c), C.c_uint64(seed),
                           C.c_uint64(env_base), C.c_uint32(counter), p(actions), C.c_longlong(obs.shape[0]))
    return actions


# ---- Blackjack / TFE / Tetris bodies -----------------------------------------------------------------
BJ_COMPAT = (("decks", np.int64, 52), ("deck_positions", np.int32, 0), ("players_cards", np.int32, 20),
             ("players_card_idx", np.int32, 0), ("player_card_sums", np.int32, 0), ("has_ace", np.uint8, 0),
             ("dealer_cards", np.int32, 20), ("dealer_card_idx", np.int32, 0), ("dealer_upcard", np.int32, 0),
             ("dealer_card_sums", np.int32, 0), ("dealer_has_ace", np.uint8, 0))


class BjCompatStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _, _ in BJ_COMPAT]


class EmuBlackjack:
    def __init__(self, n, seed=0, env_base=0):
        self.L, self.n, self.seed, self.env_base = lib(), int(n), seed, env_base
        self.st = np.zeros(n, np.uint32)
        self.deck = np.zeros((52, n), np.uint8)
        self.obs = np.zeros((n, 3), np.int32)
        self.rewards = np.zeros(n, np.int32)
        self.terminated = np.zeros(n, np.uint8)

    def reset(self, decks=None, episode=0):
        inj = np.ascontiguousarray(decks, dtype=np.uint8) if decks is not None else None
        self.L.emu_blackjack_reset(p(self.st), p(self.deck), p(inj), C.c_uint64(self.seed), C.c_uint64(self.env_base),
                                   C.c_uint32(episode), p(self.obs), p(self.rewards), p(self.terminated), C.c_int(self.n))
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        self.L.emu_blackjack_step(p(self.st), p(self.deck), p(a), p(self.obs), p(self.rewards), p(self.terminated),
                                  C.c_int(self.n))
        return self.obs, self.rewards, self.terminated.astype(bool)

    def unpack(self):
        out, s = {}, BjCompatStruct()
        for name, dt, w in BJ_COMPAT:
            out[name] = np.zeros((self.n, w) if w else (self.n,), dt)
            setattr(s, name, out[name].ctypes.data)
        self.L.emu_blackjack_unpack(p(self.st), p(self.deck), C.byref(s), C.c_int(self.n))
        return out


class EmuTFE:
    def __init__(self, n, side=4, seed=0, env_base=0):
        self.L, self.n, self.side, self.seed, self.env_base = lib(), int(n), int(side), seed, env_base
        self.packed = np.zeros(n, np.uint64)
        self.boards = np.zeros((n, side, side), np.int32)
        self.scores = np.zeros(n, np.int64)
        self.rewards = np.zeros(n, np.int32)
        self.done = np.zeros(n, np.uint8)

    def reset(self, chance4=None, counter=0):
        c = np.ascontiguousarray(chance4, dtype=np.uint32) if chance4 is not None else None
        self.L.emu_tfe_reset(p(self.packed), p(self.scores), p(self.done), p(c), C.c_uint64(self.seed),
                             C.c_uint64(self.env_base), C.c_uint32(counter), p(self.boards), C.c_int(self.n), C.c_int(self.side))
        return self.boards

    def set_boards(self, values):
        v = np.ascontiguousarray(values, dtype=np.int32)
        bad = self.L.emu_tfe_pack(p(v), p(self.packed), C.c_int(self.n), C.c_int(self.side))
        assert bad == 0

    def step(self, actions, chance2=None, auto_reset=False, reset_chance4=None, counter=0):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        c = np.ascontiguousarray(chance2, dtype=np.uint32) if chance2 is not None else None
        rc = np.ascontiguousarray(reset_chance4, dtype=np.uint32) if reset_chance4 is not None else None
        self.L.emu_tfe_step(p(self.packed), p(a), p(c), p(rc), C.c_uint64(self.seed), C.c_uint64(self.env_base),
                            C.c_uint32(counter), C.c_int(bool(auto_reset)), p(self.boards), p(self.rewards), p(self.done),
                            p(self.scores), C.c_int(self.n), C.c_int(self.side))
        return self.boards, self.rewards, self.done.astype(bool)


class EmuTetris:
    def __init__(self, n, seed=0, env_base=0):
        self.L, self.n, self.seed, self.env_base = lib(), int(n), seed, env_base
        self.rows = np.zeros((n, 20), np.uint16)
        self.piece = np.zeros(n, np.uint8)
        self.lines = np.zeros(n, np.int32)
        self.done = np.zeros(n, np.uint8)
        self.rewards = np.zeros(n, np.int32)

    def reset(self, first_piece):
        self.rows[...] = 0
        self.piece[...] = first_piece
        self.lines[...] = 0
        self.done[...] = 0
        return self.rows

    def step(self, actions, next_piece=None, auto_reset=False, reset_piece=None, counter=0):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        nx = np.ascontiguousarray(next_piece, dtype=np.uint8) if next_piece is not None else None
        rp = np.ascontiguousarray(reset_piece, dtype=np.uint8) if reset_piece is not None else None
        self.L.emu_tetris_step(p(self.rows), p(self.piece), p(self.lines), p(self.done), p(a), p(nx), p(rp),
                               C.c_uint64(self.seed), C.c_uint64(self.env_base), C.c_uint32(counter),
                               C.c_int(bool(auto_reset)), p(self.rewards), C.c_int(self.n))
        return self.rows, self.rewards, self.done.astype(bool)
