"""TEST INFRASTRUCTURE ONLY.  ctypes front-end of the C oracle (oracle/*.c).  Imported by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs; never by pulselib_b200."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_build" / "liboracle.so"
_LIB = None


def build(force: bool = False) -> Path:
    srcs = list(_DIR.glob("*.c")) + list(_DIR.glob("*.h"))
    if force or not _SO.exists() or any(s.stat().st_mtime > _SO.stat().st_mtime for s in srcs):
        r = subprocess.run(["make", "-C", str(_DIR)] + (["-B"] if force else []), capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"oracle build failed:\n{r.stdout}\n{r.stderr}")
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        build()
        _LIB = C.CDLL(str(_SO))
        _LIB.oracle_check_table7.restype = C.c_longlong
    return _LIB


def p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class PokerStruct(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_games", "n_players", "active_players", "max_players",
                                         "starting_bbs", "max_bbs", "K", "alpha")] + \
               [("w1", C.c_float), ("w2", C.c_float)] + \
               [(n, C.c_void_p) for n in ("decks", "deck_positions", "board", "pots", "stages", "stacks", "hands",
                                          "current_round_bet", "total_invested", "status", "button", "sb", "bb",
                                          "idx", "highest", "agg", "acted", "last_raise_size", "prev_invested",
                                          "is_done", "equity_dirty", "equities", "hand_ranks")] + \
               [("n_hand_ranks", C.c_int64)]


INT_FIELDS_N = ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg", "acted",
                "last_raise_size", "prev_invested")
INT_FIELDS_NP = ("stacks", "current_round_bet", "total_invested", "status")


class OraclePoker:
    """Reference-shaped PokerGPU on numpy arrays, stepped by oracle_poker.c."""

    def __init__(self, hand_ranks, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000,
                 w1=.5, w2=.5, K=20, alpha=300):
        self.L = lib()
        N, P = n_games, n_players
        self.n_games, self.n_players, self.max_players = N, P, max_players
        self.obs_size = 13 + (max_players - 1) * 3
        self.hand_ranks = np.ascontiguousarray(hand_ranks, dtype=np.int32)
        self.decks = np.zeros((N, 52), np.int32)
        self.board = np.full((N, 5), -1, np.int32)
        self.hands = np.full((N, P, 2), -1, np.int32)
        for f in INT_FIELDS_N:
            setattr(self, f, np.zeros(N, np.int32))
        for f in INT_FIELDS_NP:
            setattr(self, f, np.zeros((N, P), np.int32))
        self.is_done = np.zeros(N, np.uint8)
        self.equity_dirty = np.zeros(N, np.uint8)
        self._eq = np.zeros(N * P, np.float32)
        self.active_players = P
        self._first = True
        self.s = PokerStruct()
        for k, v in dict(n_games=N, n_players=P, active_players=P, max_players=max_players,
                         starting_bbs=starting_bbs, max_bbs=max_bbs, K=K, alpha=alpha, w1=w1, w2=w2).items():
            setattr(self.s, k, v)
        self._sync()
        self.obs = np.zeros((N, self.obs_size), np.float32)
        self.rewards = np.zeros(N, np.float32)

    @property
    def equities(self):
        return self._eq[: self.n_games * self.active_players].reshape(self.n_games, self.active_players)

    def _sync(self):
        for n in ("decks", "board", "hands", "is_done", "equity_dirty") + INT_FIELDS_N + INT_FIELDS_NP:
            setattr(self.s, n, getattr(self, n).ctypes.data)
        self.s.equities = self._eq.ctypes.data
        self.s.hand_ranks = self.hand_ranks.ctypes.data
        self.s.n_hand_ranks = self.hand_ranks.size
        self.s.active_players = self.active_players

    def tape_decks(self, seed, env_base, hand):
        self._sync()
        self.L.oracle_poker_fill_tape_decks(C.byref(self.s), C.c_uint64(seed), C.c_uint64(env_base), C.c_uint32(hand))

    def reset(self, decks=None, active_players=None, rotation=0):
        if decks is not None:
            self.decks[...] = decks
        self.active_players = active_players or self.n_players
        self._sync()
        self.L.oracle_poker_reset(C.byref(self.s), C.c_int(self.active_players), C.c_int(self._first),
                                  C.c_int(rotation))
        self._first = False
        self.L.oracle_poker_get_obs(C.byref(self.s), p(self.obs))
        return self.obs

    def step(self, actions, use_cached_equities=False):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        self._sync()
        self.L.oracle_poker_step(C.byref(self.s), p(a), p(self.obs), p(self.rewards), C.c_int(use_cached_equities))
        return self.obs, self.rewards, self.is_done.astype(bool)

    def step_fast(self, actions_i64):
        """step without re-syncing pointers or building obs (CPU-baseline timing loop)."""
        self.L.oracle_poker_step(C.byref(self.s), p(actions_i64), p(self.obs), p(self.rewards), C.c_int(0))

    def call(self, name, *args):
        self._sync()
        getattr(self.L, "oracle_poker_" + name)(C.byref(self.s), *args)


def poker_policy(obs, seat_idx, kinds, chance2, actions):
    """oracle_poker_policy: the reference's scripted agents + build_actions, in place on `actions`."""
    obs = np.ascontiguousarray(obs, dtype=np.float32)
    seat = np.ascontiguousarray(seat_idx, dtype=np.int32)
    k = np.ascontiguousarray(kinds, dtype=np.uint8)
    c = np.ascontiguousarray(chance2, dtype=np.uint32)
    assert actions.dtype == np.int64 and actions.flags.c_contiguous
    lib().oracle_poker_policy(p(obs), C.c_int(obs.shape[1]), p(seat), p(k), C.c_int(len(k)), p(c), p(actions),
                              C.c_int(obs.shape[0]))
    return actions


def policy_tape(seed, env_base, counter, n):
    out = np.zeros((n, 2), np.uint32)
    lib().oracle_policy_tape(C.c_uint64(seed), C.c_uint64(env_base), C.c_uint32(counter), p(out), C.c_int(n))
    return out


def enumerated_strength(stages, hands, board, active_players):
    """oracle_enumerated_strength_batch: brute-force mean river strength over the unseen board cards, [n, ap]."""
    st = np.ascontiguousarray(stages, dtype=np.int32)
    h = np.ascontiguousarray(hands, dtype=np.int32)
    b = np.ascontiguousarray(board, dtype=np.int32)
    out = np.zeros((len(st), int(active_players)), np.float32)
    lib().oracle_enumerated_strength_batch(p(st), p(h), p(b), C.c_int(h.shape[1]), C.c_int(int(active_players)),
                                           C.c_longlong(len(st)), p(out))
    return out


def showdown_equity(hands, status, board, active_players, samples=256, seed=0, env_base=0, counter=0, hand_ranks=None):
    """oracle_showdown_equity_batch: all-in equity of every seat with all hole cards known, [n, ap] float32.
    hands [n, P, 2], status [n, P], board [n, 5] (-1 = not dealt yet); hand_ranks: the HandRanks table (the reference's
    own showdown evaluation, a table walk) or None for the brute-force evaluator."""
    h = np.ascontiguousarray(hands, dtype=np.int32)
    st = np.ascontiguousarray(status, dtype=np.int32)
    b = np.ascontiguousarray(board, dtype=np.int32)
    hr = None if hand_ranks is None else np.ascontiguousarray(hand_ranks, dtype=np.int32)
    out = np.zeros((h.shape[0], int(active_players)), np.float32)
    lib().oracle_showdown_equity_batch(p(hr), p(h), p(st), p(b), C.c_int(h.shape[1]), C.c_int(int(active_players)),
                                       C.c_longlong(h.shape[0]), C.c_int(int(samples)), C.c_uint64(seed),
                                       C.c_uint64(env_base), C.c_uint32(counter), p(out))
    return out


def showdown_equity_preflop_exact(hands, status, hand_ranks):
    """Full enumeration of C(52 - 2 ap, 5) boards for ONE table (hands [ap, 2], status [ap]) -> float64 [ap]."""
    h = np.ascontiguousarray(hands, dtype=np.int32)
    st = np.ascontiguousarray(status, dtype=np.int32)
    hr = np.ascontiguousarray(hand_ranks, dtype=np.int32)
    out = np.zeros(h.shape[0], np.float64)
    lib().oracle_showdown_equity_preflop_exact(p(hr), p(h), p(st), C.c_int(h.shape[0]), p(out))
    return out


def hand_stats_update(vec, q_stacks, initial, stages, button, done, terminated, q_seat, active_players):
    """Numpy restatement of one step of the reference's hand bookkeeping
    (/root/reference/scripts/Poker/trainGPU_performance.py:186-206 + utils/performance.py:55-59, :170-174) folded into
    the 67-double layout of include/pulse_envs.h.  `terminated` (bool array) is updated in place."""
    done = np.asarray(done).astype(bool)
    newly = done & ~terminated                                   # dones & ~terminated_before
    terminated |= done
    d = (np.asarray(q_stacks)[newly].astype(np.float64) - np.asarray(initial)[newly].astype(np.float64))
    bucket = np.where(np.asarray(stages)[newly] >= 4, 4, np.clip(np.asarray(stages)[newly], 0, 3))
    pos = np.remainder(q_seat - np.asarray(button)[newly].astype(np.int64), active_players)
    win = d > 0
    vec[0] += d.size; vec[1] += d.sum(); vec[2] += (d * d).sum(); vec[3] += win.sum()
    for b in range(5):
        m = bucket == b
        vec[4 + b] += m.sum(); vec[9 + b] += d[m].sum(); vec[14 + b] += (win & m).sum()
    for p_ in range(16):
        m = pos == p_
        vec[19 + p_] += m.sum(); vec[35 + p_] += d[m].sum(); vec[51 + p_] += (win & m).sum()
    return vec


class OracleRing:
    """Numpy restatement of the reference's transition hand-off: the masked gathers of
    /root/reference/scripts/Poker/trainGPU.py:79-94 followed by ReplayBufferTorch.add
    (/root/reference/utils/ReplayBuffer.py:47-60), with float32 state/reward buffers."""

    def __init__(self, capacity, obs_width):
        self.capacity = capacity
        self.state_buffer = np.zeros((capacity, obs_width), np.float32)
        self.next_state_buffer = np.zeros((capacity, obs_width), np.float32)
        self.action_buffer = np.zeros(capacity, np.int64)
        self.reward_buffer = np.zeros(capacity, np.float32)
        self.terminated_buffer = np.zeros(capacity, bool)
        self.position = self.size = self.total = 0

    def append(self, state, actions, rewards, next_state, dones, seat, term_before, q_seat):
        active = (np.asarray(seat) == q_seat)
        if term_before is not None:
            active &= ~np.asarray(term_before).astype(bool)
        k = int(active.sum())
        idx = np.arange(self.position, self.position + k) % self.capacity
        self.state_buffer[idx] = state[active]; self.next_state_buffer[idx] = next_state[active]
        self.action_buffer[idx] = actions[active]; self.reward_buffer[idx] = rewards[active]
        self.terminated_buffer[idx] = np.asarray(dones)[active].astype(bool)
        self.position = (self.position + k) % self.capacity
        self.size = min(self.size + k, self.capacity)
        self.total += k
        return k


# ---- Blackjack -----------------------------------------------------------------------------------
BJ_FIELDS = ("decks", "deck_positions", "terminated", "players_cards", "players_card_idx", "player_card_sums",
             "has_ace", "dealer_cards", "dealer_card_idx", "dealer_upcard", "dealer_card_sums", "dealer_has_ace",
             "rewards", "obs")
BJ_U8 = ("terminated", "has_ace", "dealer_has_ace")


class BlackjackStruct(C.Structure):
    _fields_ = [("n", C.c_int32)] + [(n, C.c_void_p) for n in BJ_FIELDS]


class OracleBlackjack:
    """Reference-shaped BlackJack on numpy arrays, stepped by oracle_blackjack.c."""

    def __init__(self, batch_size):
        self.L = lib()
        N = self.batch_size = int(batch_size)
        shapes = dict(decks=(N, 52), players_cards=(N, 20), dealer_cards=(N, 20), obs=(N, 3))
        self.s = BlackjackStruct()
        self.s.n = N
        for f in BJ_FIELDS:
            arr = np.zeros(shapes.get(f, (N,)), np.uint8 if f in BJ_U8 else np.int32)
            setattr(self, f, arr)
            setattr(self.s, f, arr.ctypes.data)

    def tape_decks(self, seed, env_base, episode):
        self.L.oracle_blackjack_tape_decks(C.c_uint64(seed), C.c_uint64(env_base), C.c_uint32(episode),
                                           C.c_int(self.batch_size), p(self.decks))
        return self.decks

    def reset(self, decks=None):
        if decks is not None:
            self.decks[...] = decks
        self.L.oracle_blackjack_reset(C.byref(self.s))
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        self.L.oracle_blackjack_step(C.byref(self.s), p(a))
        return self.obs, self.rewards, self.terminated.astype(bool)


# ---- 2048 / TFE ----------------------------------------------------------------------------------
STREAM_TFE_SPAWN, STREAM_TFE_RESET = 3, 4


class OracleTFE:
    """Batch of reference-rule 2048 boards (int32 tile values), stepped by oracle_tfe.c."""

    def __init__(self, n_envs, side=4):
        self.L = lib()
        self.n_envs, self.side = int(n_envs), int(side)
        self.boards = np.zeros((self.n_envs, side, side), np.int32)
        self.scores = np.zeros(self.n_envs, np.int64)
        self.rewards = np.zeros(self.n_envs, np.int32)
        self.done = np.zeros(self.n_envs, np.uint8)

    def tape(self, seed, env_base, stream, counter, words):
        out = np.zeros((self.n_envs, words), np.uint32)
        self.L.oracle_tfe_tape_chance(C.c_uint64(seed), C.c_uint64(env_base), C.c_uint32(stream), C.c_uint32(counter),
                                      C.c_int(self.n_envs), C.c_int(words), p(out))
        return out

    def reset(self, chance4):
        c = np.ascontiguousarray(chance4, dtype=np.uint32)
        self.L.oracle_tfe_reset(p(self.boards), p(c), p(self.scores), p(self.done), C.c_int(self.n_envs), C.c_int(self.side))
        return self.boards

    def step(self, actions, chance2, auto_reset=False, reset_chance4=None):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        c = np.ascontiguousarray(chance2, dtype=np.uint32)
        rc = np.ascontiguousarray(reset_chance4, dtype=np.uint32) if reset_chance4 is not None else None
        self.L.oracle_tfe_step(p(self.boards), p(a), p(c), p(self.scores), p(self.rewards), p(self.done),
                               C.c_int(self.n_envs), C.c_int(self.side), C.c_int(bool(auto_reset)), p(rc))
        return self.boards, self.rewards, self.done.astype(bool)


# ---- Tetris --------------------------------------------------------------------------------------
class OracleTetris:
    """The Tetris spec (oracle_tetris.c): rows uint16 [N,20], bit x of rows[y] = cell (x, y), row 0 bottom."""

    def __init__(self, n_envs):
        self.L = lib()
        self.n_envs = int(n_envs)
        self.rows = np.zeros((self.n_envs, 20), np.uint16)
        self.piece = np.zeros(self.n_envs, np.uint8)
        self.lines = np.zeros(self.n_envs, np.int32)
        self.done = np.zeros(self.n_envs, np.uint8)
        self.rewards = np.zeros(self.n_envs, np.int32)

    def tape_pieces(self, seed, env_base, counter, word):
        out = np.zeros(self.n_envs, np.uint8)
        self.L.oracle_tetris_tape_pieces(C.c_uint64(seed), C.c_uint64(env_base), C.c_uint32(counter), C.c_int(word),
                                         C.c_int(self.n_envs), p(out))
        return out

    def reset(self, first_piece):
        fp = np.ascontiguousarray(first_piece, dtype=np.uint8)
        self.L.oracle_tetris_reset(p(self.rows), p(self.piece), p(self.lines), p(self.done), p(fp), C.c_int(self.n_envs))
        return self.rows

    def step(self, actions, next_piece, auto_reset=False, reset_piece=None):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        nx = np.ascontiguousarray(next_piece, dtype=np.uint8)
        rp = np.ascontiguousarray(reset_piece, dtype=np.uint8) if reset_piece is not None else nx
        self.L.oracle_tetris_step(p(self.rows), p(self.piece), p(self.lines), p(self.done), p(a), p(nx), p(rp),
                                  p(self.rewards), C.c_int(bool(auto_reset)), C.c_int(self.n_envs))
        return self.rows, self.rewards, self.done.astype(bool)
