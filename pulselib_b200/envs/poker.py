"""PokerGPU behind the reference's gym.Env surface, stepped by one fused CUDA launch.

Drop-in for /root/reference/environments/Poker/PokerGPU.py:8-633: same constructor, same tensor
attributes (shapes/dtypes of SURVEY.md 8(a)), same internal method names, same return tuples.  All
state lives in torch CUDA tensors that are handed by pointer to libpulse_envs.so
(include/pulse_envs.h: pulse_poker_reset / pulse_poker_step).  Tests of the reference overwrite
these tensors and rebind them; the pointer table is therefore rebuilt from the live attributes on
every call.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib, handranks
from .._lib import c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces

PH_EQUITIES, PH_EXECUTE, PH_ADVANCE, PH_FOLD, PH_SHOWDOWN, PH_CLEAR, PH_REWARD, PH_OBS = 1, 2, 4, 8, 16, 32, 64, 128
PH_STEP, PH_NEWLY_IS_DONE, PH_RAW_REWARD = 255, 256, 512

_STATE_FIELDS = ("decks", "deck_positions", "board", "pots", "stages", "stacks", "hands", "current_round_bet",
                 "total_invested", "status", "button", "sb", "bb", "idx", "highest", "agg", "acted",
                 "last_raise_size", "prev_invested", "is_done", "equity_dirty", "equities", "hand_ranks")


class _State(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _STATE_FIELDS] + [("n_hand_ranks", C.c_int64)]


class _Config(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_games", "n_players", "active_players", "max_players", "starting_bbs",
                                         "max_bbs", "K", "alpha")] + [("w1", C.c_float), ("w2", C.c_float)]


_TABLES: dict = {}


def device_hand_ranks(device: torch.device) -> torch.Tensor:
    """The 130 MB table, uploaded once per device and shared by every env on it."""
    key = (device.type, device.index)
    if key not in _TABLES:
        _TABLES[key] = torch.from_numpy(handranks.load_table()).to(device)
    return _TABLES[key]


class PokerGPU(gym.Env):
    metadata = {"render.modes": ["human"]}
    NUM_ACTIONS = 13
    ACTIVE, FOLDED, ALLIN, SITOUT = 0, 1, 2, 3
    STATE_SPACE = 28
    MIN_EQUITY_RANK, MAX_EQUITY_RANK = 4145.0, 36874.0
    MAX_FLOP_EQUITY, MIN_FLOP_EQUITY = 823779.0, 74359.0
    MIN_TURN_RIVER_EQUITY, MAX_TURN_RIVER_EQUITY = 4109.0, 36874.0

    def __init__(self, device, agents, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000,
                 w1=.5, w2=.5, K=20, alpha=300, seed=0, env_base=0, hand_ranks=None):
        super().__init__()
        self.device = _lib.require_cuda(device)
        self._lib = _lib.load()
        if not 2 <= int(n_players) <= 10:
            raise ValueError("n_players must be in 2..10")
        if int(max_players) < int(n_players) or int(max_players) > 16:
            raise ValueError("max_players must be in n_players..16")
        dev = self.device
        self.w1 = torch.tensor(w1, device=dev, dtype=torch.float32)
        self.w2 = torch.tensor(w2, device=dev, dtype=torch.float32)
        self.K = torch.tensor(K, device=dev, dtype=torch.int32)
        self.alpha = torch.tensor(alpha, device=dev, dtype=torch.int32)
        self.agents = agents
        self.n_players, self.n_games = int(n_players), int(n_games)
        self.starting_bbs, self.max_bbs, self.max_players = int(starting_bbs), int(max_bbs), int(max_players)
        self.action_space = spaces.Discrete(self.NUM_ACTIONS)
        self.raise_fractions = torch.tensor([0.25, 0.33, 0.50, 0.75, 1.00, 1.50, 2.00, 3.00, 4.00], device=dev)
        self.obs_size = 13 + ((self.max_players - 1) * 3)
        self.observation_space = spaces.Box(low=0, high=10000, shape=(self.obs_size,), dtype=np.float32)
        self.stacks = None
        # the reference raises FileNotFoundError when HandRanks.dat is absent (PokerGPU.py:47-49);
        # here the table is regenerated instead (pulselib_b200/handranks.py) unless one is given
        self.hand_ranks = device_hand_ranks(dev) if hand_ranks is None else \
            torch.as_tensor(hand_ranks, dtype=torch.int32, device=dev).contiguous()
        self.last_raise_size = None
        self.g = torch.arange(self.n_games, device=dev)
        self.is_truncated = torch.zeros(self.n_games, dtype=torch.bool, device=dev)
        self.bb_amounts = torch.ones(self.n_games, device=dev, dtype=torch.int32)
        self.offset_cards = torch.arange(52, device=dev, dtype=torch.int32)
        self.seed_value, self.env_base, self._hand_counter = int(seed), int(env_base), 0
        self.active_players = self.n_players

    # w1/w2/K/alpha are 0-dim tensors in the reference and tests rebind them; cache host copies at
    # assignment so step() never has to read them back from the device
    def __setattr__(self, name, value):
        if name in ("w1", "w2", "K", "alpha"):
            host = float(value) if name in ("w1", "w2") else int(value)
            object.__setattr__(self, "_host_" + name, host)
        object.__setattr__(self, name, value)

    def set_agents(self, agents):
        self.agents = agents

    # ---- pointer table -------------------------------------------------------------------------
    def _tensor(self, name, dtype):
        t = getattr(self, name)
        if t.dtype != dtype or not t.is_contiguous() or t.device != self.device:
            t = t.to(device=self.device, dtype=dtype).contiguous()
            object.__setattr__(self, name, t)
        return t

    def _state(self):
        s = _State()
        for name in _STATE_FIELDS:
            dtype = torch.bool if name in ("is_done", "equity_dirty") else \
                torch.float32 if name == "equities" else torch.int32
            setattr(s, name, self._tensor(name, dtype).data_ptr())
        s.n_hand_ranks = self.hand_ranks.numel()
        return s

    def _config(self):
        return _Config(self.n_games, self.n_players, int(self.active_players), self.max_players, self.starting_bbs,
                       self.max_bbs, self._host_K, self._host_alpha, self._host_w1, self._host_w2)

    def _launch(self, phases, actions=None, actor_idx=None, obs=None, rewards=None, done_out=None):
        s, cfg = self._state(), self._config()
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_poker_step(
                C.byref(s), C.byref(cfg), ptr(actions), ptr(actor_idx), c_u32(phases), ptr(obs), ptr(rewards),
                ptr(done_out), _lib.cuda_stream(self.device)), "pulse_poker_step")

    def _actions(self, actions):
        a = torch.as_tensor(actions, device=self.device)
        if a.dtype != torch.int64 or not a.is_contiguous():
            a = a.to(torch.int64).contiguous()
        if tuple(a.shape) != (self.n_games,):
            raise ValueError(f"actions must have shape {(self.n_games,)}, got {tuple(a.shape)}")
        return a

    # ---- reset (PokerGPU.py:73-157) --------------------------------------------------------------
    def reset(self, seed=None, options=None, rotation=0):
        if seed is not None:
            self.seed_value, self._hand_counter = int(seed), 0
        options = options or {}
        N, P, dev = self.n_games, self.n_players, self.device
        if options.get("active_players", False):
            candidate = torch.randint(2, P + 1, (1,), device=dev).item()
        else:
            candidate = P
        q_seat = options.get("q_agent_seat", 0)
        self.active_players = int(min(max(candidate, q_seat + 1), P))
        AP = self.active_players

        prefixed = options.get("prefixed_decks")
        if prefixed is not None:
            prefixed = torch.as_tensor(prefixed, dtype=torch.int32, device=dev)
            if tuple(prefixed.shape) != (N, 52):
                raise ValueError(f"prefixed_decks must have shape {(N, 52)}, got {tuple(prefixed.shape)}")
            prefixed = prefixed.contiguous()

        first = self.stacks is None
        i32 = dict(dtype=torch.int32, device=dev)
        if first:
            self.stacks = torch.empty((N, P), **i32)
        # fresh tensors every reset, as in the reference
        self.decks = torch.empty((N, 52), **i32)
        for name in ("deck_positions", "pots", "stages", "sb", "bb", "idx", "highest", "agg", "acted",
                     "last_raise_size", "prev_invested"):
            setattr(self, name, torch.empty(N, **i32))
        self.board = torch.empty((N, 5), **i32)
        self.hands = torch.empty((N, P, 2), **i32)
        self.current_round_bet = torch.empty((N, P), **i32)
        self.total_invested = torch.empty((N, P), **i32)
        self.status = torch.empty((N, P), **i32)
        had_button = hasattr(self, "button_pos")
        if not had_button:
            self.button = torch.zeros(N, **i32)
        else:
            self.button = self.button.clone()
        self.is_done = torch.empty(N, dtype=torch.bool, device=dev)
        self.equity_dirty = torch.empty(N, dtype=torch.bool, device=dev)
        self.equities = torch.empty((N, AP), dtype=torch.float32, device=dev)
        self.obs = torch.empty((N, self.obs_size), dtype=torch.float32, device=dev)
        self.prev_stacks = torch.zeros(N, **i32)
        self.raise_amounts = torch.zeros(N, **i32)
        self.active_player_idx = torch.arange(AP, **i32)

        stack_rotation = 0 if first else int(options.get("rotation", rotation))
        s, cfg = self._state(), self._config()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_poker_reset(
                C.byref(s), C.byref(cfg), ptr(prefixed), c_u64(self.seed_value), c_u64(self.env_base),
                c_u32(self._hand_counter), c_i(stack_rotation), c_i((1 if first else 0) | (0 if had_button else 2)), ptr(self.obs),
                _lib.cuda_stream(dev)), "pulse_poker_reset")
        self._hand_counter += 1
        self.button_pos = self.button[0]
        return self.obs, self.get_info()

    # ---- step (PokerGPU.py:527-633): one fused launch ---------------------------------------------
    def step(self, actions):
        actions = self._actions(actions)
        N, dev = self.n_games, self.device
        rewards = torch.empty(N, dtype=torch.float32, device=dev)
        done = torch.empty(N, dtype=torch.bool, device=dev)  # the reference rebinds is_done (:623)
        phases = PH_STEP
        if "calculate_equities" in self.__dict__:  # replaced per instance (benchmarks/tests do this)
            if self.equity_dirty.any():
                self.calculate_equities()
            phases &= ~PH_EQUITIES
        self._launch(phases, actions=actions, obs=self.obs, rewards=rewards, done_out=done)
        self.is_done = done
        return self.obs, rewards, self.is_done, self.is_truncated, self.get_info()

    # ---- the reference's internal methods, each one phase of the same kernel ----------------------
    def get_obs(self):
        self._launch(PH_OBS, obs=self.obs)
        return self.obs

    def get_info(self):
        return {"active_players": self.active_players, "stacks": self.stacks, "seat_idx": self.idx}

    def calculate_equities(self):
        self._launch(PH_EQUITIES)

    def execute_actions(self, actions):
        self._launch(PH_EXECUTE, actions=self._actions(actions))

    def resolve_fold_winners(self):
        self._launch(PH_FOLD | PH_NEWLY_IS_DONE)

    def resolve_terminated_games(self):
        self._launch(PH_SHOWDOWN | PH_NEWLY_IS_DONE)

    def poker_reward_gpu(self, actions, actor_idx):
        rewards = torch.empty(self.n_games, dtype=torch.float32, device=self.device)
        actor = torch.as_tensor(actor_idx, device=self.device).to(torch.int32).contiguous()
        self._launch(PH_REWARD | PH_RAW_REWARD, actions=self._actions(actions), actor_idx=actor, rewards=rewards)
        return rewards

    # small helpers the reference exposes; plain tensor indexing, not on the step path
    def post_blinds(self):  # PokerGPU.py:188-199
        g, bb = self.g, self.bb.long()
        self.stacks[g, bb] -= self.bb_amounts
        self.current_round_bet[g, bb] += self.bb_amounts
        self.total_invested[g, bb] += self.bb_amounts
        self.pots += self.bb_amounts
        self.status[g, bb] = torch.where(self.stacks[g, bb] == 0, self.ALLIN, self.ACTIVE).to(torch.int32)

    def deal_players_cards(self, n_cards):  # PokerGPU.py:201-206
        pos = self.deck_positions.long().unsqueeze(1) + torch.arange(n_cards, device=self.device).unsqueeze(0)
        cards = torch.gather(self.decks, 1, pos)
        self.deck_positions += n_cards
        return cards

    def deal_cards(self, g, n_cards):  # PokerGPU.py:208-214
        g = torch.as_tensor(g, device=self.device).long()
        pos = self.deck_positions[g].long().unsqueeze(1) + torch.arange(n_cards, device=self.device).unsqueeze(0)
        cards = self.decks[g.unsqueeze(1), pos]
        self.deck_positions[g] += n_cards
        return cards.to(torch.int32)

    def close(self):
        return None
