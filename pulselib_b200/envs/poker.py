"""PokerGPU behind the reference's gym.Env surface, stepped by one fused CUDA launch.

Drop-in for /root/reference/environments/Poker/PokerGPU.py:8-633: same constructor, same tensor
attributes (shapes/dtypes of SURVEY.md 8(a)), same internal method names, same return tuples.

Two classes:

* `PokerGPUCompat` keeps every reference attribute as its own int32 torch tensor and hands them by pointer to
  libpulse_envs.so (include/pulse_envs.h: pulse_poker_reset / pulse_poker_step).  Tests of the reference overwrite
  these tensors and rebind them; the pointer table is therefore rebuilt from the live attributes on every call.
* `PokerGPU` -- what `gym.make("Pulse-Poker-GPU-v1")` builds -- runs on the packed throughput layout
  (envs/poker_packed.py, csrc/poker_packed*.cu*) and **mirrors on access** (SURVEY.md 7.3 item 4): reset / step /
  info need no reference-shaped tensor at all; the first read of `env.stages`, `env.board`, ... between two calls
  unpacks the whole state into fresh reference-shaped tensors (one launch, pulse_poker_unpack_full); if the caller
  then writes into one of them, rebinds one, or calls one of the reference's internal methods
  (`execute_actions`, `calculate_equities`, ...), the env carries on from those tensors on the reference-shaped
  kernels until the next reset().  Writes are detected through torch's per-tensor version counter, so reads cost
  one unpack and nothing else.
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


class PokerGPUCompat(gym.Env):
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


_MIRROR = ("decks", "deck_positions", "board", "pots", "stages", "stacks", "hands", "current_round_bet",
           "total_invested", "status", "button", "sb", "bb", "idx", "highest", "agg", "acted", "last_raise_size",
           "prev_invested", "is_done", "equity_dirty", "equities")
_LIVE = ("stacks", "idx", "prev_invested", "is_done")       # real tensors the packed kernels keep current
_DERIVED = ("button_pos", "prev_stacks", "raise_amounts", "active_player_idx")
_MIRROR_SET = frozenset(_MIRROR)
_LAZY_SET = frozenset(_MIRROR + _DERIVED) - frozenset(_LIVE)


class PokerGPU(PokerGPUCompat):
    """`Pulse-Poker-GPU-v1`: the reference's env surface on the packed throughput layout, mirror on access.

    Packed mode (default after every reset): `stacks` is a live [N, P] view of the engine's seat-major stack array,
    `idx` / `prev_invested` / `is_done` are the tensors the step kernel writes, `obs` the observation buffer; every
    other reference attribute appears when it is first read (a snapshot that is dropped by the next reset/step).
    Reference-shaped mode: entered when a mirrored tensor was written or rebound, an internal method is called, or
    `calculate_equities` was replaced on the instance; also for good when `hand_ranks` is given (a caller-supplied
    table defines the hand values, the packed engine computes them) or `layout="compat"`."""

    def __init__(self, device, agents, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000,
                 w1=.5, w2=.5, K=20, alpha=300, seed=0, env_base=0, hand_ranks=None, layout="packed"):
        d = self.__dict__
        d["_mode"], d["_eng"], d["_eng_live"] = "compat", None, False
        d["_ver"], d["_materialised"], d["_touched"] = {}, None, False
        super().__init__(device, agents, n_players, max_players, n_games, starting_bbs, max_bbs, w1, w2, K, alpha,
                         seed, env_base, hand_ranks)
        if layout not in ("packed", "compat"):
            raise ValueError("layout must be 'packed' or 'compat'")
        d["_force_compat"] = hand_ranks is not None or layout == "compat"

    # ---- attribute plumbing ----------------------------------------------------------------------------------
    def __getattr__(self, name):               # only reached when normal lookup fails
        d = self.__dict__
        if name in _LAZY_SET and d.get("_mode") == "packed" and d.get("_eng_live"):
            self._materialise(decks=name == "decks")
            if name in d:
                return d[name]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        d = self.__dict__
        if name in _MIRROR_SET and d.get("_mode") == "packed" and d.get("_eng_live"):
            self._materialise(decks=True)      # the rest of the state must exist when this one is overridden
            d["_touched"] = True
        super().__setattr__(name, value)

    def __delattr__(self, name):
        d = self.__dict__
        if name in _LAZY_SET and d.get("_mode") == "packed" and d.get("_eng_live"):
            self._materialise(decks=True)      # e.g. `del env.button_pos` in a white-box test
            d["_touched"] = True
        super().__delattr__(name)

    def packed_engine(self):
        """The PokerGPUPacked engine while it is authoritative (helpers such as HandStats read it directly)."""
        if self._mode == "packed" and self._eng_live:
            self._sync()
        return self._eng if self._mode == "packed" and self._eng_live else None

    def _record_live(self):
        d, ver = self.__dict__, self.__dict__["_ver"]
        for n in _LIVE:
            t = d[n]
            ver[n] = (t, t._version)

    def _materialise(self, decks=False):
        """Snapshot of the packed state as reference-shaped tensors, one launch.  The decks (52 tape draws per table to
        re-derive, or a copy of the prefixed ones) only when somebody asks for them."""
        d = self.__dict__
        eng = d["_eng"]
        if d["_materialised"] is None:
            names = tuple(n for n in _MIRROR if n not in _LIVE and n != "decks")
            out = eng.unpack_into(eng.alloc_mirror(names), full=True)   # no decks tensor given: none written
            N, i32 = self.n_games, dict(dtype=torch.int32, device=self.device)
            out["button_pos"] = out["button"][0]
            out["prev_stacks"], out["raise_amounts"] = torch.zeros(N, **i32), torch.zeros(N, **i32)
            out["active_player_idx"] = torch.arange(int(self.active_players), **i32)
            d["_materialised"] = names + _DERIVED
        elif decks and "decks" not in d:
            out = {}
        else:
            return
        if decks and "decks" not in d:
            out.update(eng.unpack_into(eng.alloc_mirror(("decks",), full=True), full=True))
            d["_materialised"] += ("decks",)
        for n, t in out.items():
            d[n] = t
            if n in _MIRROR_SET:
                d["_ver"][n] = (t, t._version)

    def _sync(self):
        """Start of every call in packed mode: did the caller write to / rebind any exposed tensor?  Yes -> carry on
        from the reference-shaped tensors.  No -> drop the snapshot (the next packed launch makes it stale)."""
        d = self.__dict__
        changed = d["_touched"]
        if not changed:
            for n, (t, v) in d["_ver"].items():
                if d.get(n) is not t or t._version != v:
                    changed = True
                    break
        if changed:
            self._go_compat()
            return
        if d["_materialised"] is not None:
            for n in d["_materialised"]:
                d.pop(n, None)
                d["_ver"].pop(n, None)
            d["_materialised"] = None

    def _go_compat(self):
        d = self.__dict__
        if d["_mode"] != "packed" or not d["_eng_live"]:
            return
        self._materialise(decks=True)
        d["_mode"], d["_touched"], d["_materialised"] = "compat", False, None
        d["_ver"] = {}
        d["stacks"] = d["stacks"].contiguous()   # [N, P] row-major copy of the seat-major view
        if d["prev_invested"] is None:
            d["prev_invested"] = torch.zeros(self.n_games, dtype=torch.int32, device=self.device)

    def repack(self) -> bool:
        """Back onto the packed throughput layout after the reference-shaped tensors were edited (white-box tests,
        hand-written positions): pulse_poker_pack.  Returns False -- and carries on reference-shaped -- when some table
        is not expressible in the packed layout (e.g. a deck position the street does not imply), when
        `calculate_equities` was replaced on the instance, or when the env was built reference-shaped for good."""
        d = self.__dict__
        if d["_mode"] == "packed":
            return bool(d["_eng_live"])
        if d["_force_compat"] or d["_eng"] is None or "calculate_equities" in d or d.get("stacks") is None:
            return False
        eng, N = d["_eng"], self.n_games
        eng.active_players = int(self.active_players)
        for n in ("w1", "w2", "K", "alpha"):
            setattr(eng, n, d["_host_" + n])
        if eng.pack_from(self._state()) != 0:
            return False
        eng._stack.copy_(self._tensor("stacks", torch.int32).t())       # seats beyond active_players keep their chips
        eng._prefixed = self._tensor("decks", torch.int32)              # the mirror hands these decks back
        eng.obs = d["obs"]
        done = self._tensor("is_done", torch.bool)
        for n in _MIRROR + _DERIVED:
            d.pop(n, None)
        d["_ver"], d["_materialised"], d["_touched"] = {}, None, False
        d["stacks"], d["idx"], d["prev_invested"] = eng._stack.t(), eng._idx, eng._prev_inv
        d["is_done"] = done
        d["_mode"], d["_eng_live"] = "packed", True
        self._record_live()
        return True

    # ---- reset / step -------------------------------------------------------------------------------------------
    def reset(self, seed=None, options=None, rotation=0):
        d = self.__dict__
        if d["_force_compat"]:
            return super().reset(seed=seed, options=options, rotation=rotation)
        from .poker_packed import PokerGPUPacked

        N, dev = self.n_games, self.device
        if d["_eng"] is None:
            d["_eng"] = PokerGPUPacked(dev, agents=None, n_players=self.n_players, max_players=self.max_players,
                                       n_games=N, starting_bbs=self.starting_bbs, max_bbs=self.max_bbs, seed=self.seed_value,
                                       env_base=self.env_base, live_info=True, track_prev_invested=True)
        eng = d["_eng"]
        if d["_mode"] == "packed" and d["_eng_live"]:
            self._sync()
        flags = 0
        if d["_mode"] == "compat":
            # what persists across resets in the reference: the stacks (:101-110) and the button (:121)
            stacks = d.get("stacks")
            if stacks is None:
                flags |= 1
            else:
                eng._stack.copy_(self._tensor("stacks", torch.int32).t())
            if "button_pos" in d:
                eng._tbl[:, 0] = (self._tensor("button", torch.int32) & 15) << 11
            else:
                flags |= 2
        for n in ("w1", "w2", "K", "alpha"):
            setattr(eng, n, d["_host_" + n])
        eng._hand_counter = self._hand_counter
        eng.obs = torch.empty((N, self.obs_size), dtype=torch.float32, device=dev)   # fresh every reset, as the reference
        if seed is not None:
            d["seed_value"] = int(seed)
        eng.reset(seed=seed, options=options, rotation=rotation, first_flags=flags)
        d["_hand_counter"] = eng._hand_counter
        for n in _MIRROR + _DERIVED:
            d.pop(n, None)
        d["_ver"], d["_materialised"], d["_touched"] = {}, None, False
        d["active_players"] = eng.active_players
        d["stacks"], d["idx"], d["prev_invested"] = eng._stack.t(), eng._idx, eng._prev_inv
        d["is_done"] = torch.zeros(N, dtype=torch.bool, device=dev)
        d["obs"] = eng.obs
        d["_mode"], d["_eng_live"] = "packed", True
        self._record_live()
        return self.obs, self.get_info()

    def step(self, actions):
        d = self.__dict__
        if d["_mode"] == "packed" and d["_eng_live"]:
            if "calculate_equities" in d:      # replaced per instance: it must be what step() calls
                self._go_compat()
            else:
                self._sync()
        if d["_mode"] == "compat":
            return super().step(actions)
        a = self._actions(actions)
        eng, N, dev = d["_eng"], self.n_games, self.device
        rewards = torch.empty(N, dtype=torch.float32, device=dev)
        done = torch.empty(N, dtype=torch.bool, device=dev)   # the reference rebinds is_done every step (:623)
        for n in ("w1", "w2", "K", "alpha"):
            setattr(eng, n, d["_host_" + n])
        eng.step_into(a, d["obs"], rewards, done)
        d["is_done"] = done
        d["_ver"]["is_done"] = (done, done._version)
        return d["obs"], rewards, done, self.is_truncated, self.get_info()

    # ---- the reference's internal methods: white-box callers work on the reference-shaped tensors ----------------
    def get_obs(self):
        if self._mode == "packed" and self._eng_live:
            self._sync()
            if self._mode == "packed":
                return self.obs            # current by construction: nothing was written since the last launch
        return super().get_obs()

    def calculate_equities(self):
        self._go_compat()
        return super().calculate_equities()

    def execute_actions(self, actions):
        self._go_compat()
        return super().execute_actions(actions)

    def resolve_fold_winners(self):
        self._go_compat()
        return super().resolve_fold_winners()

    def resolve_terminated_games(self):
        self._go_compat()
        return super().resolve_terminated_games()

    def poker_reward_gpu(self, actions, actor_idx):
        self._go_compat()
        return super().poker_reward_gpu(actions, actor_idx)
