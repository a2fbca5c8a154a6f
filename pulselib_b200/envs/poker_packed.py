"""PokerGPU on the packed HBM layout: same rules, same reset/step API and return tuples as
`pulselib_b200.envs.poker.PokerGPU` (the drop-in for /root/reference/environments/Poker/PokerGPU.py),
but the per-table state is kept in the 52 B (heads-up) packed structure-of-arrays form described in
csrc/poker_packed_body.cuh instead of one int32 tensor per field.  This is the throughput engine:
training loops that only consume (obs, reward, done, info) can use it unchanged -- info["stacks"] is a
live [N, P] view of the seat-major stack array and info["seat_idx"] a live int32 [N] tensor the step
kernel keeps current, so `info["seat_idx"] == q_seat` and build_actions
(/root/reference/environments/Poker/utils.py:108-123) work as they do on the reference.  Code that
reaches into env.status / env.board etc. gets read-only snapshots through `unpack()` / attribute access;
the registered drop-in `Pulse-Poker-GPU-v1` (envs/poker.py::PokerGPU) wraps this engine and adds
writable mirrors of every reference attribute.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib
from .._lib import c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces
from .poker import _Config, _State, _STATE_FIELDS, device_hand_ranks


class _Packed(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("tbl", "tbl2", "stack", "crb", "inv", "hole", "ev", "hand_ctr", "hand_ranks")] + \
               [("n_hand_ranks", C.c_int64), ("idx", C.c_void_p), ("prev_invested", C.c_void_p)]


_SNAPSHOT = ("stacks", "status", "idx", "agg", "acted", "highest", "last_raise_size", "current_round_bet",
             "total_invested", "pots", "stages", "board", "hands", "deck_positions", "button", "is_done")


class PokerGPUPacked(gym.Env):
    metadata = {"render.modes": ["human"]}
    NUM_ACTIONS = 13
    ACTIVE, FOLDED, ALLIN, SITOUT = 0, 1, 2, 3

    def __init__(self, device, agents=None, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000,
                 w1=.5, w2=.5, K=20, alpha=300, seed=0, env_base=0, auto_reset=False, live_info=True,
                 track_prev_invested=False):
        super().__init__()
        self.device = dev = _lib.require_cuda(device)
        self._lib = _lib.load()
        if not 2 <= int(n_players) <= 10:
            raise ValueError("n_players must be in 2..10")
        if int(max_players) < int(n_players) or int(max_players) > 16:
            raise ValueError("max_players must be in n_players..16")
        self.agents = agents
        self.n_players, self.n_games, self.max_players = int(n_players), int(n_games), int(max_players)
        self.starting_bbs, self.max_bbs = int(starting_bbs), int(max_bbs)
        self.w1, self.w2, self.K, self.alpha = float(w1), float(w2), int(K), int(alpha)
        self.auto_reset = bool(auto_reset)
        self.seed_value, self.env_base, self._hand_counter = int(seed), int(env_base), 0
        self.action_space = spaces.Discrete(self.NUM_ACTIONS)
        self.obs_size = 13 + (self.max_players - 1) * 3
        self.observation_space = spaces.Box(low=0, high=10000, shape=(self.obs_size,), dtype=np.float32)
        self.active_players = self.n_players
        N, P = self.n_games, self.n_players
        self.hand_ranks = device_hand_ranks(dev)
        self._tbl = torch.zeros((N, 4), dtype=torch.int32, device=dev)
        self._tbl2 = torch.zeros((N, 2), dtype=torch.int32, device=dev)
        self._stack = torch.zeros((P, N), dtype=torch.int32, device=dev)
        self._crb = torch.zeros((P, N), dtype=torch.int32, device=dev)
        self._inv = torch.zeros((P, N), dtype=torch.int32, device=dev)
        self._hole = torch.zeros((P, N), dtype=torch.int16, device=dev)
        self._hand_ctr = torch.zeros(N, dtype=torch.int32, device=dev)
        self._ev = torch.zeros((P, N), dtype=torch.int64, device=dev)
        self.obs = torch.zeros((N, self.obs_size), dtype=torch.float32, device=dev)
        self.is_truncated = torch.zeros(N, dtype=torch.bool, device=dev)
        self._first = self._restart_tape = True
        # live mirrors the step kernel keeps current: the seat to act (info["seat_idx"], 4 B per table-step) and,
        # for the drop-in class only, the reference's prev_invested scratch (PokerGPU.py:539)
        self._idx = torch.zeros(N, dtype=torch.int32, device=dev) if live_info else None
        self._prev_inv = torch.zeros(N, dtype=torch.int32, device=dev) if track_prev_invested else None
        self._prefixed = None
        self._stepped = False
        self._packed = _Packed(self._tbl.data_ptr(), self._tbl2.data_ptr(), self._stack.data_ptr(),
                               self._crb.data_ptr(), self._inv.data_ptr(), self._hole.data_ptr(), self._ev.data_ptr(),
                               self._hand_ctr.data_ptr(), self.hand_ranks.data_ptr(), self.hand_ranks.numel(),
                               self._idx.data_ptr() if live_info else None,
                               self._prev_inv.data_ptr() if track_prev_invested else None)

    def set_agents(self, agents):
        self.agents = agents

    def _config(self):
        return _Config(self.n_games, self.n_players, int(self.active_players), self.max_players, self.starting_bbs,
                       self.max_bbs, self.K, self.alpha, self.w1, self.w2)

    def state_bytes_per_table(self) -> int:
        return 24 + 22 * self.n_players  # 14 B/seat touched every step + 8 B/seat of per-deal hand values

    def reset(self, seed=None, options=None, rotation=0, first_flags=None):
        """first_flags (internal, used by the drop-in class): overrides bits 0-1 of pulse_poker_packed_reset's
        first_reset (bit 0 fresh stacks, bit 1 button 0)."""
        if seed is not None:
            self.seed_value, self._hand_counter, self._restart_tape = int(seed), 0, True
        options = options or {}
        N, P, dev = self.n_games, self.n_players, self.device
        if options.get("active_players", False):
            candidate = torch.randint(2, P + 1, (1,), device=dev).item()
        else:
            candidate = P
        self.active_players = int(min(max(candidate, options.get("q_agent_seat", 0) + 1), P))
        prefixed = options.get("prefixed_decks")
        if prefixed is not None:
            prefixed = torch.as_tensor(prefixed, dtype=torch.int32, device=dev)
            if tuple(prefixed.shape) != (N, 52):
                raise ValueError(f"prefixed_decks must have shape {(N, 52)}, got {tuple(prefixed.shape)}")
            prefixed = prefixed.contiguous()
        flags = (3 if self._first else 0) if first_flags is None else int(first_flags)
        rot = 0 if flags & 1 else int(options.get("rotation", rotation))
        cfg = self._config()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_poker_packed_reset(
                C.byref(self._packed), C.byref(cfg), ptr(prefixed), c_u64(self.seed_value), c_u64(self.env_base),
                c_u32(self._hand_counter), c_i(rot), c_i(flags | (4 if self._restart_tape else 0)),
                ptr(self.obs), _lib.cuda_stream(dev)), "pulse_poker_packed_reset")
        self._first = self._restart_tape = False
        self._hand_counter += 1
        self._prefixed = prefixed   # the mirror's decks come from here (or from the tape)
        self._stepped = False
        return self.obs, self.get_info()

    def step_into(self, actions, obs, rewards, done):
        """step() writing into caller-provided tensors (no allocation; used by rollout loops)."""
        cfg = self._config()
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_poker_packed_step(
                C.byref(self._packed), C.byref(cfg), ptr(actions), ptr(obs), ptr(rewards), ptr(done),
                c_i(1 if self.auto_reset else 0), c_u64(self.seed_value), c_u64(self.env_base),
                _lib.cuda_stream(self.device)), "pulse_poker_packed_step")
        self._stepped = True

    def step(self, actions):
        a = torch.as_tensor(actions, device=self.device)
        if a.dtype != torch.int64 or not a.is_contiguous():
            a = a.to(torch.int64).contiguous()
        if tuple(a.shape) != (self.n_games,):
            raise ValueError(f"actions must have shape {(self.n_games,)}, got {tuple(a.shape)}")
        rewards = torch.empty(self.n_games, dtype=torch.float32, device=self.device)
        done = torch.empty(self.n_games, dtype=torch.bool, device=self.device)
        self.step_into(a, self.obs, rewards, done)
        self._done = done
        return self.obs, rewards, done, self.is_truncated, self.get_info()

    def step_policy(self, seat_kinds, actions=None, chance2=None, counter=None, return_actions=True, obs_out=None,
                    return_seat=False):
        """step() with the scripted seats' rules evaluated inside the step kernel: one
        launch does build_actions + step.  seat_kinds: one PULSE_AGENT_* per seat (pulselib_b200.policy);
        `actions` int64 [N] is read only for EXTERNAL seats.  Returns step()'s tuple with the played
        actions added to info["actions"] (and, with return_seat, the seat that acted in info["seat_idx_before"]).
        obs_out: write the observation there instead of self.obs (lets a loop ping-pong state / next_state
        without the reference's state.clone())."""
        kb = bytes(seat_kinds)
        if len(kb) != self.n_players:
            raise ValueError(f"seat_kinds must have {self.n_players} entries")
        N, dev = self.n_games, self.device
        if actions is not None:
            actions = torch.as_tensor(actions, device=dev)
            if actions.dtype != torch.int64 or not actions.is_contiguous():
                actions = actions.to(torch.int64).contiguous()
            if tuple(actions.shape) != (N,):
                raise ValueError(f"actions must have shape {(N,)}, got {tuple(actions.shape)}")
        elif any(k == 0 for k in kb):
            raise ValueError("actions are required when a seat is EXTERNAL")
        if chance2 is not None:
            chance2 = torch.as_tensor(chance2, device=dev)
            if tuple(chance2.shape) != (N, 2):
                raise ValueError(f"chance2 must have shape {(N, 2)}")
            chance2 = chance2.to(torch.int64).to(torch.int32).contiguous()
        if counter is None:
            counter = self._policy_counter = getattr(self, "_policy_counter", -1) + 1
        rewards = torch.empty(N, dtype=torch.float32, device=dev)
        done = torch.empty(N, dtype=torch.bool, device=dev)
        chosen = torch.empty(N, dtype=torch.int64, device=dev) if return_actions else None
        seat = torch.empty(N, dtype=torch.int32, device=dev) if return_seat else None
        obs = self.obs if obs_out is None else obs_out
        cfg = self._config()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_poker_packed_step_policy(
                C.byref(self._packed), C.byref(cfg), ptr(actions), C.c_char_p(kb), ptr(chance2), c_u32(counter),
                ptr(chosen), ptr(seat), ptr(obs), ptr(rewards), ptr(done), c_i(1 if self.auto_reset else 0),
                c_u64(self.seed_value), c_u64(self.env_base), _lib.cuda_stream(dev)), "pulse_poker_packed_step_policy")
        self._done = done
        self._stepped = True
        info = self.get_info()
        info["actions"] = chosen
        if return_seat:
            info["seat_idx_before"] = seat
        return obs, rewards, done, self.is_truncated, info

    def rollout_policy(self, seat_kinds, n_steps, counter=None, return_actions=True):
        """K steps in ONE launch with every seat played by a device rule (pulse_poker_packed_rollout): returns
        (final obs [N, W], rewards float32 [K, N], dones bool [K, N], actions int64 [K, N] or None) -- what K
        step_policy() calls with counters counter .. counter + K - 1 would return one after another."""
        kb = bytes(seat_kinds)
        if len(kb) != self.n_players:
            raise ValueError(f"seat_kinds must have {self.n_players} entries")
        N, K, dev = self.n_games, int(n_steps), self.device
        if counter is None:
            counter = getattr(self, "_policy_counter", -1) + 1
        self._policy_counter = counter + K - 1
        rewards = torch.empty((K, N), dtype=torch.float32, device=dev)
        done = torch.empty((K, N), dtype=torch.bool, device=dev)
        chosen = torch.empty((K, N), dtype=torch.int64, device=dev) if return_actions else None
        cfg = self._config()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_poker_packed_rollout(
                C.byref(self._packed), C.byref(cfg), C.c_char_p(kb), c_u32(counter), c_i(K), c_i(1 if self.auto_reset else 0),
                c_u64(self.seed_value), c_u64(self.env_base), ptr(self.obs), ptr(rewards), ptr(done), ptr(chosen),
                _lib.cuda_stream(dev)), "pulse_poker_packed_rollout")
        self._stepped = True
        return self.obs, rewards, done, chosen

    def rollout_actions(self, actions):
        """K steps in ONE launch on supplied actions (pulse_poker_packed_rollout_actions): actions uint8 [K, N] ->
        (obs after the last step, rewards float32 [K, N], dones bool [K, N]) -- what K step() calls would return one
        after another (no auto_reset: finished tables idle, as in the reference)."""
        a = torch.as_tensor(actions, device=self.device)
        if a.dim() != 2 or a.shape[1] != self.n_games:
            raise ValueError(f"actions must have shape (K, {self.n_games}), got {tuple(a.shape)}")
        if a.dtype != torch.uint8 or not a.is_contiguous():
            a = torch.where(a < 0, torch.full_like(a, 255), a.clamp(max=13)).to(torch.uint8).contiguous()
        K, N, dev = int(a.shape[0]), self.n_games, self.device
        rewards = torch.empty((K, N), dtype=torch.float32, device=dev)
        done = torch.empty((K, N), dtype=torch.bool, device=dev)
        cfg = self._config()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_poker_packed_rollout_actions(
                C.byref(self._packed), C.byref(cfg), ptr(a), c_i(K), ptr(self.obs), ptr(rewards), ptr(done),
                _lib.cuda_stream(dev)), "pulse_poker_packed_rollout_actions")
        self._stepped = True
        return self.obs, rewards, done

    def enumerated_strength(self):
        """float32 [N, active_players]: river hand strength averaged over the board cards still to come, per seat
        (see include/pulse_envs.h).  An optional reward input; step() keeps the reference's proxy."""
        out = torch.empty((self.n_games, int(self.active_players)), dtype=torch.float32, device=self.device)
        cfg = self._config()
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_poker_packed_enumerated_strength(C.byref(self._packed), C.byref(cfg), ptr(out),
                                                                       _lib.cuda_stream(self.device)),
                       "pulse_poker_packed_enumerated_strength")
        return out

    def showdown_equity(self, preflop_samples=256, counter=None):
        """float32 [n_games, active_players]: every seat's all-in equity with ALL hole cards known (SURVEY 8(f) N4;
        /root/reference/rl/depth_notes.txt:316-321 asks for it).  Flop, turn and river are enumerated exactly; preflop
        averages `preflop_samples` boards drawn from the chance tape at `counter` (default: a counter of this env
        that advances with every call, so repeated calls average independent samples)."""
        if counter is None:
            counter = self._equity_counter = getattr(self, "_equity_counter", -1) + 1
        out = torch.empty((self.n_games, self.active_players), dtype=torch.float32, device=self.device)
        cfg = self._config()
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_poker_packed_showdown_equity(
                C.byref(self._packed), C.byref(cfg), ptr(out), c_i(int(preflop_samples)), c_u64(self.seed_value),
                c_u64(self.env_base), c_u32(int(counter) & 0xFFFFFFFF), _lib.cuda_stream(self.device)),
                "pulse_poker_packed_showdown_equity")
        return out

    def get_info(self):
        # what build_actions and the trainers read (environments/Poker/utils.py:108-123, scripts/Poker/trainGPU.py:75-96):
        # real tensors -- `info["seat_idx"] == q_seat` is a bool tensor, `info["stacks"][:, q_seat]` a row of the
        # seat-major stack array
        seat = self._idx if self._idx is not None else self.unpack(("idx",))["idx"]
        return {"active_players": self.active_players, "stacks": self._stack.t(), "seat_idx": seat}

    def alloc_mirror(self, fields=None, full=False) -> dict:
        """Fresh reference-shaped tensors (shapes / dtypes of SURVEY.md 8(a)) for `fields`."""
        N, P, dev = self.n_games, self.n_players, self.device
        AP = int(self.active_players)
        i32 = dict(dtype=torch.int32, device=dev)
        shapes = {n: (N,) for n in ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg",
                                    "acted", "last_raise_size", "prev_invested")}
        shapes.update({n: (N, P) for n in ("stacks", "current_round_bet", "total_invested", "status")})
        shapes.update(board=(N, 5), hands=(N, P, 2))
        if full:
            shapes["decks"] = (N, 52)
        if fields is not None and "decks" in fields:
            shapes["decks"] = (N, 52)
        names = fields if fields is not None else tuple(shapes) + ("is_done", "equity_dirty", "equities")
        out = {}
        for n in names:
            if n in ("is_done", "equity_dirty"):
                out[n] = torch.empty(N, dtype=torch.bool, device=dev)
            elif n == "equities":
                out[n] = torch.empty((N, AP), dtype=torch.float32, device=dev)
            else:
                out[n] = torch.empty(shapes[n], **i32)
        return out

    def unpack_into(self, out: dict, full=False):
        """Fills the given reference-shaped tensors from the packed state (one launch).  full: also decks (re-drawn
        from the tape or copied from the last reset's prefixed decks) -- pulse_poker_unpack_full."""
        s = _State()
        for name in _STATE_FIELDS:
            t = self.hand_ranks if name == "hand_ranks" else out.get(name)
            setattr(s, name, None if t is None else t.data_ptr())
        s.n_hand_ranks = self.hand_ranks.numel()
        cfg, dev = self._config(), self.device
        with _lib.device_guard(dev):
            if full:
                _lib.check(self._lib.pulse_poker_unpack_full(
                    C.byref(self._packed), C.byref(s), C.byref(cfg), ptr(self._prefixed), c_u64(self.seed_value),
                    c_u64(self.env_base), c_i(0 if self._stepped else 1), _lib.cuda_stream(dev)), "pulse_poker_unpack_full")
            else:
                _lib.check(self._lib.pulse_poker_unpack(C.byref(self._packed), C.byref(s), C.byref(cfg),
                                                        _lib.cuda_stream(dev)), "pulse_poker_unpack")
        return out

    def pack_from(self, state) -> int:
        """Reference-shaped state -> the packed layout (pulse_poker_pack), e.g. after edits made through the reference's
        attributes.  `state`: a ctypes _State (envs/poker.py) of device tensors of the reference's shapes / dtypes.
        Returns the number of tables the packed layout cannot express (0 = all good); reads that count back, so it
        synchronises the stream."""
        bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        cfg = self._config()
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_poker_pack(C.byref(state), C.byref(self._packed), C.byref(cfg), ptr(bad),
                                                  _lib.cuda_stream(self.device)), "pulse_poker_pack")
        self._stepped = True
        return int(bad.item())

    def unpack(self, fields=None) -> dict:
        """Reference-shaped int32 snapshot of the packed state (one kernel launch)."""
        return self.unpack_into(self.alloc_mirror(fields if fields is not None else _SNAPSHOT))

    def __getattr__(self, name):
        if name in _SNAPSHOT:
            return self.unpack()[name]
        raise AttributeError(name)

    def close(self):
        return None
