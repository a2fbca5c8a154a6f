"""BlackJack behind the reference's gym.Env surface
(/root/reference/environments/blackjack/blackjack.py:9-186), one fused CUDA launch per step.

Same constructor, same return tuples (`step` returns `(obs int32 [N,3], rewards int32 [N], terminated
bool [N], None, {})`, with obs / rewards / terminated the same tensor objects every call until the next
reset, blackjack.py:186), same attribute names.  The engine keeps 4 B of packed state and a 52 B deck
per env (csrc/blackjack_body.cuh); the reference's per-card tensors (`players_cards`, `dealer_cards`,
`decks`, sums, flags) are rebuilt on demand by `pulse_blackjack_unpack` when an attribute is read.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib
from .._lib import c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces

_COMPAT = (("decks", torch.int64, 52), ("deck_positions", torch.int32, 0), ("players_cards", torch.int32, 20),
           ("players_card_idx", torch.int32, 0), ("player_card_sums", torch.int32, 0), ("has_ace", torch.bool, 0),
           ("dealer_cards", torch.int32, 20), ("dealer_card_idx", torch.int32, 0), ("dealer_upcard", torch.int32, 0),
           ("dealer_card_sums", torch.int32, 0), ("dealer_has_ace", torch.bool, 0))


class _Compat(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _, _ in _COMPAT]


class BlackJack(gym.Env):
    metadata = {"render.modes": ["human"]}
    NUM_ACTIONS = 2  # hit, stand
    WIN_REWARD, LOSS_REWARD = 1, -1

    def __init__(self, device, batch_size, seed=0, env_base=0):
        super().__init__()
        self.device = _lib.require_cuda(device)
        self._lib = _lib.load()
        self.batch_size = int(batch_size)
        self.action_space = spaces.Discrete(self.NUM_ACTIONS)
        self.obs_size = 3
        self.observation_space = spaces.Box(low=0, high=10000, shape=(self.obs_size,), dtype=np.float32)
        self.g = torch.arange(self.batch_size, device=self.device, dtype=torch.int32)
        self.seed_value, self.env_base, self._episode = int(seed), int(env_base), 0
        self._st = self._deck = None
        self._view = None

    # ---- reference API ------------------------------------------------------------------------
    def reset(self, seed=None, options=None):
        """options["decks"]: optional [N,52] permutations of 0..51 to deal from (chance injection);
        otherwise decks come from the chance tape (seed, env_base + i, episode)."""
        if seed is not None:
            self.seed_value, self._episode = int(seed), 0
        n, dev = self.batch_size, self.device
        injected = None
        if options and options.get("decks") is not None:
            injected = torch.as_tensor(options["decks"], device=dev)
            if tuple(injected.shape) != (n, 52):
                raise ValueError(f"decks must have shape {(n, 52)}, got {tuple(injected.shape)}")
            injected = injected.to(torch.uint8).contiguous()
        # fresh tensors on every reset, like the reference (:31-44)
        self._st = torch.empty(n, dtype=torch.int32, device=dev)
        self._deck = torch.empty((52, n), dtype=torch.uint8, device=dev)
        self.obs = torch.empty((n, 3), dtype=torch.int32, device=dev)
        self.rewards = torch.empty(n, dtype=torch.int32, device=dev)
        self.terminated = torch.empty(n, dtype=torch.bool, device=dev)
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_blackjack_reset(
                ptr(self._st), ptr(self._deck), ptr(injected), c_u64(self.seed_value), c_u64(self.env_base),
                c_u32(self._episode), ptr(self.obs), ptr(self.rewards), ptr(self.terminated), c_i(n),
                _lib.cuda_stream(dev)), "pulse_blackjack_reset")
        self._episode += 1
        self._view = None
        return self.obs, self.get_info()

    def step(self, actions):
        if self._st is None:
            raise _lib.PulseError("BlackJack.step called before reset")
        n, dev = self.batch_size, self.device
        actions = torch.as_tensor(actions, device=dev)
        assert actions.shape == (n,)  # blackjack.py:114
        if actions.dtype != torch.int64 or not actions.is_contiguous():
            actions = actions.to(torch.int64).contiguous()
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_blackjack_step(
                ptr(self._st), ptr(self._deck), ptr(actions), ptr(self.obs), ptr(self.rewards), ptr(self.terminated),
                c_i(n), _lib.cuda_stream(dev)), "pulse_blackjack_step")
        self._view = None
        return self.obs, self.rewards, self.terminated, None, self.get_info()

    def rollout(self, actions, reset_first=False, n_hands=1):
        """K steps in ONE launch (pulse_blackjack_rollout): actions int64 or uint8 [K, N] -> (obs int32 [K, N, 3], rewards
        int32 [K, N], terminated bool [K, N]), exactly what K step() calls would return one after another.  With
        reset_first the hand is dealt first (= reset()) and the deal's observation is returned as a fourth value.
        n_hands > 1: actions [n_hands * K, N]; every hand after the first (and the first with reset_first) starts with a
        fresh deal -- what n_hands x (reset(), K x step()) would return; the deals' observations come back [n_hands, N, 3].
        The env's obs / rewards / terminated attributes are left at the last step's values."""
        n, dev = self.batch_size, self.device
        actions = torch.as_tensor(actions, device=dev)
        if actions.dtype not in (torch.uint8, torch.int64):
            actions = actions.to(torch.int64)
        actions = actions.contiguous()
        H = int(n_hands)
        if actions.dim() != 2 or actions.shape[1] != n or H < 1 or actions.shape[0] % H:
            raise ValueError(f"actions must have shape (n_hands * K, {n}), got {tuple(actions.shape)}")
        KT = int(actions.shape[0])
        if self._st is None:
            if not reset_first:
                raise _lib.PulseError("BlackJack.rollout called before reset")
            self._st = torch.empty(n, dtype=torch.int32, device=dev)
            self._deck = torch.empty((52, n), dtype=torch.uint8, device=dev)
        obs = torch.empty((KT, n, 3), dtype=torch.int32, device=dev)
        rewards = torch.empty((KT, n), dtype=torch.int32, device=dev)
        term = torch.empty((KT, n), dtype=torch.bool, device=dev)
        deals = H - (0 if reset_first else 1)
        first_obs = torch.empty((H, n, 3) if H > 1 else (n, 3), dtype=torch.int32, device=dev) if deals else None
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_blackjack_rollout(
                ptr(self._st), ptr(self._deck), ptr(actions), c_i(actions.element_size()), c_i(KT // H), c_i(H),
                c_i(1 if reset_first else 0), c_u64(self.seed_value), c_u64(self.env_base), c_u32(self._episode),
                ptr(first_obs), ptr(obs), ptr(rewards), ptr(term), c_i(n), _lib.cuda_stream(dev)), "pulse_blackjack_rollout")
        self._episode += deals
        if KT:
            self.obs, self.rewards, self.terminated = obs[-1], rewards[-1], term[-1]
        self._view = None
        return (obs, rewards, term, first_obs) if deals else (obs, rewards, term)

    def get_obs(self):
        return self.obs

    def get_info(self):
        return {}

    def close(self):
        self._st = self._deck = self._view = None

    # ---- the reference's tensor attributes, rebuilt from the packed state on demand ------------------
    def unpack(self) -> dict:
        if self._view is None:
            n, dev = self.batch_size, self.device
            out, s = {}, _Compat()
            for name, dt, w in _COMPAT:
                out[name] = torch.empty((n, w) if w else (n,), dtype=dt, device=dev)
                setattr(s, name, out[name].data_ptr())
            with _lib.device_guard(dev):
                _lib.check(self._lib.pulse_blackjack_unpack(ptr(self._st), ptr(self._deck), C.byref(s), c_i(n),
                                                            _lib.cuda_stream(dev)), "pulse_blackjack_unpack")
            self._view = out
        return self._view

    def __getattr__(self, name):
        if name in _COMPAT_NAMES and self.__dict__.get("_st") is not None:
            return self.unpack()[name]
        raise AttributeError(name)


_COMPAT_NAMES = frozenset(n for n, _, _ in _COMPAT)
