"""Scripted PokerGPU opponents and `build_actions` on the device.

Drop-in for /root/reference/environments/Poker/utils.py:108-123 (`build_actions`) together with the four
scripted GPU agents of /root/reference/environments/Poker/Player.py:79-176: same call signature, same
resulting `actions` tensor.  The reference makes one masked Python call per agent type per step (boolean
masks, gathers, scatters, a randint per type); here every scripted seat of the whole batch is one kernel
launch (`pulse_poker_policy`), and only agents that live on the host side (the Q network, anything with a
Python `action`) are still called the reference's way.

Random draws come from the engine's Philox tape (seed, env_base + table, stream 2, counter) -- pass a new
`counter` every step -- or are injected as `chance2` uint32 [N, 2] (see csrc/poker_policy.cuh).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import c_i, c_u32, c_u64, ptr

EXTERNAL, RANDOM, HEURISTIC_HANDS, TIGHT_AGGRESSIVE, LOOSE_PASSIVE, SMALL_BALL = range(6)
_KIND_BY_NAME = {"random": RANDOM, "heuristic_hands": HEURISTIC_HANDS, "tight_aggressive": TIGHT_AGGRESSIVE,
                 "loose_passive": LOOSE_PASSIVE, "small_ball": SMALL_BALL}


def agent_kind(agent_type) -> int:
    """PokerAgentType member (or its string value) -> PULSE_AGENT_* ; unknown / host-side types -> EXTERNAL."""
    name = getattr(agent_type, "value", agent_type)
    return _KIND_BY_NAME.get(str(name).lower(), EXTERNAL)


def seat_kinds(agent_types) -> bytes:
    kinds = [agent_kind(t) for t in agent_types]
    if not 1 <= len(kinds) <= 16:
        raise ValueError("between 1 and 16 seats")
    return bytes(kinds)


def scripted_actions(state, curr_players, kinds, actions, *, seed=0, env_base=0, counter=0, chance2=None):
    """One launch: actions[g] = rule of seat curr_players[g] for every table whose seat is scripted; other
    tables keep what `actions` holds.  state float32 [N, W], curr_players int [N], actions int64 [N]."""
    dev = _lib.require_cuda(state.device)
    lib = _lib.load()
    if state.dtype != torch.float32 or not state.is_contiguous():
        state = state.to(torch.float32).contiguous()
    seat = curr_players if curr_players.dtype == torch.int32 and curr_players.is_contiguous() \
        else curr_players.to(torch.int32).contiguous()
    if actions.dtype != torch.int64 or not actions.is_contiguous() or actions.device != state.device:
        raise ValueError("actions must be a contiguous int64 tensor on the state's device")
    n = state.shape[0]
    if tuple(seat.shape) != (n,) or tuple(actions.shape) != (n,):
        raise ValueError(f"curr_players and actions must have shape {(n,)}")
    if chance2 is not None:
        chance2 = torch.as_tensor(chance2, device=dev)
        if tuple(chance2.shape) != (n, 2):
            raise ValueError(f"chance2 must have shape {(n, 2)}")
        chance2 = chance2.to(torch.int64).to(torch.int32).contiguous() if chance2.dtype not in (torch.int32, torch.uint32) \
            else chance2.contiguous()
    kb = bytes(kinds)
    with _lib.device_guard(dev):
        _lib.check(lib.pulse_poker_policy(ptr(state), c_i(state.shape[1]), ptr(seat), C.c_char_p(kb), c_i(len(kb)),
                                          ptr(chance2), c_u64(seed), c_u64(env_base), c_u32(counter), ptr(actions),
                                          C.c_longlong(n), _lib.cuda_stream(dev)), "pulse_poker_policy")
    return actions


_counters: dict = {}


def build_actions(state, actions, curr_players, agents, agent_types, device, epsilon=0.1, *, seed=0, env_base=0,
                  counter=None, chance2=None):
    """Same contract as the reference's build_actions (utils.py:108-123): fills `actions` in place.

    Host-side agent types are called exactly as the reference calls them (`get_actions(state[mask])` for
    qlearning, `action(state[mask])` otherwise); every scripted GPU type and `random` is one kernel launch."""
    kinds = seat_kinds(agent_types)
    groups: dict = {}
    for seat, t in enumerate(agent_types):
        if kinds[seat] == EXTERNAL:
            groups.setdefault(getattr(t, "value", t), []).append(seat)
    for name, seats in groups.items():
        mask = torch.zeros_like(curr_players, dtype=torch.bool)
        for s in seats:
            mask |= curr_players == s
        agent = agents[seats[0]]
        if str(name).lower() == "qlearning":
            actions[mask] = agent.get_actions(state[mask])
        else:
            actions[mask] = agent.action(state[mask])
    if any(k != EXTERNAL for k in kinds):
        if counter is None:  # a fresh tape block per call
            key = (int(seed), int(env_base))
            counter = _counters.get(key, 0)
            _counters[key] = counter + 1
        scripted_actions(state, curr_players, kinds, actions, seed=seed, env_base=env_base, counter=counter,
                         chance2=chance2)
    return actions
