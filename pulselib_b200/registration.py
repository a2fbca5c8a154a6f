"""Registers the engine's env classes under the reference's ids
(/root/reference/environments/__init__.py:3-31), so `gym.make("Pulse-Poker-GPU-v1", ...)` and friends
resolve to the CUDA-backed classes.  With the real gymnasium installed it registers there; otherwise
with the stand-in in gym_compat.  The reference's `Pulse-2048-v2` entry point names a module that no
longer exists (`environments.TFE:TFE`, :5); here it resolves."""
from __future__ import annotations

from . import gym_compat
from .gym_compat import gym

_IDS = {
    "Pulse-2048-v2": "pulselib_b200.envs.tfe:TFE",
    "Pulse-Poker-GPU-v1": "pulselib_b200.envs.poker:PokerGPU",
    "Pulse-Blackjack-Standard": "pulselib_b200.envs.blackjack:BlackJack",
    "Pulse-Particle-2d": "pulselib_b200.envs.particle2d:Particle2D",
    # engine extensions (no reference counterpart)
    "Pulse-2048-Batched-v1": "pulselib_b200.envs.tfe:TFEBatched",
    "Pulse-Tetris-Batched-v1": "pulselib_b200.envs.tetris:TetrisBatched",
    "Pulse-Poker-GPU-Packed-v1": "pulselib_b200.envs.poker_packed:PokerGPUPacked",
}


def register_all(override: bool = True) -> None:
    gym_compat.install_shim()
    from gymnasium.envs.registration import register, registry  # the real package or the stand-in

    for env_id, entry in _IDS.items():
        if env_id in registry:
            if not override:
                continue
            del registry[env_id]
        register(id=env_id, entry_point=entry)


__all__ = ["register_all", "gym"]
