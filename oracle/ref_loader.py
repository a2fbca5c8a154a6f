"""TEST INFRASTRUCTURE ONLY.  Imports the reference's own env files by path from /root/reference
(never copied) so the C oracle can be validated against the real thing and golden vectors
generated.  Only usable in the build container: /root/reference does not exist on the GPU box."""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path
from unittest import mock

REF = Path("/root/reference")


def available() -> bool:
    return (REF / "environments" / "Poker" / "PokerGPU.py").exists()


def _load(name: str, rel: str):
    from pulselib_b200 import gym_compat

    gym_compat.install_shim()
    spec = importlib.util.spec_from_file_location(name, REF / rel)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def poker_module():
    return _load("_ref_pokergpu", "environments/Poker/PokerGPU.py")


def make_poker(hand_ranks_np, **kw):
    """Reference PokerGPU on CPU with `hand_ranks_np` standing in for HandRanks.dat (the same
    Path.exists/stat + torch.from_file patch the reference's own test uses,
    tests/poker/test_poker_gpu_prefixed_decks.py:14-34)."""
    import torch

    mod = poker_module()
    table = torch.from_numpy(hand_ranks_np)

    class _Stat:
        st_size = table.numel() * 4

    with mock.patch.object(Path, "exists", lambda self: True), \
            mock.patch.object(Path, "stat", lambda self, **k: _Stat()), \
            mock.patch.object(torch, "from_file", lambda **k: table):
        env = mod.PokerGPU(device=torch.device("cpu"), agents=[], **kw)
    return env


def blackjack_module():
    return _load("_ref_blackjack", "environments/blackjack/blackjack.py")


def particle2d_module():
    return _load("_ref_particle2d", "environments/Particle2D/Particle2D.py")


def tfe_module():
    return _load("_ref_tfe", "environments/2048/TFE.py")
