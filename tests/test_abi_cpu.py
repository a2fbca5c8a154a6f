"""CPU: the C-ABI library loads and exports every symbol include/pulse_envs.h declares (no compute calls
without a GPU), argument validation works on the host side, the registration ids resolve, and the env
classes refuse a CPU device instead of falling back."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "pulse_envs.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pulse_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from pulselib_b200 import _lib

    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 24
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/pulse_envs.h but not exported"
    assert lib.pulse_version() >= 100


def test_host_side_argument_validation():
    from pulselib_b200 import _lib

    lib = _lib.load()
    lib.pulse_last_error.restype = C.c_char_p
    null = C.c_void_p(0)
    # n_envs == 0 is a no-op that needs no device; bad sizes are rejected before any launch
    assert lib.pulse_tfe_step(null, null, 8, null, null, C.c_uint64(0), C.c_uint64(0), 0, null, 0, null, null, null, null, 0, 4, null) == 0
    assert lib.pulse_tfe_step(null, null, 8, null, null, C.c_uint64(0), C.c_uint64(0), 0, null, 0, null, null, null, null, 1, 5, null) == -1
    assert b"side" in lib.pulse_last_error()
    assert lib.pulse_tfe_step(null, null, 4, null, null, C.c_uint64(0), C.c_uint64(0), 0, null, 0, null, null, null, null, 1, 4, null) == -1
    assert lib.pulse_blackjack_step(null, null, null, null, null, null, 3, null) == -1
    assert lib.pulse_tetris_step(null, null, null, null, null, 8, null, null, C.c_uint64(0), C.c_uint64(0), 0, null, 0, null, -1, null) == -1
    assert lib.pulse_particle2d_step(null, null, null, null, null, null, null, C.c_float(0.1), 10, 0, null) == 0


def test_registration_ids_resolve():
    import importlib

    from pulselib_b200 import registration

    registration.register_all()
    import gymnasium as gym

    for env_id in ("Pulse-2048-v2", "Pulse-Poker-GPU-v1", "Pulse-Blackjack-Standard", "Pulse-Particle-2d"):
        spec = gym.registry[env_id]   # the reference's ids (environments/__init__.py:3-31)
        mod, _, attr = spec.entry_point.partition(":")
        assert hasattr(importlib.import_module(mod), attr)


@pytest.mark.parametrize("make", [
    lambda: __import__("pulselib_b200.envs.blackjack", fromlist=["x"]).BlackJack("cpu", 4),
    lambda: __import__("pulselib_b200.envs.tfe", fromlist=["x"]).TFEBatched("cpu", 4),
    lambda: __import__("pulselib_b200.envs.tetris", fromlist=["x"]).TetrisBatched("cpu", 4),
    lambda: __import__("pulselib_b200.envs.poker_packed", fromlist=["x"]).PokerGPUPacked("cpu", n_games=4),
])
def test_no_cpu_fallback(make):
    from pulselib_b200._lib import PulseError

    with pytest.raises(PulseError):
        make()


def test_product_never_imports_oracle():
    for path in (ROOT / "pulselib_b200").rglob("*.py"):
        text = path.read_text()
        assert "import oracle" not in text and "from oracle" not in text, path
