"""TEST INFRASTRUCTURE ONLY.  Lets the CPU suite run the product's Python env classes (mode switching, mirror on
access, argument checking -- the host logic) without a GPU: inside `hostemu_backend()` the ctypes handle the classes
call is swapped for libhostemu.so -- the very same per-table kernel bodies compiled for the host, entry points named
emu_* with the signatures of pulse_* minus the trailing stream -- and the CUDA-only guards accept CPU tensors.
The product never takes this path: it lives under tests/, is installed only by this context manager, and
tests/test_abi_cpu.py checks that constructing an env without it on a CPU device still raises."""
from __future__ import annotations

import contextlib
import ctypes as C

from . import emu

_MAP = {
    "pulse_poker_reset": "emu_poker_reset", "pulse_poker_step": "emu_poker_step",
    "pulse_poker_packed_reset": "emu_poker_packed_reset", "pulse_poker_packed_step": "emu_poker_packed_step",
    "pulse_poker_unpack": "emu_poker_unpack", "pulse_poker_unpack_full": "emu_poker_unpack_full",
}


class _Backend:
    def __init__(self):
        self._lib = emu.lib()

    def __getattr__(self, name):
        if name in _MAP:
            return getattr(self._lib, _MAP[name])     # the extra trailing stream argument is ignored by the C callee
        raise AttributeError(f"hostemu shim has no {name}")

    def pulse_last_error(self):
        return b"hostemu"


class _NoGuard:
    def __init__(self, dev):
        pass

    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


@contextlib.contextmanager
def hostemu_backend():
    import torch

    from pulselib_b200 import _lib

    saved = (_lib.load, _lib.require_cuda, _lib.cuda_stream, _lib.device_guard, _lib._LIB)
    backend = _Backend()
    _lib.load = lambda build_if_missing=True: backend
    _lib.require_cuda = lambda device: torch.device("cpu")
    _lib.cuda_stream = lambda device: C.c_void_p(0)
    _lib.device_guard = _NoGuard
    try:
        yield backend
    finally:
        _lib.load, _lib.require_cuda, _lib.cuda_stream, _lib.device_guard, _lib._LIB = saved
