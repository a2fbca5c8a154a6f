"""ctypes binding of libpulse_envs.so (include/pulse_envs.h).  There is no CPU fallback: if the
library cannot be loaded or built, importing an env class works but constructing one raises."""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

_PKG = Path(__file__).resolve().parent
_LIB = None
_LOCK = threading.Lock()

c_ptr = C.c_void_p
c_i, c_u32, c_u64, c_f, c_sz = C.c_int, C.c_uint32, C.c_uint64, C.c_float, C.c_size_t


class PulseError(RuntimeError):
    pass


def lib_path() -> Path:
    return _PKG / "libpulse_envs.so"


def load(build_if_missing: bool = True):
    """Returns the loaded CDLL; builds it with nvcc first when it is missing or stale."""
    global _LIB
    if _LIB is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = lib_path()
        if build_if_missing:
            from . import build as _build

            try:
                if _build.is_stale():
                    _build.build()
            except Exception as e:  # no nvcc: fine if a built library is already there
                if not path.exists():
                    raise PulseError(f"libpulse_envs.so is missing and could not be built: {e}") from e
        if not path.exists():
            raise PulseError(f"{path} not found; run `python -m pulselib_b200.build`")
        lib = C.CDLL(str(path))
        lib.pulse_last_error.restype = C.c_char_p
        lib.pulse_version.restype = c_i
        _LIB = lib
        return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().pulse_last_error().decode("utf-8", "replace")
        raise PulseError(f"{what or 'libpulse_envs'} failed (code {rc}): {msg}")


def ptr(t):
    """Device (or host) address of a contiguous torch tensor / numpy array, or NULL for None."""
    if t is None:
        return c_ptr(0)
    if hasattr(t, "data_ptr"):
        if not t.is_contiguous():
            raise PulseError("non-contiguous tensor passed to libpulse_envs")
        return c_ptr(t.data_ptr())
    return c_ptr(t.ctypes.data)


def cuda_stream(device):
    import torch

    return c_ptr(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(device):
    """Normalises `device` and refuses anything but a CUDA device (the engine is CUDA-only)."""
    import torch

    dev = torch.device(device)
    if dev.type != "cuda":
        raise PulseError(
            f"pulselib_b200 envs run on sm_100a CUDA kernels only; got device={device!r}. "
            "There is deliberately no CPU fallback.")
    if not torch.cuda.is_available():
        raise PulseError("CUDA is not available in this process")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


class device_guard:
    """cudaSetDevice for the duration of a call when the env lives on a non-current device."""

    def __init__(self, dev):
        import torch

        self._torch = torch
        self._idx = dev.index
        self._prev = None

    def __enter__(self):
        cur = self._torch.cuda.current_device()
        if cur != self._idx:
            self._prev = cur
            self._torch.cuda.set_device(self._idx)

    def __exit__(self, *exc):
        if self._prev is not None:
            self._torch.cuda.set_device(self._prev)
        return False
