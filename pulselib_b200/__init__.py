"""pulselib_b200: B200-native batched env-step engine behind Pulselib's env API.

Python env classes (`pulselib_b200.envs`) hand torch tensors through a ctypes C-ABI
(`include/pulse_envs.h`, `libpulse_envs.so`) to hand-written sm_100a CUDA kernels.
"""
__version__ = "0.1.0"

from . import gym_compat  # noqa: F401
