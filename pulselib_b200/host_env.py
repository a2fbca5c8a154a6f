"""PokerGPU with host (numpy) buffers: the end-to-end path of the engine.

`PokerHostEnv` keeps the reference's step contract -- actions in, float32 observations [N, 13 + 3 (M - 1)], float32
rewards and bool dones out (/root/reference/environments/Poker/PokerGPU.py:527-633) -- with every array in HOST
memory.  It binds `pulse_poker_host_*` (include/pulse_envs.h, csrc/host_env.cu): 1 B per table goes to the GPU, one
packed 28 B record per LIVE table comes back, and a pool of native threads expands the records into `self.obs`,
`self.rewards`, `self.dones` -- the same three arrays on every call, as the reference reuses `self.obs`.
bench.py's `e2e` number is this class."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, handranks
from .envs.poker import _Config


_PINNED_KEEPALIVE = []


class PokerHostEnv:
    NUM_ACTIONS = 13

    def __init__(self, device=0, n_players=6, max_players=10, n_games=100, starting_bbs=100, max_bbs=1000, w1=.5, w2=.5,
                 K=20, alpha=300, seed=0, env_base=0, chunks=8, threads=0):
        import torch

        if not torch.cuda.is_available():
            raise _lib.PulseError("PokerHostEnv needs a CUDA device: the engine has no CPU fallback")
        self._lib = _lib.load()
        dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
        self.device_index = dev.index if dev.index is not None else torch.cuda.current_device()
        self.n_players, self.max_players, self.n_games = int(n_players), int(max_players), int(n_games)
        self.active_players = self.n_players
        self.obs_size = 13 + 3 * (self.max_players - 1)
        cfg = _Config(self.n_games, self.n_players, self.n_players, self.max_players, int(starting_bbs), int(max_bbs),
                      int(K), int(alpha), float(w1), float(w2))
        table = handranks.load_table()
        self._h = C.c_void_p()
        _lib.check(self._lib.pulse_poker_host_create(C.byref(self._h), C.byref(cfg), C.c_void_p(table.ctypes.data),
                                                     C.c_size_t(table.size), C.c_uint64(int(seed)), C.c_uint64(int(env_base)),
                                                     C.c_int(self.device_index), C.c_int(int(chunks)), C.c_int(int(threads))),
                   "pulse_poker_host_create")
        # 64-byte aligned so heads-up rows (64 B) are whole cache lines for the expander's streaming stores
        self.obs = _aligned((self.n_games, self.obs_size), np.float32)
        self.rewards = _aligned((self.n_games,), np.float32)
        self.dones = _aligned((self.n_games,), np.uint8)

    def reset(self, active_players=None, rotation=0, expand=True):
        """expand=False: records only (for step_records() consumers): self.obs is not written"""
        ap = self.n_players if active_players is None else int(active_players)
        _lib.check(self._lib.pulse_poker_host_reset(self._h, C.c_int(ap), C.c_int(int(rotation)),
                                                    _p(self.obs) if expand else C.c_void_p(0),
                                                    _p(self.rewards), _p(self.dones)), "pulse_poker_host_reset")
        self.active_players = ap
        return self.obs, {"active_players": ap}

    def step(self, actions):
        a = np.ascontiguousarray(actions)
        if a.dtype not in (np.uint8, np.int64):
            a = a.astype(np.int64)
        if a.shape != (self.n_games,):
            raise ValueError(f"actions must have shape {(self.n_games,)}, got {a.shape}")
        _lib.check(self._lib.pulse_poker_host_step(self._h, _p(a), C.c_int(a.dtype.itemsize), _p(self.obs), _p(self.rewards),
                                                   _p(self.dones)), "pulse_poker_host_step")
        return self.obs, self.rewards, self.dones.view(np.bool_), False, {"active_players": self.active_players}

    def step_records(self, actions):
        """step() without the float32 expansion (pulse_poker_host_step_records): only `self.dones` is updated; the step's
        packed records stay in the env's pinned buffer -- see records() / live_mask().  Do not mix with step() between
        resets."""
        a = np.ascontiguousarray(actions)
        if a.dtype not in (np.uint8, np.int64):
            a = a.astype(np.int64)
        if a.shape != (self.n_games,):
            raise ValueError(f"actions must have shape {(self.n_games,)}, got {a.shape}")
        _lib.check(self._lib.pulse_poker_host_step_records(self._h, _p(a), C.c_int(a.dtype.itemsize), _p(self.dones)),
                   "pulse_poker_host_step_records")
        return self.dones.view(np.bool_)

    def records(self):
        """[(first_table, n_tables, rows uint32 [n_rows, words])] per chunk of the last call: zero-copy views of the pinned
        record buffer, valid until the next call.  Rows are in table order; a table has one iff its live_mask() bit is set."""
        words = self.info()["row_bytes"] // 4
        out = []
        for c in range(self.info()["chunks"]):
            rows, n, first, cnt = C.POINTER(C.c_uint32)(), C.c_longlong(), C.c_longlong(), C.c_longlong()
            _lib.check(self._lib.pulse_poker_host_records(self._h, C.c_int(c), C.byref(rows), C.byref(n), C.byref(first),
                                                          C.byref(cnt)), "pulse_poker_host_records")
            view = np.ctypeslib.as_array(rows, shape=(max(n.value, 1), words))[: n.value]
            out.append((first.value, cnt.value, view))
        return out

    def live_mask(self) -> np.ndarray:
        """bool [N]: the table was live when the last call began (= it has a record in that call's output)"""
        self._lib.pulse_poker_host_live_mask.restype = C.POINTER(C.c_uint32)
        p = self._lib.pulse_poker_host_live_mask(self._h)
        words = np.ctypeslib.as_array(p, shape=((self.n_games + 31) // 32,))
        return np.unpackbits(words.view(np.uint8), bitorder="little")[: self.n_games].astype(bool)

    def info(self) -> dict:
        v = [C.c_int() for _ in range(4)] + [C.c_longlong(), C.c_longlong()]
        _lib.check(self._lib.pulse_poker_host_info(self._h, *[C.byref(x) for x in v]), "pulse_poker_host_info")
        return dict(row_bytes=v[0].value, obs_width=v[1].value, chunks=v[2].value, threads=v[3].value,
                    last_h2d_bytes=v[4].value, last_d2h_bytes=v[5].value)

    def timing(self, reset=False) -> dict:
        out = (C.c_double * 6)()
        _lib.check(self._lib.pulse_poker_host_timing(self._h, out, C.c_int(1 if reset else 0)), "pulse_poker_host_timing")
        keys = ("calls", "stage_s", "submit_s", "expand_wall_s", "worker_wait_s", "worker_expand_s")
        return dict(zip(keys, (float(v) for v in out)))

    @staticmethod
    def pinned_bytes(shape):
        """A page-locked uint8 numpy array (e.g. [steps, n_games] of actions): `step` hands rows of it to the copy
        engine where they lie instead of staging them first."""
        import torch

        t = torch.empty(tuple(int(v) for v in np.atleast_1d(shape)), dtype=torch.uint8).pin_memory()
        a = t.numpy()
        _PINNED_KEEPALIVE.append(t)      # the array does not own the pinned allocation
        return a

    def chunk_timing(self):
        """diagnostic: [(seconds until chunk c's records were seen, seconds until its slice was expanded)] summed over
        the calls since timing(reset=True), as observed by the last worker, measured from the start of each call's
        expansion phase"""
        out = (C.c_double * 128)()
        n = self._lib.pulse_poker_host_chunk_timing(self._h, out, C.c_int(64))
        return [(out[2 * c], out[2 * c + 1]) for c in range(min(n, 64))]

    def close(self):
        if getattr(self, "_h", None):
            self._lib.pulse_poker_host_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _p(a):
    return C.c_void_p(a.ctypes.data)


def _aligned(shape, dtype, align=64):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    raw = np.zeros(n + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n].view(dtype).reshape(shape)


def expand_rows(rows: np.ndarray, max_players: int, vectorised: bool = True):
    """Packed step records (uint32 [n, words]) -> (obs float32 [n, W], reward float32 [n], done uint8 [n]); host only.
    vectorised: through the host env's own expander (AVX2 when available), else the portable definition."""
    lib = _lib.load()
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    n, W = rows.shape[0], 13 + 3 * (max_players - 1)
    obs, rew, done = _aligned((n, W), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint8)
    _lib.check(lib.pulse_poker_wire_expand_host(_p(rows), C.c_longlong(n), C.c_int(max_players), C.c_int(1 if vectorised else 0),
                                                _p(obs), _p(rew), _p(done)), "pulse_poker_wire_expand_host")
    return obs, rew, done


def expand_isa() -> str:
    lib = _lib.load()
    lib.pulse_poker_host_expand_isa.restype = C.c_char_p
    return lib.pulse_poker_host_expand_isa().decode()


class TFEHostEnv:
    """Batched 2048 with host (numpy) buffers: `pulse_tfe_host_*` (include/pulse_envs.h, csrc/tfe_host_env.cu).

    The reference's step contract per board -- int32 tile values [side, side], reward, done
    (/root/reference/environments/2048/TFE.py:152-189) -- for N boards whose arrays live in host memory: 1 B of action
    per board goes to the GPU, the packed 8 B board + reward + done come back, and native threads rebuild the tile
    values in `self.obs`.  Same trajectories as `TFEBatched(seed, env_base, auto_reset)` stepped from the tape."""

    def __init__(self, device=0, n_envs=1024, side=4, seed=0, env_base=0, auto_reset=True, chunks=4, threads=0):
        import torch

        if not torch.cuda.is_available():
            raise _lib.PulseError("TFEHostEnv needs a CUDA device: the engine has no CPU fallback")
        if int(side) not in (2, 3, 4):
            raise ValueError("side must be 2, 3 or 4")
        self._lib = _lib.load()
        dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
        self.device_index = dev.index if dev.index is not None else torch.cuda.current_device()
        self.n_envs, self.side = int(n_envs), int(side)
        self._h = C.c_void_p()
        _lib.check(self._lib.pulse_tfe_host_create(C.byref(self._h), C.c_longlong(self.n_envs), C.c_int(self.side),
                                                   C.c_int(1 if auto_reset else 0), C.c_uint64(int(seed)),
                                                   C.c_uint64(int(env_base)), C.c_int(self.device_index),
                                                   C.c_int(int(chunks)), C.c_int(int(threads))), "pulse_tfe_host_create")
        self.obs = _aligned((self.n_envs, self.side, self.side), np.int32)   # 4 x 4 boards: one 64-byte line each
        self.rewards = _aligned((self.n_envs,), np.int32)
        self.dones = _aligned((self.n_envs,), np.uint8)

    def reset(self):
        _lib.check(self._lib.pulse_tfe_host_reset(self._h, _p(self.obs), _p(self.dones)), "pulse_tfe_host_reset")
        self.rewards[:] = 0
        return self.obs, {}

    def step(self, actions):
        a = np.ascontiguousarray(actions)
        if a.dtype not in (np.uint8, np.int64):
            a = a.astype(np.int64)
        if a.shape != (self.n_envs,):
            raise ValueError(f"actions must have shape {(self.n_envs,)}, got {a.shape}")
        _lib.check(self._lib.pulse_tfe_host_step(self._h, _p(a), C.c_int(a.dtype.itemsize), _p(self.obs), _p(self.rewards),
                                                 _p(self.dones)), "pulse_tfe_host_step")
        return self.obs, self.rewards, self.dones.view(np.bool_), False, {}

    def boards(self) -> np.ndarray:
        """the packed boards of the last call (uint64 [N], a view of the env's pinned buffer)"""
        self._lib.pulse_tfe_host_boards.restype = C.POINTER(C.c_uint64)
        return np.ctypeslib.as_array(self._lib.pulse_tfe_host_boards(self._h), shape=(self.n_envs,))

    def scores(self) -> np.ndarray:
        out = np.zeros(self.n_envs, np.int64)
        _lib.check(self._lib.pulse_tfe_host_scores(self._h, _p(out)), "pulse_tfe_host_scores")
        return out

    def info(self) -> dict:
        v = [C.c_int(), C.c_int(), C.c_longlong(), C.c_longlong()]
        _lib.check(self._lib.pulse_tfe_host_info(self._h, *[C.byref(x) for x in v]), "pulse_tfe_host_info")
        return dict(chunks=v[0].value, threads=v[1].value, last_h2d_bytes=v[2].value, last_d2h_bytes=v[3].value)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.pulse_tfe_host_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def expand_boards(boards: np.ndarray, side: int = 4) -> np.ndarray:
    """Packed 2048 boards (uint64 [n]) -> int32 tile values [n, side, side] through the host env's expander; host only."""
    lib = _lib.load()
    b = np.ascontiguousarray(boards, dtype=np.uint64)
    out = _aligned((b.shape[0], side, side), np.int32)
    _lib.check(lib.pulse_tfe_expand_host(_p(b), C.c_longlong(b.shape[0]), C.c_int(int(side)), _p(out)), "pulse_tfe_expand_host")
    return out
