"""2048 behind the reference's env surface (/root/reference/environments/2048/TFE.py:111-189).

`TFEBatched` is the engine: N boards stepped by one fused CUDA launch (pulse_tfe_step).  The reference
has no batched form -- it steps one numpy board per object and batches with gym.make_vec(..., "sync")
(scripts/TFE/train.py:36) -- so the batched class applies the reference's per-board rules to every row
and adds the vector-env conventions: `auto_reset` (a finished board restarts on its next step, like
gymnasium's next-step autoreset) and one global step counter that keys the chance tape.

`TFE(board_height, board_width)` is the registered `Pulse-2048-v2` single-board env with the
reference's exact signature and return types (numpy int32 board, python int reward, bool done, False,
{'score': int}); it is a batch of one on the GPU.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .._lib import c_i, c_u32, c_u64, ptr
from ..gym_compat import gym, spaces


class TFEBatched(gym.Env):
    metadata = {"render_modes": ["human"]}

    def __init__(self, device, n_envs, side=4, seed=0, env_base=0, auto_reset=True, return_obs=True, device_counter=False):
        super().__init__()
        self.device = _lib.require_cuda(device)
        self._lib = _lib.load()
        if int(side) not in (2, 3, 4):
            raise ValueError("side must be 2, 3 or 4 (square boards; 4 bits per cell in one uint64)")
        self.n_envs, self.side = int(n_envs), int(side)
        self.n = self.m = self.side
        self.seed_value, self.env_base, self.auto_reset = int(seed), int(env_base), bool(auto_reset)
        self.return_obs = bool(return_obs)
        self.action_space = spaces.Discrete(4)  # 0 left, 1 up, 2 right, 3 down (TFE.py:154-178)
        self.observation_space = spaces.Box(low=0, high=np.inf, shape=(self.side, self.side), dtype=np.int32)
        n, dev = self.n_envs, self.device
        self.boards = torch.zeros(n, dtype=torch.int64, device=dev)  # packed exponents, 4 bits per cell
        self.total_score = torch.zeros(n, dtype=torch.int64, device=dev)
        self.done = torch.zeros(n, dtype=torch.bool, device=dev)
        self.rewards = torch.zeros(n, dtype=torch.int32, device=dev)
        self.obs = torch.zeros((n, self.side, self.side), dtype=torch.int32, device=dev) if return_obs else None
        self._t = 0
        self._started = False
        # device_counter: the tape counter lives on the device and is advanced by each step launch, so a
        # captured step can be replayed from a CUDA graph (pulselib_b200.graphs) with fresh chance
        self._ctr = torch.zeros(2, dtype=torch.int32, device=dev) if device_counter else None
        with _lib.device_guard(dev):  # row-move table of this device (built once)
            _lib.check(self._lib.pulse_tfe_init(_lib.cuda_stream(dev)), "pulse_tfe_init")

    def _chance(self, c, words):
        """uint32 draws [N, words] as an int32-typed device tensor (same bits)."""
        if c is None:
            return None
        if isinstance(c, torch.Tensor) and c.dtype == torch.int32:
            t = c.to(self.device)
        else:
            a = c.cpu().numpy() if isinstance(c, torch.Tensor) else np.asarray(c)
            t = torch.from_numpy(np.ascontiguousarray(a.astype(np.uint32)).view(np.int32)).to(self.device)
        if tuple(t.shape) != (self.n_envs, words):
            raise ValueError(f"chance must have shape {(self.n_envs, words)}, got {tuple(t.shape)}")
        return t.contiguous()

    def reset(self, seed=None, options=None):
        """options["chance"]: optional uint32 [N,4] spawn draws (two spawns) instead of the tape."""
        if seed is not None:
            self.seed_value, self._t = int(seed), 0
        if self._ctr is not None and self._started:     # fold the device counter back into the host one
            self._t += int(self._ctr[0].item())
            self._ctr.zero_()
        chance = self._chance((options or {}).get("chance"), 4)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tfe_reset(
                ptr(self.boards), ptr(self.total_score), ptr(self.done), ptr(chance), c_u64(self.seed_value),
                c_u64(self.env_base), c_u32(self._t), ptr(self.obs), c_i(self.n_envs), c_i(self.side),
                _lib.cuda_stream(self.device)), "pulse_tfe_reset")
        self._t += 1
        self._started = True
        return self.get_obs(), self.get_info()

    def step(self, actions, chance=None, reset_chance=None):
        """actions: uint8 or int64 [N] in 0..3.  Returns (obs int32 [N,side,side] (same tensor every call,
        as the reference returns its board array itself, TFE.py:130-132, :189), reward int32 [N], done
        bool [N], False, {'score': int64 [N]})."""
        if not self._started:
            raise _lib.PulseError("TFEBatched.step called before reset")
        actions = torch.as_tensor(actions, device=self.device)
        if tuple(actions.shape) != (self.n_envs,):
            raise ValueError(f"actions must have shape {(self.n_envs,)}, got {tuple(actions.shape)}")
        if actions.dtype not in (torch.uint8, torch.int64) or not actions.is_contiguous():
            actions = actions.to(torch.int64).contiguous()
        chance, reset_chance = self._chance(chance, 2), self._chance(reset_chance, 4)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tfe_step(
                ptr(self.boards), ptr(actions), c_i(actions.element_size()), ptr(chance), ptr(reset_chance),
                c_u64(self.seed_value), c_u64(self.env_base), c_u32(self._t), ptr(self._ctr), c_i(self.auto_reset),
                ptr(self.obs),
                ptr(self.rewards), ptr(self.done), ptr(self.total_score), c_i(self.n_envs), c_i(self.side),
                _lib.cuda_stream(self.device)), "pulse_tfe_step")
        if self._ctr is None:
            self._t += 1
        return self.get_obs(), self.rewards, self.done, False, self.get_info()

    def rollout(self, actions, return_obs=None):
        """K steps in ONE launch (pulse_tfe_rollout) for an open-loop action tape: actions uint8 / int64 [K, N] ->
        (obs int32 [K, N, side, side] or None, reward int32 [K, N], done bool [K, N]) -- exactly what K step() calls
        would return one after another (chance from the tape at the same counters)."""
        if not self._started:
            raise _lib.PulseError("TFEBatched.rollout called before reset")
        if self._ctr is not None:
            raise _lib.PulseError("rollout keeps the tape counter on the host: construct the env without device_counter")
        actions = torch.as_tensor(actions, device=self.device)
        if actions.dim() != 2 or actions.shape[1] != self.n_envs:
            raise ValueError(f"actions must have shape (K, {self.n_envs}), got {tuple(actions.shape)}")
        if actions.dtype not in (torch.uint8, torch.int64) or not actions.is_contiguous():
            actions = actions.to(torch.int64).contiguous()
        K, n, dev = int(actions.shape[0]), self.n_envs, self.device
        want_obs = self.return_obs if return_obs is None else bool(return_obs)
        obs = torch.empty((K, n, self.side, self.side), dtype=torch.int32, device=dev) if want_obs else None
        rew = torch.empty((K, n), dtype=torch.int32, device=dev)
        done = torch.empty((K, n), dtype=torch.bool, device=dev)
        with _lib.device_guard(dev):
            _lib.check(self._lib.pulse_tfe_rollout(
                ptr(self.boards), ptr(actions), c_i(actions.element_size()), c_i(K), c_u64(self.seed_value),
                c_u64(self.env_base), c_u32(self._t), c_i(self.auto_reset), ptr(obs), ptr(rew), ptr(done), ptr(self.done),
                ptr(self.total_score), c_i(n), c_i(self.side), _lib.cuda_stream(dev)), "pulse_tfe_rollout")
        self._t += K
        if K:
            self.rewards = rew[-1]
            if want_obs and self.obs is not None:
                self.obs = obs[-1]
        return obs, rew, done

    def get_obs(self):
        return self.obs if self.return_obs else self.boards

    def get_info(self):
        return {"score": self.total_score}

    def board_values(self) -> torch.Tensor:
        """int32 [N, side, side] tile values, unpacked from the packed boards."""
        out = torch.empty((self.n_envs, self.side, self.side), dtype=torch.int32, device=self.device)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tfe_unpack(ptr(self.boards), ptr(out), c_i(self.n_envs), c_i(self.side),
                                                  _lib.cuda_stream(self.device)), "pulse_tfe_unpack")
        return out

    def set_board_values(self, values) -> None:
        """Loads int32 [N, side, side] tile values (0 or powers of two up to 32768)."""
        v = torch.as_tensor(values, device=self.device).to(torch.int32).contiguous()
        if tuple(v.shape) != (self.n_envs, self.side, self.side):
            raise ValueError("values must have shape [n_envs, side, side]")
        bad = torch.zeros(1, dtype=torch.int32, device=self.device)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tfe_pack(ptr(v), ptr(self.boards), ptr(self.done), ptr(bad), c_i(self.n_envs), c_i(self.side),
                                                _lib.cuda_stream(self.device)), "pulse_tfe_pack")
        if int(bad.item()):
            raise ValueError("board values must be 0 or a power of two in 2..32768")
        self._started = True

    def domain_check(self, raise_on_overflow=True):
        """(boards holding a 32768 tile, boards holding two or more).  The packed board stores tile exponents up to 15;
        two 32768 tiles would stay unmerged where the reference makes 65536 (csrc/tfe_body.cuh) -- this is the flag
        that says a batch is about to leave the domain in which the engine is bit-exact."""
        counts = torch.zeros(2, dtype=torch.int32, device=self.device)
        with _lib.device_guard(self.device):
            _lib.check(self._lib.pulse_tfe_domain_check(ptr(self.boards), ptr(counts), c_i(self.n_envs),
                                                        _lib.cuda_stream(self.device)), "pulse_tfe_domain_check")
        one, two = (int(v) for v in counts.tolist())
        if two and raise_on_overflow:
            raise _lib.PulseError(f"{two} board(s) hold two 32768 tiles: the next merge would leave the packed domain "
                                  "(tiles above 32768 are not representable in 4 bits per cell)")
        return one, two

    def is_game_over(self):
        return self.done

    def close(self):
        pass


class TFE(gym.Env):
    """`Pulse-2048-v2`: one board, the reference's signature and return types (TFE.py:111-189)."""

    def __init__(self, board_height, board_width, device="cuda", seed=0):
        super().__init__()
        if board_height != board_width:
            # the reference's rotate buffer is zeros_like(board) (TFE.py:127): only squares work there too
            raise ValueError("TFE boards must be square")
        self.n, self.m = board_height, board_width
        self._env = TFEBatched(device, 1, side=board_height, seed=seed, auto_reset=False)
        self.action_space = self._env.action_space
        self.observation_space = self._env.observation_space
        self.board = np.zeros((self.n, self.m), dtype=np.int32)
        self.total_score = 0
        self.render_mode = "human"

    def get_obs(self):
        return self.board

    def get_info(self):
        return {"score": self.total_score}

    def is_game_over(self):
        return bool(self._env.done[0].item())

    def reset(self, seed=None, options=None):
        obs, _ = self._env.reset(seed=seed, options=options)
        self.board = obs[0].cpu().numpy()
        self.total_score = 0
        return self.get_obs(), self.get_info()

    def step(self, action):
        a = torch.tensor([int(action)], dtype=torch.int64, device=self._env.device)
        obs, rew, done, _, info = self._env.step(a)
        np.copyto(self.board, obs[0].cpu().numpy())   # same array object every step, like the reference (:181)
        if self.board.max() >= 32768:                 # a single board: the audit is free here
            self._env.domain_check()
        self.total_score = int(info["score"][0].item())
        return self.get_obs(), int(rew[0].item()), bool(done[0].item()), False, self.get_info()
