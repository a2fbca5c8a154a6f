"""Generates tests/golden/poker_replay.npz with the reference's OWN replay buffer
(/root/reference/utils/ReplayBuffer.py::ReplayBufferTorch, imported in place; build container only) fed the way the
reference's training loop selects rows (/root/reference/scripts/Poker/trainGPU.py:79-94: q_mask = seat_idx == q_seat,
active = q_mask & ~terminated_before, five masked gathers).  Stored: the per-step inputs and the ring's final
contents, position and size.  The reference's buffers are int32 and only accept int32 sources, so the rows are cast
(truncation toward zero) before `add`; the replay compares after the same cast.

    python tests/golden/make_replay_golden.py
"""
import importlib.util
import sys
import types
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent / "poker_replay.npz"
sys.modules.setdefault("turtle", types.SimpleNamespace(position=None))   # `from turtle import position` (unused) needs tkinter
spec = importlib.util.spec_from_file_location("_ref_replay", "/root/reference/utils/ReplayBuffer.py")
mod = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mod)


def main():
    rng = np.random.default_rng(20260401)
    rec = {}
    for tag, N, W, P, q_seat, steps, cap in (("wrap", 1200, 16, 2, 1, 9, 2048), ("six", 900, 28, 6, 2, 7, 8000)):
        buf = mod.ReplayBufferTorch("", cap, "cpu", W, 1)
        terminated = torch.zeros(N, dtype=torch.bool)
        state = torch.from_numpy(rng.integers(-1, 300, (N, W)).astype(np.float32))
        tr = {k: [] for k in ("state", "next_state", "actions", "rewards", "dones", "seat", "term_before")}
        for t in range(steps):
            seat = torch.from_numpy(rng.integers(0, P, N).astype(np.int32))
            actions = torch.from_numpy(rng.integers(0, 13, N))
            rewards = torch.from_numpy((rng.standard_normal(N) * 30).astype(np.float32))
            next_state = torch.from_numpy(rng.integers(-1, 300, (N, W)).astype(np.float32))
            dones = torch.from_numpy(rng.random(N) < 0.2) | terminated
            # --- trainGPU.py:79-94 ---
            q_mask = seat == q_seat
            active_games = q_mask & ~terminated
            term_before = terminated.clone()
            terminated = terminated | dones
            if active_games.any():
                i32 = torch.int32  # ReplayBufferTorch's buffers are int32 and index_put wants matching dtypes
                buf.add(state[active_games].to(i32), actions[active_games].unsqueeze(1).to(i32),
                        rewards[active_games].to(i32), next_state[active_games].to(i32), dones[active_games].to(i32))
            for k, v in (("state", state), ("next_state", next_state), ("actions", actions), ("rewards", rewards),
                         ("dones", dones), ("seat", seat), ("term_before", term_before)):
                tr[k].append(v.numpy().copy())
            state = next_state
        for k, v in tr.items():
            rec[f"{tag}/{k}"] = np.stack(v)
        rec[f"{tag}/meta"] = np.array([N, W, P, q_seat, steps, cap], np.int64)
        rec[f"{tag}/state_buffer"] = buf.state_buffer.numpy()
        rec[f"{tag}/next_state_buffer"] = buf.next_state_buffer.numpy()
        rec[f"{tag}/action_buffer"] = buf.action_buffer.numpy()[:, 0]
        rec[f"{tag}/reward_buffer"] = buf.reward_buffer.numpy()
        rec[f"{tag}/terminated_buffer"] = buf.terminated_buffer.numpy()
        rec[f"{tag}/position_size"] = np.array([buf.position, buf.size], np.int64)
        print(tag, "position/size", buf.position, buf.size)
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, OUT.stat().st_size)


if __name__ == "__main__":
    main()
