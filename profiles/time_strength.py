"""Times pulse_poker_packed_enumerated_strength (8(f) N4) on 1M heads-up tables per street (CUDA events)."""
import json

import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pulselib_b200.envs.poker_packed import PokerGPUPacked  # noqa: E402

dev = torch.device("cuda:0")
N = 1 << 20
env = PokerGPUPacked(dev, seed=3, n_players=2, max_players=2, n_games=N, starting_bbs=100, max_bbs=300, w1=.5, w2=.3,
                     K=100, alpha=50)
env.reset()
call = torch.ones(N, dtype=torch.int64, device=dev)
out = {}
for t in range(8):
    stages = env.unpack()["stages"]
    hist = torch.bincount(stages.to(torch.int64), minlength=6).tolist()
    env.enumerated_strength()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        s = env.enumerated_strength()
    b.record()
    torch.cuda.synchronize()
    out[f"after_{t}_calls"] = {"stage_hist": hist, "ms": a.elapsed_time(b) / 3, "mean": float(s.mean())}
    env.step(call)
print(json.dumps(out))
