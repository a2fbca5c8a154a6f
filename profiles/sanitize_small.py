"""Small driver for compute-sanitizer (memcheck / racecheck / initcheck): every kernel family once, at ragged
sizes, so the tools finish in a minute.  Usage on the GPU box:
    compute-sanitizer --tool memcheck python profiles/sanitize_small.py"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from pulselib_b200 import policy  # noqa: E402
from pulselib_b200.envs.blackjack import BlackJack  # noqa: E402
from pulselib_b200.envs.particle2d import Particle2D  # noqa: E402
from pulselib_b200.envs.poker import PokerGPU  # noqa: E402
from pulselib_b200.envs.poker_packed import PokerGPUPacked  # noqa: E402
from pulselib_b200.envs.tetris import TetrisBatched  # noqa: E402
from pulselib_b200.envs.tfe import TFEBatched  # noqa: E402
from pulselib_b200.hand_stats import HandStats  # noqa: E402
from pulselib_b200.replay import TransitionRing  # noqa: E402

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev)
g.manual_seed(0)
for P, M, N in ((2, 2, 1000), (6, 6, 777), (10, 16, 301)):
    env = PokerGPUPacked(dev, n_players=P, max_players=M, n_games=N, seed=1, auto_reset=True)
    hs, ring = HandStats(env, 0), TransitionRing(dev, 2 * N, env.obs_size)
    obs, _ = env.reset()
    hs.begin_hand()
    prev = obs.clone()
    kinds = bytes([0] + [3 + (s % 3) for s in range(1, P)])
    for t in range(12):
        a = torch.randint(0, 13, (N,), dtype=torch.int64, device=dev, generator=g)
        if t % 2:
            o, r, d, _, info = env.step_policy(kinds, actions=a, return_seat=True)
            ring.append(prev, info["actions"], r, o, d, info["seat_idx_before"], None, 0)
        else:
            o, r, d, _, _ = env.step(a)
        hs.update(d)
        prev = o.clone()
    compat = PokerGPU(dev, [], n_players=P, max_players=M, n_games=N)
    o, info = compat.reset()
    for t in range(8):
        a = torch.zeros(N, dtype=torch.int64, device=dev)
        policy.build_actions(o, a, info["seat_idx"], [None] * P, ["random"] + ["small_ball"] * (P - 1), dev)
        o, r, d, _, info = compat.step(a)
for N in (1, 1000):
    e = TFEBatched(dev, N, side=4, seed=1); e.reset()
    t = TetrisBatched(dev, N, seed=1); t.reset()
    b = BlackJack(dev, N, seed=1); b.reset()
    p = Particle2D(dev, N, seed=1); p.reset()
    for k in range(6):
        e.step(torch.randint(0, 4, (N,), dtype=torch.int64, device=dev, generator=g))
        t.step(torch.randint(0, 40, (N,), dtype=torch.int64, device=dev, generator=g))
        b.step(torch.randint(0, 2, (N,), dtype=torch.int64, device=dev, generator=g))
        p.step(torch.rand((N, 2), device=dev, generator=g) * 2 - 1)
torch.cuda.synchronize()
print("sanitize_small: done")
