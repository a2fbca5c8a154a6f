"""Times the REFERENCE's own CPU path (cerredz/Pulselib imported in place from /root/reference; torch eager on
torch.device("cpu"), numba for 2048) in the build container -- BASELINE.md section 2 / SURVEY.md 8(d) "CPU baseline beside it".
The reference cannot travel to the GPU box, so these numbers are measured here (core count stated) and quoted next to
the C+OpenMP port that bench.py times on the GPU box's host cores.

    python profiles/time_reference_cpu.py > profiles/r2_reference_cpu.json
"""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import ref_loader  # noqa: E402
from pulselib_b200 import handranks  # noqa: E402

CFG = dict(starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)
out = {"host": {"cpu_count": os.cpu_count(), "torch_threads": torch.get_num_threads(),
                "model": next((l.split(":")[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")), "?")},
       "reference": "cerredz/Pulselib imported in place from /root/reference", "results": {}}


def poker(P, N, hand_len, min_s=20.0):
    hr = handranks.load_table()
    env = ref_loader.make_poker(hr, n_players=P, max_players=P, n_games=N, **CFG)
    rng = np.random.default_rng(0)
    acts = torch.from_numpy(rng.integers(0, 13, (hand_len, N)).astype(np.int64))
    env.reset(options={"active_players": False, "rotation": 0})
    for i in range(3):
        env.step(acts[i])
    steps, spent = 0, 0.0
    while spent < min_s:
        t0 = time.perf_counter()
        env.reset(options={"active_players": False, "rotation": 0})
        for i in range(hand_len):
            env.step(acts[i])
        spent += time.perf_counter() - t0
        steps += hand_len * N
    return {"env_steps_per_s": steps / spent, "tables": N, "n_players": P, "hand_len": hand_len,
            "what": "PokerGPU.reset + step (PokerGPU.py:73-157, :527-633), uniform random actions 0..12, whole-batch hands"}


def blackjack(N=4096, min_s=10.0):
    mod = ref_loader.blackjack_module()
    env = mod.BlackJack(device=torch.device("cpu"), batch_size=N)
    rng = np.random.default_rng(1)
    acts = torch.from_numpy(rng.integers(0, 2, (8, N)).astype(np.int64))
    steps, spent, k = 0, 0.0, 0
    env.reset()
    while spent < min_s:
        t0 = time.perf_counter()
        if k % 8 == 0:
            env.reset()
        env.step(acts[k % 8])
        spent += time.perf_counter() - t0
        steps += N; k += 1
    return {"env_steps_per_s": steps / spent, "envs": N, "what": "BlackJack.step (blackjack.py:179-186), reset every 8 steps"}


def particle(N=1 << 24, min_s=10.0):
    mod = ref_loader.particle2d_module()
    env = mod.Particle2D(device=torch.device("cpu"), batch_size=N, dt=0.1, max_steps=200)
    env.reset()
    a = torch.rand(N, 2) * 2 - 1
    env.step(a)
    steps, spent = 0, 0.0
    while spent < min_s:
        t0 = time.perf_counter()
        env.step(a)
        spent += time.perf_counter() - t0
        steps += N
    return {"env_steps_per_s": steps / spent, "envs": N, "what": "Particle2D.step (Particle2D.py:22-30)"}


def tfe(boards=256, steps_each=400):
    mod = ref_loader.tfe_module()
    rng = np.random.default_rng(2)
    env = mod.TFE(4, 4)
    env.reset(); env.step(0)                      # numba compile
    t0 = time.perf_counter()
    n = 0
    for b in range(boards):
        env = mod.TFE(4, 4)
        env.reset()
        for a in rng.integers(0, 4, steps_each):
            _, _, done, _, _ = env.step(int(a))
            n += 1
            if done:
                env.reset()
    spent = time.perf_counter() - t0
    return {"env_steps_per_s": n / spent, "boards": boards, "what": "TFE.step (TFE.py:152-189), one board per object, Python loop, one core"}


ONLY = set(sys.argv[1:])   # optional: names to (re)measure; the others are left out of the printed file
for name, fn in (("poker_hu_256", lambda: poker(2, 256, 16, min_s=10.0)), ("poker_hu_65536", lambda: poker(2, 65536, 16)), ("poker_hu_1048576", lambda: poker(2, 1 << 20, 16)),
                 ("poker_6max_65536", lambda: poker(6, 65536, 24)), ("poker_10max_65536", lambda: poker(10, 65536, 32)),
                 ("blackjack_4096", blackjack), ("particle2d_16M", particle), ("tfe_4x4_single_board_loop", tfe)):
    if ONLY and name not in ONLY:
        continue
    t0 = time.time()
    out["results"][name] = fn()
    out["results"][name]["seconds"] = round(time.time() - t0, 1)
    print(name, out["results"][name], file=sys.stderr, flush=True)
print(json.dumps(out, indent=1))
