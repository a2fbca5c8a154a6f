"""Per-step digests of whole env batches, used to chain the 1M-trajectory parity claim to the reference itself.

tests/golden/make_digests_1m.py runs the reference (imported in place from /root/reference) head to head with the
C oracle on >= 1,048,576 seeded trajectories per integer env and writes one 64-bit digest per output array per
step to tests/golden/digests_1m.json.  The CPU suite recomputes the same digests from the oracle alone, the GPU
suite from the CUDA path (tests/test_digests_1m.py, tests/test_gpu_million.py): three implementations, one file.

A digest is a sum over envs (mod 2^64) of a mixed hash of (global env id, the env's row), so it is independent of
how the batch was split into runs or GPUs -- the reference is stepped in 65,536-table runs, the engine in one batch
-- and sensitive to which env holds which row.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1


def _splitmix(i: int) -> int:
    z = (i * 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return (z ^ (z >> 31)) & _M64


def _as_u64_matrix(a) -> np.ndarray:
    a = np.asarray(a)
    if a.dtype == np.float32:
        a = np.ascontiguousarray(a).view(np.uint32)
    elif a.dtype == np.bool_:
        a = a.astype(np.uint8)
    a = a.reshape(a.shape[0], -1)
    if a.dtype.kind == "i":
        a = a.astype(np.int64).view(np.uint64)
    return a.astype(np.uint64, copy=False)


def digest(arrays, env_base: int = 0) -> int:
    """arrays: one array or a sequence of arrays with the same leading (env) dimension; float32 is hashed by bit
    pattern, integers by value (sign-extended to 64 bits)."""
    if isinstance(arrays, np.ndarray) or hasattr(arrays, "dtype"):
        arrays = [arrays]
    mats = [_as_u64_matrix(a) for a in arrays]
    n = mats[0].shape[0]
    row = np.zeros(n, np.uint64)
    col = 0
    with np.errstate(over="ignore"):
        for m in mats:
            assert m.shape[0] == n
            c = np.array([_splitmix(col + j + 1) | 1 for j in range(m.shape[1])], np.uint64)
            row += m @ c          # integer matmul wraps mod 2^64
            col += m.shape[1]
        g = np.arange(env_base, env_base + n, dtype=np.uint64)
        row += g * np.uint64(0x9E3779B97F4A7C15)
        row ^= row >> np.uint64(31)
        row *= np.uint64(0x7FB5D329728EA185)
        row ^= row >> np.uint64(27)
        row *= np.uint64(0x81DADEF4BC2DD44D)
        row ^= row >> np.uint64(33)
        return int(row.sum(dtype=np.uint64))


def add(a: int, b: int) -> int:
    return (a + b) & _M64


def hexd(v: int) -> str:
    return f"{v & _M64:016x}"


POKER_STATE = ("deck_positions", "pots", "stages", "button", "idx", "highest", "agg", "acted", "last_raise_size",
               "stacks", "current_round_bet", "total_invested", "status", "board", "hands")


def poker_state_digest(get, env_base: int) -> int:
    """get(name) -> array; the integer state of the reference's tensor attributes"""
    return digest([np.asarray(get(f)).astype(np.int64) for f in POKER_STATE], env_base)


# ---- the 1M-trajectory cases (shared by the generator and by both checking suites) ---------------------------
CASES = {
    # 8 runs x 65,536 tables x 2 consecutive hands (second hand: persistent stacks, rotation 1, moved button)
    "poker_hu": dict(P=2, M=2, runs=8, n_run=65536, hands=2, steps=16, seed=20260401, rot=1,
                     cfg=dict(starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)),
    "poker_6max": dict(P=6, M=6, runs=8, n_run=65536, hands=2, steps=24, seed=20260402, rot=1,
                       cfg=dict(starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)),
    # the reference's training default (PokerGPU.py:20): ten seats, a different table size every hand
    # (options["active_players"], :76-80), stacks rolled by 3 seats between hands; 8 runs x 32,768 tables x 4 hands
    "poker_10max": dict(P=10, M=10, runs=8, n_run=32768, hands=4, steps=32, seed=20260403, rot=3, aps=[10, 7, 4, 9],
                        cfg=dict(starting_bbs=20, max_bbs=1000, w1=.5, w2=.5, K=20, alpha=300)),
    # 16 runs x 65,536 envs x 1 episode of 10 steps
    "blackjack": dict(runs=16, n_run=65536, steps=10, seed=5),
    # 1,048,576 single-board reference envs x 10 steps
    "tfe_4x4": dict(n=1 << 20, steps=10, seed=2, side=4),
}


def case_actions(name: str, hand: int = 0) -> np.ndarray:
    """[steps, total envs] int64 action tape of a case (one seeded numpy stream per case and hand)."""
    c = CASES[name]
    rng = np.random.default_rng([c["seed"], 1000, hand])
    if name.startswith("poker"):
        return rng.integers(0, 13, (c["steps"], c["runs"] * c["n_run"])).astype(np.int64)
    if name == "blackjack":
        return rng.integers(0, 2, (c["steps"], c["runs"] * c["n_run"])).astype(np.int64)
    return rng.integers(0, 4, (c["steps"], c["n"])).astype(np.int64)


# ---- drivers shared by the CPU (oracle) and GPU (CUDA) checks --------------------------------------------------
def load_digests():
    import json
    from pathlib import Path

    return json.loads((Path(__file__).resolve().parent / "golden" / "digests_1m.json").read_text())


def check_poker_case(name, engine, reward_rtol=1e-5):
    """engine: reset(hand, active_players, rotation) -> obs array; step(actions) -> (obs, reward, done) arrays;
    get(field) -> array.  One batch of runs * n_run tables, global table ids 0 .. N-1.  Asserts every per-step digest
    of tests/golden/digests_1m.json (written while the reference itself was stepping the same trajectories)."""
    c, want = CASES[name], load_digests()[name]
    P = c["P"]
    for h in range(c["hands"]):
        acts = case_actions(name, h)
        ap = c["aps"][h] if "aps" in c else P
        obs = engine.reset(h, ap, (c["rot"] * h) % P)
        assert hexd(digest(obs)) == want["hands"][h]["reset_obs"], f"{name} hand {h}: reset observation digest"
        for t in range(c["steps"]):
            obs, rew, done = engine.step(acts[t])
            w, tag = want["hands"][h]["steps"][t], f"{name} hand {h} step {t}"
            assert hexd(digest(obs)) == w["obs"], tag + ": observation digest differs from the reference's"
            assert hexd(digest(np.asarray(done).astype(np.uint8))) == w["done"], tag + ": done digest"
            assert hexd(poker_state_digest(engine.get, 0)) == w["state"], tag + ": state digest"
            assert int(np.asarray(done).sum()) == w["done_count"], tag
            rs = float(np.asarray(rew).astype(np.float64).sum())
            assert abs(rs - w["reward_sum"]) <= reward_rtol * max(1.0, abs(w["reward_sum"])) + 1e-3 * abs(c["cfg"]["alpha"]), tag + " reward sum"
    return want["trajectories"]
