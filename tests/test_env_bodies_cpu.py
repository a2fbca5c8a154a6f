"""CPU: the per-env bodies of the CUDA kernels (csrc/*_body.cuh, compiled for the host by tests/hostemu)
against the reference's golden trajectories and against the C oracle on long seeded rollouts.  This is the
same code the GPU runs; the `-m gpu` tests repeat the comparison through the C ABI on the device."""
from pathlib import Path

import numpy as np
import pytest

from oracle import pyoracle
from tests.hostemu.emu import EmuBlackjack, EmuTFE, EmuTetris
from tests.test_env_oracles import BJ_FIELDS

GOLD = Path(__file__).resolve().parent / "golden"


def test_blackjack_body_matches_reference_golden():
    g = np.load(GOLD / "blackjack.npz")
    N, episodes, steps, seed, env_base = (int(x) for x in g["meta"])
    env = EmuBlackjack(N, seed=seed, env_base=env_base)
    for e in range(episodes):
        obs = env.reset(episode=e)                       # decks from the device-side tape
        np.testing.assert_array_equal(env.unpack()["decks"], g["decks"][e])
        np.testing.assert_array_equal(obs, g["reset_obs"][e])
        for t in range(steps):
            obs, rew, term = env.step(g["actions"][e, t])
            np.testing.assert_array_equal(obs, g["obs"][e, t])
            np.testing.assert_array_equal(rew, g["rewards"][e, t])
            np.testing.assert_array_equal(term, g["terminated"][e, t])
            u = env.unpack()
            for k in BJ_FIELDS:
                if k in u:
                    np.testing.assert_array_equal(u[k].astype(g[k].dtype), g[k][e, t], err_msg=f"{k} e{e} t{t}")


def test_blackjack_body_matches_oracle_random_and_injected():
    N = 20000
    rng = np.random.default_rng(11)
    env, orc = EmuBlackjack(N, seed=5, env_base=123), pyoracle.OracleBlackjack(N)
    for e in range(4):
        if e % 2:
            decks = np.argsort(rng.random((N, 52)), axis=1).astype(np.int32)
            decks[:50] = 0                                  # degenerate all-aces decks (not permutations)
            np.testing.assert_array_equal(env.reset(decks=decks), orc.reset(decks))
        else:
            orc.tape_decks(5, 123, e)
            np.testing.assert_array_equal(env.reset(episode=e), orc.reset())
        for t in range(14):
            a = rng.integers(0, 3 if t % 3 == 0 else 2, N)  # action 2 is a no-op in the reference
            if t < 12:
                a[:2000] = 0                                # hit until bust
            o1, r1, t1 = env.step(a)
            o2, r2, t2 = orc.step(a)
            np.testing.assert_array_equal(o1, o2)
            np.testing.assert_array_equal(r1, r2)
            np.testing.assert_array_equal(t1, t2)
        u = env.unpack()
        for k in ("players_cards", "dealer_cards", "deck_positions", "players_card_idx", "dealer_card_idx",
                  "player_card_sums", "dealer_card_sums", "has_ace", "dealer_has_ace", "dealer_upcard"):
            np.testing.assert_array_equal(u[k], getattr(orc, k), err_msg=k)


@pytest.mark.parametrize("name", ["tfe_4x4", "tfe_3x3"])
def test_tfe_body_matches_reference_golden(name):
    g = np.load(GOLD / f"{name}.npz")
    side, N, steps, seed, env_base = (int(x) for x in g["meta"])
    env = EmuTFE(N, side, seed=seed, env_base=env_base)
    np.testing.assert_array_equal(env.reset(counter=0), g["reset_boards"])          # chance from the tape
    for t in range(steps):
        boards, rew, done = env.step(g["actions"][t], counter=t)
        np.testing.assert_array_equal(boards, g["boards"][t], err_msg=f"t{t}")
        np.testing.assert_array_equal(rew, g["rewards"][t])
        np.testing.assert_array_equal(done, g["done"][t].astype(bool))
        np.testing.assert_array_equal(env.scores, g["scores"][t])


@pytest.mark.parametrize("side", [2, 3, 4])
def test_tfe_body_matches_oracle_autoreset(side):
    N, steps = 3000, 600
    rng = np.random.default_rng(side)
    env, orc = EmuTFE(N, side, seed=9, env_base=77), pyoracle.OracleTFE(N, side)
    np.testing.assert_array_equal(env.reset(counter=0), orc.reset(orc.tape(9, 77, pyoracle.STREAM_TFE_RESET, 0, 4)))
    resets = 0
    for t in range(1, steps):
        a = rng.integers(0, 4, N)
        resets += int(orc.done.sum())
        b2, r2, d2 = orc.step(a, orc.tape(9, 77, pyoracle.STREAM_TFE_SPAWN, t, 2), auto_reset=True,
                              reset_chance4=orc.tape(9, 77, pyoracle.STREAM_TFE_RESET, t, 4))
        b1, r1, d1 = env.step(a, auto_reset=True, counter=t)
        np.testing.assert_array_equal(b1, b2, err_msg=f"t{t}")
        np.testing.assert_array_equal(r1, r2)
        np.testing.assert_array_equal(d1, d2)
        np.testing.assert_array_equal(env.scores, orc.scores)
    assert resets > N // 2


def test_tfe_row_table_and_slots():
    """The staged row-move table (built at hashed slots, read through TfeTableRows) against the computed squash on
    all 65,536 rows; the board-wide slot hash against the per-row one; the slot map is a bijection."""
    import ctypes as C

    from tests.hostemu import emu

    L = emu.lib()
    assert L.emu_tfe_table_check() == 0
    rng = np.random.default_rng(3)
    boards = rng.integers(0, 2**64, 1 << 16, dtype=np.uint64)
    L.emu_tfe_slot_check.argtypes = [C.c_void_p, C.c_int]
    assert L.emu_tfe_slot_check(boards.ctypes.data_as(C.c_void_p), C.c_int(len(boards))) == 0


def test_tfe_body_large_tiles():
    """Hand-built boards with big tiles (random play never gets there): every move, both engines."""
    rng = np.random.default_rng(0)
    N = 4096
    exps = rng.integers(0, 15, (N, 4, 4))
    exps[rng.random((N, 4, 4)) < 0.3] = 0
    vals = np.where(exps > 0, 1 << exps, 0).astype(np.int32)
    for a in range(4):
        env, orc = EmuTFE(N, 4), pyoracle.OracleTFE(N, 4)
        env.set_boards(vals)
        orc.boards[...] = vals
        ch = rng.integers(0, 2**32, (N, 2), dtype=np.uint64).astype(np.uint32)
        b1, r1, d1 = env.step(np.full(N, a), ch)
        b2, r2, d2 = orc.step(np.full(N, a), ch)
        np.testing.assert_array_equal(b1, b2)
        np.testing.assert_array_equal(r1, r2)
        np.testing.assert_array_equal(d1, d2)
        np.testing.assert_array_equal(env.scores, orc.scores)


def test_tetris_body_matches_spec_oracle():
    N, steps = 4000, 400
    rng = np.random.default_rng(4)
    env, orc = EmuTetris(N, seed=4, env_base=31), pyoracle.OracleTetris(N)
    first = orc.tape_pieces(4, 31, 0, 1)
    orc.reset(first)
    env.reset(first)
    games = 0
    for t in range(1, steps):
        a = rng.integers(0, 40, N)
        if t % 7 == 0:
            a[:100] = rng.integers(-5, 60, 100)             # out-of-range actions are clamped
        rows2, r2, d2 = orc.step(a, orc.tape_pieces(4, 31, t, 0), auto_reset=True, reset_piece=orc.tape_pieces(4, 31, t, 1))
        rows1, r1, d1 = env.step(a, auto_reset=True, counter=t)  # pieces from the device-side tape
        np.testing.assert_array_equal(rows1, rows2, err_msg=f"t{t}")
        np.testing.assert_array_equal(r1, r2)
        np.testing.assert_array_equal(d1, d2)
        np.testing.assert_array_equal(env.piece, orc.piece)
        np.testing.assert_array_equal(env.lines, orc.lines)
        games += int(d2.sum())
    assert games > N


def test_tetris_body_line_clears_with_guided_play():
    """Random play almost never clears a line; flat I and O pieces laid side by side do, every few steps."""
    N = 64
    env, orc = EmuTetris(N), pyoracle.OracleTetris(N)
    first = np.zeros(N, np.uint8)
    env.reset(first)
    orc.reset(first)
    total = 0
    seq = [(0, 0), (0, 4), (1, 8)]                          # (piece, column): I at 0, I at 4, O at 8 -> rows fill up
    for t in range(60):
        piece, col = seq[t % 3]
        nxt = np.full(N, seq[(t + 1) % 3][0], np.uint8)
        a = np.full(N, col)
        rows1, r1, d1 = env.step(a, nxt, reset_piece=nxt)
        rows2, r2, d2 = orc.step(a, nxt)
        np.testing.assert_array_equal(rows1, rows2)
        np.testing.assert_array_equal(r1, r2)
        total += int(r2[0])
    assert total == 20 and not d2.any()
