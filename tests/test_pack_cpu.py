"""CPU suite for pulse_poker_pack: reference-shaped tensors -> packed layout, on the host-built body.
Round trip: a packed table unpacked into the reference's tensors and packed again must give back the same words --
on every street, with seats sitting out, finished and settled hands included -- and an edit made through the
reference's attributes must survive the round trip and drive the next step exactly as the reference-shaped engine
would.  Tables the packed layout cannot express are counted, not silently bent."""
import numpy as np
import pytest

PACKED = ("tbl", "tbl2", "pstack", "pcrb", "pinv", "phole", "live_idx")


def _play(hr_table, P, AP, n, steps, seed):
    from tests.hostemu.emu import EmuPokerPacked

    env = EmuPokerPacked(hr_table, seed=seed, env_base=9, n_players=P, max_players=P, n_games=n, starting_bbs=25, max_bbs=60)
    env.reset(hand=0, active_players=AP, rotation=1)
    rng = np.random.default_rng(seed)
    for _ in range(steps):
        a = rng.integers(0, 13, n)
        a[rng.random(n) < 0.6] = 1
        env.step(a)
    return env


@pytest.mark.parametrize("P,AP,steps", [(2, 2, 0), (2, 2, 3), (2, 2, 9), (6, 6, 7), (6, 4, 12), (10, 7, 15), (10, 10, 40)])
def test_unpack_then_pack_is_the_identity(P, AP, steps, hr_table):
    n = 96
    env = _play(hr_table, P, AP, n, steps, 31 * P + steps)
    want = {f: getattr(env, f).copy() for f in PACKED}
    want_ev = env.pev[:AP].copy()
    for f in PACKED:                      # scramble: everything must come back from the mirror
        getattr(env, f)[...] = 0x5A
    env.pev[...] = 0
    assert env.pack_from_mirror() == 0
    for f in PACKED:
        np.testing.assert_array_equal(getattr(env, f), want[f], err_msg=f)
    np.testing.assert_array_equal(env.pev[:AP], want_ev, err_msg="hand-value words")
    if steps in (7, 9, 12, 15):
        assert len(np.unique(env.stages)) >= 3
    if steps == 40:
        assert np.all(env.stages == 5)      # every hand settled: the board comes back from the five visible cards


def test_edit_through_the_mirror_then_pack_then_step(hr_table):
    """the white-box pattern of the reference's tests (env.stacks[0] = ..., env.pots[...] = ...): edit, pack, step on
    the packed layout == the same edit stepped on the reference-shaped layout"""
    from tests.hostemu.emu import EmuPoker

    P, n = 6, 64
    env = _play(hr_table, P, P, n, 5, 77)
    ref = EmuPoker(hr_table, n_players=P, max_players=P, n_games=n, starting_bbs=25, max_bbs=60)
    for f in ("decks", "board", "hands", "is_done", "equity_dirty", "_eq") + tuple(__import__("tests.hostemu.emu", fromlist=["N_FIELDS"]).N_FIELDS) + \
            tuple(__import__("tests.hostemu.emu", fromlist=["NP_FIELDS"]).NP_FIELDS):
        getattr(ref, f)[...] = getattr(env, f)
    ref.active_players = env.active_players
    ref._first = False
    for e in (env, ref):
        e.stacks[::3, 1] += 17
        e.pots[::5] += 4
        e.last_raise_size[1::4] = 3
    assert env.pack_from_mirror() == 0
    rng = np.random.default_rng(5)
    for t in range(12):
        a = rng.integers(0, 13, n)
        o1, r1, d1 = env.step(a)
        o2, r2, d2 = ref.step(a)
        np.testing.assert_array_equal(o1, o2, err_msg=f"step {t} obs")
        np.testing.assert_array_equal(d1, d2, err_msg=f"step {t} done")
        np.testing.assert_allclose(r1, r2, rtol=2e-6, atol=1e-4, err_msg=f"step {t} reward")
        for f in ("stacks", "pots", "stages", "status", "current_round_bet", "total_invested", "board"):
            np.testing.assert_array_equal(getattr(env, f), getattr(ref, f), err_msg=f"step {t} {f}")


def test_unrepresentable_tables_are_counted(hr_table):
    env = _play(hr_table, 2, 2, 32, 4, 3)
    assert env.pack_from_mirror() == 0
    env.deck_positions[3] += 1            # a deck position the street does not imply
    env.board[5, 4] = 7 if env.stages[5] < 3 else env.board[5, 4]
    env.hands[9, 0, 0] = 0                # not a card
    env.status[11, 1] = 7
    env.stages[13] = 6
    bad = env.pack_from_mirror()
    expect = 4 + (1 if env.stages[5] < 3 else 0)
    assert bad == expect, (bad, expect)
