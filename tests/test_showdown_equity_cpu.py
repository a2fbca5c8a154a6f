"""CPU suite for the showdown equity (second half): the kernel body compiled for the host
(csrc/poker_packed_body.cuh::packed_showdown_equity_part) against the oracle's restatement of the definition
(oracle/oracle_core.c::oracle_showdown_equity), plus known answers that need neither.

No reference code computes this quantity (the reference's notes ask for it, rl/depth_notes.txt:316-321), so the pins
are: the oracle's showdown comparison is the reference's own table walk over HandRanks; textbook match-ups; the full
enumeration of all 1,712,304 heads-up preflop boards."""
import numpy as np
import pytest

from oracle import pyoracle


def _card(rank, suit):
    """rank 2..14, suit 0..3 -> the table's card id (1..52, rank-major)"""
    return 4 * (rank - 2) + suit + 1


def _play(hr_table, P, n, steps, seed, auto_reset=False):
    from tests.hostemu.emu import EmuPokerPacked

    env = EmuPokerPacked(hr_table, auto_reset=auto_reset, seed=seed, env_base=500, n_players=P, max_players=P, n_games=n)
    env.reset(hand=0)
    rng = np.random.default_rng(seed)
    for _ in range(steps):
        a = rng.integers(1, 13, n)                 # few folds: keep several contenders around
        a[rng.random(n) < 0.08] = 0
        env.step(a)
    return env


@pytest.mark.parametrize("P,steps", [(2, 0), (2, 3), (2, 7), (6, 0), (6, 9), (6, 20), (10, 14)])
def test_host_body_matches_oracle_on_played_tables(P, steps, hr_table):
    n = 48
    env = _play(hr_table, P, n, steps, 100 + P + steps)
    got = env.showdown_equity(samples=64, counter=3, lanes=1)
    want = pyoracle.showdown_equity(env.hands, env.status, env.board, P, samples=64, seed=env.seed, env_base=env.env_base,
                                    counter=3, hand_ranks=hr_table)
    np.testing.assert_array_equal(got, want)
    # the boards of a table split over 32 lanes as in the kernel, and over an odd count: same integers, same floats
    np.testing.assert_array_equal(env.showdown_equity(samples=64, counter=3, lanes=32), got)
    np.testing.assert_array_equal(env.showdown_equity(samples=64, counter=3, lanes=7), got)
    # rows are a partition of the pot; folded seats hold nothing
    folded = (env.status[:, :P] == 1) | (env.status[:, :P] == 3)
    assert np.all(got[folded] == 0.0)
    np.testing.assert_allclose(got.sum(1), (~folded).any(1).astype(np.float32), atol=1e-6)
    if (P, steps) in ((2, 3), (6, 9), (10, 14)):
        assert len(np.unique(env.stages)) > 2, "the batch should span several streets"


def test_oracle_table_walk_and_bruteforce_evaluators_agree(hr_table):
    env = _play(hr_table, 6, 6, 9, 5)
    kw = dict(samples=8, seed=9, env_base=1, counter=2)
    a = pyoracle.showdown_equity(env.hands, env.status, env.board, 6, hand_ranks=hr_table, **kw)
    b = pyoracle.showdown_equity(env.hands, env.status, env.board, 6, hand_ranks=None, **kw)
    np.testing.assert_array_equal(a, b)


def test_known_answers(hr_table):
    S, H, D, Cl = 0, 1, 2, 3
    hands = np.array([[[_card(14, S), _card(14, H)], [_card(13, S), _card(13, H)]]], np.int32)   # AA v KK
    status = np.zeros((1, 2), np.int32)
    # river: the kings hit their set
    board = np.array([[_card(13, D), _card(7, Cl), _card(2, D), _card(9, S), _card(4, H)]], np.int32)
    np.testing.assert_array_equal(pyoracle.showdown_equity(hands, status, board, 2, hand_ranks=hr_table), [[0.0, 1.0]])
    # turn, aces ahead: kings need one of two kings out of 44 unseen cards
    board[0, 4] = -1
    board[0, 0] = _card(3, D)
    e = pyoracle.showdown_equity(hands, status, board, 2, hand_ranks=hr_table)
    np.testing.assert_allclose(e, [[42 / 44, 2 / 44]], rtol=1e-7)
    # a board that plays for both: royal flush on the table -> a chop on every street that shows it
    board = np.array([[_card(10, D), _card(11, D), _card(12, D), _card(13, D), _card(14, D)]], np.int32)
    np.testing.assert_array_equal(pyoracle.showdown_equity(hands, status, board, 2, hand_ranks=hr_table), [[0.5, 0.5]])
    # everybody else folded
    status6 = np.array([[1, 1, 0, 1, 1, 1]], np.int32)
    hands6 = np.arange(1, 13, dtype=np.int32).reshape(1, 6, 2)
    none = np.full((1, 5), -1, np.int32)
    np.testing.assert_array_equal(pyoracle.showdown_equity(hands6, status6, none, 6, samples=4), [[0, 0, 1, 0, 0, 0]])


def test_sampled_preflop_converges_to_full_enumeration(hr_table):
    """AA v KK preflop: all 1,712,304 boards enumerated (about 0.82 / 0.18), against 65,536 sampled boards from the
    kernel body (standard error 0.0015 per seat) and against two half-size runs at different tape counters."""
    from tests.hostemu.emu import EmuPokerPacked

    S, H = 0, 1
    hands = np.array([[_card(14, S), _card(14, H)], [_card(13, S), _card(13, H)]], np.int32)
    exact = pyoracle.showdown_equity_preflop_exact(hands, np.zeros(2, np.int32), hr_table)
    assert abs(exact.sum() - 1.0) < 1e-12 and 0.81 < exact[0] < 0.83
    deck = np.concatenate([hands.reshape(-1), np.setdiff1d(np.arange(1, 53), hands.reshape(-1))]).astype(np.int32)
    env = EmuPokerPacked(hr_table, seed=3, env_base=0, n_players=2, max_players=2, n_games=1)
    env.reset(decks=deck[None, :], hand=0)
    # the four cards on top of the deck are A A K K; however the deal hands them out, the yardstick is the full
    # enumeration for the hands as dealt
    assert sorted(env.hands[0].reshape(-1).tolist()) == sorted(hands.reshape(-1).tolist())
    exact = pyoracle.showdown_equity_preflop_exact(env.hands[0], np.zeros(2, np.int32), hr_table)
    got = env.showdown_equity(samples=65536, counter=0, lanes=32)[0]
    np.testing.assert_allclose(got, exact, atol=0.008)          # five standard errors
    want = pyoracle.showdown_equity(env.hands, env.status, env.board, 2, samples=65536, seed=3, env_base=0, counter=0,
                                    hand_ranks=hr_table)[0]
    np.testing.assert_array_equal(got, want)
    a = env.showdown_equity(samples=32768, counter=1, lanes=32)[0]
    b = env.showdown_equity(samples=32768, counter=2, lanes=32)[0]
    assert not np.array_equal(a, b), "different tape counters must give different samples"
    np.testing.assert_allclose(a, exact, atol=0.011)
    np.testing.assert_allclose(b, exact, atol=0.011)
