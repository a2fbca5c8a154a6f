"""CPU suite for the enumerated hand strength: the kernel body (compiled for the host) against the
oracle's brute-force enumeration, and -- the one point where the reference defines the quantity -- against the
reference's river equity (PokerGPU.py:477-481, held by the oracle that the reference's own tests pin)."""
import ctypes as C

import numpy as np

from oracle import pyoracle
from tests.hostemu import emu


def _body(stage, h0, h1, board):
    f = emu.lib().emu_enumerated_strength
    f.restype = C.c_float
    fb = 0
    for i, c in enumerate(board):
        fb |= int(c) << (6 * i)
    return f(C.c_int(stage), C.c_int(int(h0)), C.c_int(int(h1)), C.c_uint32(fb))


def test_kernel_body_equals_bruteforce_enumeration():
    rng = np.random.default_rng(4)
    n = 40
    cards = np.stack([rng.choice(52, 7, replace=False) + 1 for _ in range(n)]).astype(np.int32)
    for stage in (0, 1, 2, 3, 4, 5):
        stages = np.full(n, stage, np.int32)
        want = pyoracle.enumerated_strength(stages, cards[:, None, :2], cards[:, 2:], 1)[:, 0]
        got = np.array([_body(stage, c[0], c[1], c[2:]) for c in cards], np.float32)
        np.testing.assert_array_equal(got, want, err_msg=f"stage {stage}")
        if stage in (0, 4, 5):
            assert np.all(got == 0.5)
        else:
            assert np.all((got >= 0) & (got <= 1)) and got.std() > 0


def test_river_strength_is_the_reference_river_equity(hr_table):
    kw = dict(n_players=4, max_players=4, n_games=300, starting_bbs=50, max_bbs=200, w1=.5, w2=.3, K=100, alpha=50)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    orc.tape_decks(9, 0, 0)
    orc.reset(None)
    orc.stages[:] = 3
    orc.board[:] = orc.decks[:, 40:45]          # hole cards come from the front of the deck: these five are unseen
    orc.equity_dirty[:] = 1
    orc.call("calculate_equities")
    got = pyoracle.enumerated_strength(orc.stages, orc.hands, orc.board, 4)
    np.testing.assert_array_equal(got, orc.equities[:, :4])


def test_known_values():
    # royal flush on the river: the top of the scale; the nuts cannot get worse on later streets either
    royal = [52, 48, 44, 40, 36, 1, 2]          # A K Q J T of one suit (card = rank * 4 + suit + 1), 2c 2d
    assert _body(3, royal[0], royal[1], royal[2:]) > 0.99
    assert _body(1, royal[0], royal[1], royal[2:]) > 0.99
    # quads on the flop only improve or hold: the mean over run-outs is at least the river value of any one of them
    flop = _body(1, 49, 50, [51, 52, 1, 0, 0])
    assert flop >= 0.85
