"""Pins of the regenerated HandRanks table: size, constants the reference
hard-codes (PokerGPU.py:13-18), category counts, and agreement of table walk, computed evaluator and
brute force."""
import ctypes as C
import itertools

import numpy as np

from oracle import pyoracle
from pulselib_b200 import _lib


def test_table_size_and_range(hr_table):
    assert hr_table.size == 32487834 and hr_table.dtype == np.int32
    assert hr_table.max() < hr_table.size and hr_table.min() >= 0


def test_all_seven_card_hands(hr_table, oracle_lib):
    cc = (C.c_longlong * 10)()
    mn, mx = C.c_int(), C.c_int()
    oracle_lib.oracle_check_table7(pyoracle.p(hr_table), 1, 0, cc, C.byref(mn), C.byref(mx))
    assert list(cc)[1:] == [23294460, 58627800, 31433400, 6461620, 6180020, 4047644, 3473184, 224848, 41584]
    assert (mn.value, mx.value) == (4145, 36874)  # MIN_EQUITY_RANK, MAX_EQUITY_RANK
    bad = oracle_lib.oracle_check_table7(pyoracle.p(hr_table), 211, 1, cc, C.byref(mn), C.byref(mx))
    assert bad == 0


def test_five_card_hands_and_flop_lookup_range(hr_table):
    cards = np.array(list(itertools.combinations(range(1, 53), 5)), dtype=np.int64)
    p = np.full(len(cards), 53, dtype=np.int64)
    for i in range(5):
        p = hr_table[p + cards[:, i]]
    v5 = hr_table[p]
    assert np.bincount(v5 >> 12).tolist()[1:] == [1302540, 1098240, 123552, 54912, 10200, 5108, 3744, 624, 40]
    flop = hr_table[v5]  # the reference's HR[HR[p5]] (PokerGPU.py:521)
    valid = flop > 53
    assert (flop[valid].min(), flop[valid].max()) == (74359, 823779)  # MIN/MAX_FLOP_EQUITY


def test_six_card_minimum(hr_table):
    rng = np.random.default_rng(0)
    worst = np.array([[4, 7, 10, 13, 18, 23]], dtype=np.int64)  # 2s 3h 4d 5c 6d.. replaced below
    # lowest possible 6-card hand is 8-7-5-4-3 high with a deuce: value 4096 + 13
    cards = np.array([[1, 6, 11, 16, 22, 27]], dtype=np.int64)  # 2c 3d 4h 5s 7d 8h
    p = np.full(1, 53, dtype=np.int64)
    for i in range(6):
        p = hr_table[p + cards[:, i]]
    assert hr_table[p][0] == 4109  # MIN_TURN_RIVER_EQUITY
    hands = np.array([rng.choice(52, 6, replace=False) + 1 for _ in range(200000)], dtype=np.int64)
    p = np.full(len(hands), 53, dtype=np.int64)
    for i in range(6):
        p = hr_table[p + hands[:, i]]
    assert hr_table[p].min() >= 4109


def test_computed_evaluator_matches_table_and_bruteforce(hr_table, oracle_lib):
    lib = _lib.load()
    rng = np.random.default_rng(1)
    for n in (5, 6, 7):
        for _ in range(4000):
            c = (rng.choice(52, n, replace=False) + 1).astype(np.int32)
            walk = lib.pulse_handranks_walk(pyoracle.p(hr_table), pyoracle.p(c), n)
            assert lib.pulse_eval_cards(pyoracle.p(c), n) == walk
            assert oracle_lib.oracle_eval_bruteforce(pyoracle.p(c), n) == walk


def test_philox_known_answers(oracle_lib):
    """Random123 kat_vectors for philox4x32-10."""
    lib = _lib.load()
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        out = (C.c_uint32 * 4)()
        oracle_lib.oracle_philox4x32_10((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        assert tuple(out) == want
        # product-side tape: counter = {env lo, env hi, counter, stream}, key = seed
        env = ctr[0] | (ctr[1] << 32)
        seed = key[0] | (key[1] << 32)
        out2 = (C.c_uint32 * 4)()
        assert lib.pulse_tape_block(C.c_uint64(seed), C.c_uint64(env), C.c_uint32(ctr[3]), C.c_uint32(ctr[2]), out2) == 0
        assert tuple(out2) == want


def test_computed_evaluator_equals_table_walk_on_every_hand(hr_table):
    """csrc/poker_eval.cuh (three small look-up tables + integer arithmetic) against the reference's own evaluation --
    the HandRanks walk (PokerGPU.py:437-444, :497-500) -- on EVERY 5-, 6- and 7-card hand: 156,742,040 hands."""
    from tests.hostemu import emu

    f = emu.lib().emu_eval_vs_table
    f.restype = C.c_longlong
    for n, hands in ((5, 2598960), (6, 20358520), (7, 133784560)):
        seen = C.c_longlong()
        assert f(emu.p(hr_table), n, 1, C.byref(seen)) == 0, f"{n}-card hands"
        assert seen.value == hands
