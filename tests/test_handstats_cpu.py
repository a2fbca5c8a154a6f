"""CPU suite for the hand-level statistics row: the numpy restatement of the reference's
per-step bookkeeping, folded into the 67-double vector and summarised by the product's `summarize`, must give the
numbers the reference's own utils/performance.py produced for the same trace (tests/golden/poker_handstats.npz,
made by tests/golden/make_handstats_golden.py)."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu

Z = np.load(pu.GOLDEN / "poker_handstats.npz")


def replay(tag, update):
    N, P, AP, q_seat, hands, steps = (int(v) for v in Z[f"{tag}/meta"])
    k = 0
    for h in range(hands):
        terminated = np.zeros(N, bool)
        for t in range(steps):
            update(Z[f"{tag}/stacks"][k], Z[f"{tag}/initial"][k], Z[f"{tag}/stages"][k], Z[f"{tag}/button"][k],
                   Z[f"{tag}/done"][k], terminated, q_seat, AP, first=(t == 0))
            k += 1
    return AP


def check_summary(tag, vec, AP):
    from pulselib_b200.hand_stats import summarize

    s = summarize(vec, AP)
    mean, win_rate, hands, bb100, lcb, seat_bal = Z[f"{tag}/summary"]
    assert s["hands_completed"] == int(hands)
    # the reference reduces in float32
    np.testing.assert_allclose([s["mean_bb_delta"], s["hand_win_rate"], s["field_bb_per_100"], s["lcb95_bb_per_100"],
                                s["seat_balanced_bb_per_100"]], [mean, win_rate, bb100, lcb, seat_bal], rtol=2e-4, atol=2e-4)
    np.testing.assert_allclose([s["by_stage"][n]["win_percentage"] for n in ("preflop", "flop", "turn", "river", "showdown")],
                               Z[f"{tag}/street_win_pct"], rtol=1e-5, atol=1e-7)
    for p in range(AP):
        m = s["by_position"].get(f"position_{p}", {"hands": 0, "wins": 0})
        assert m["hands"] == int(Z[f"{tag}/position_hands"][p]) and m["wins"] == int(Z[f"{tag}/position_wins"][p])


@pytest.mark.parametrize("tag", ["six", "hu"])
def test_oracle_hand_stats_match_reference_metrics(tag):
    vec = np.zeros(67)

    def update(stacks, initial, stages, button, done, terminated, q_seat, AP, first):
        pyoracle.hand_stats_update(vec, stacks[:, q_seat], initial, stages, button, done, terminated, q_seat, AP)

    AP = replay(tag, update)
    check_summary(tag, vec, AP)
