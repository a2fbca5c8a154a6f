"""CPU suite: the compact step record of the host-buffer path (pulselib_b200/csrc/poker_wire.cuh).  The device packs
one record per live table, host threads expand it into the reference's float32 observation row; here the host-compiled
kernel bodies pack and the PRODUCT library's pure-host expander (pulse_poker_wire_expand_host) unpacks: every row must
come back bit-identical to the row the float32 path writes (pk_obs), with reward and done."""
import numpy as np
import pytest

from tests.hostemu.emu import EmuPokerPacked


@pytest.mark.parametrize("P,M,ap", [(2, 2, 2), (4, 4, 3), (6, 6, 6), (6, 6, 2), (10, 10, 10), (9, 10, 7)])
def test_wire_rows_round_trip(P, M, ap, hr_table):
    from pulselib_b200 import host_env

    N = 1500
    env = EmuPokerPacked(hr_table, n_players=P, max_players=M, n_games=N, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                         K=100, alpha=50, seed=9)
    rng = np.random.default_rng(P * 7 + M)
    for hand in range(3):
        env.reset(None, active_players=ap, rotation=hand % P, hand=hand)
        env.rewards[...] = 0
        for t in range(40):
            rows = env.wire_rows()
            for vec in (True, False):      # the host env's vectorised expander and the portable definition
                obs, rew, done = host_env.expand_rows(rows, M, vectorised=vec)
                np.testing.assert_array_equal(obs, env.obs, err_msg=f"hand {hand} step {t} obs (vectorised={vec})")
                np.testing.assert_array_equal(rew.view(np.uint32), env.rewards.view(np.uint32))
                np.testing.assert_array_equal(done, env.is_done)
            env.step(rng.integers(0, 13, N))


def test_expander_isa_is_reported():
    from pulselib_b200 import host_env

    assert host_env.expand_isa() in ("avx2", "portable")
