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


@pytest.mark.parametrize("P,M,N", [(2, 2, 1500), (2, 2, 977), (6, 6, 333), (10, 10, 65)])
def test_host_env_bookkeeping_through_whole_hands(P, M, N, hr_table):
    """The host env's expander exactly as the env drives it (live mask, compacted records, reward zeroing on the step
    after a table finishes), without a GPU: the host-compiled kernel bodies step the tables and pack the records of the
    tables that were live when the step began; pulse_poker_wire_expand_step_host must leave the caller's arrays equal to
    what the float32 path returns, for batches that are not multiples of 32 -- with guard words behind every array."""
    import ctypes as C

    from pulselib_b200 import _lib, host_env

    lib = _lib.load()
    env = EmuPokerPacked(hr_table, n_players=P, max_players=M, n_games=N, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3,
                         K=100, alpha=50, seed=21)
    W, G = 13 + 3 * (M - 1), (N + 31) // 32
    pad = 64
    obs = host_env._aligned((N + pad, W), np.float32); obs[...] = -7.0
    rew = np.full(N + pad, -7.0, np.float32); done = np.full(N + pad, 9, np.uint8)
    live, prev = np.zeros(G + 1, np.uint32), np.zeros(G + 1, np.uint32)
    p = lambda a: C.c_void_p(a.ctypes.data)
    rng = np.random.default_rng(N)

    def expand(rows, fresh):
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        rc = lib.pulse_poker_wire_expand_step_host(p(rows), p(live), p(prev), C.c_longlong(N), C.c_int(M), C.c_int(fresh),
                                                   p(obs), p(rew), p(done))
        assert rc == 0
        assert np.all(obs[N:] == -7.0) and np.all(rew[N:] == -7.0) and np.all(done[N:] == 9), "wrote past the batch"

    for hand in range(3):
        env.reset(None, hand=hand)
        env.rewards[...] = 0
        expand(env.wire_rows(), 1)
        np.testing.assert_array_equal(obs[:N], env.obs)
        assert not done[:N].any() and not rew[:N].any()
        for t in range(30):
            was_live = ~env.is_done.astype(bool)
            bits = np.zeros(G * 32, bool); bits[:N] = was_live
            np.testing.assert_array_equal(np.packbits(bits.reshape(G, 32), axis=1, bitorder="little").view(np.uint32).ravel(), live[:G])
            o, r, d = env.step(rng.integers(0, 13, N))
            expand(env.wire_rows()[was_live], 0)
            np.testing.assert_array_equal(obs[:N], o, err_msg=f"hand {hand} step {t} obs")
            np.testing.assert_array_equal(rew[:N].view(np.uint32), r.view(np.uint32), err_msg=f"hand {hand} step {t} reward")
            np.testing.assert_array_equal(done[:N].astype(bool), d)


# ---- 2048 host path: packed boards -> int32 tile values ------------------------------------------------------------
@pytest.mark.parametrize("side", [4, 3, 2])
def test_tfe_board_expander_matches_numpy(side):
    """csrc/host_expand.cpp::pulse_tfe_expand_range (AVX2 for aligned 4 x 4 boards, portable otherwise) against the
    definition: cell (r, c) holds exponent e at bits 16 r + 4 c, tile value = 2^e, 0 for e = 0."""
    from pulselib_b200.host_env import expand_boards

    rng = np.random.default_rng(side)
    n = 4099
    e = rng.integers(0, 16, (n, 4, 4)).astype(np.uint64)
    e[:, side:, :] = 0
    e[:, :, side:] = 0
    e[0] = 0
    e[1, :side, :side] = 15
    b = np.zeros(n, np.uint64)
    for r in range(4):
        for c in range(4):
            b |= e[:, r, c] << np.uint64(16 * r + 4 * c)
    ex = e[:, :side, :side].astype(np.int64)
    want = np.where(ex > 0, 1 << ex, 0).astype(np.int32)
    np.testing.assert_array_equal(expand_boards(b, side), want)
