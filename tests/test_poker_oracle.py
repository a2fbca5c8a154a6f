"""CPU suite: the oracle against the reference's golden vectors; the host-compiled kernel bodies
against the oracle; HandRanks pins."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu
from tests.hostemu.emu import EmuPoker, EmuPokerPacked

GOLDEN_CASES = ["poker_hu", "poker_6max", "poker_10max_varap"]


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_matches_reference_golden(name, hr_table):
    pu.replay_golden(name, pyoracle.OraclePoker(hr_table, **pu.golden_kwargs(name)))


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_kernel_body_hostbuild_matches_reference_golden(name, hr_table):
    pu.replay_golden(name, EmuPoker(hr_table, **pu.golden_kwargs(name)))


@pytest.mark.parametrize("P,M,N,vary,rot", [(2, 2, 3000, False, 0), (6, 6, 1500, False, 1), (10, 10, 800, True, 3),
                                              (3, 10, 800, True, 2), (4, 4, 800, False, 1), (7, 9, 600, True, 5)])
def test_kernel_body_hostbuild_matches_oracle_random(P, M, N, vary, rot, hr_table):
    kw = dict(n_players=P, max_players=M, n_games=N, starting_bbs=50, max_bbs=200, w1=.5, w2=.3, K=100, alpha=50)
    rng = np.random.default_rng(P * 100 + M)
    pu.random_rollout_compare(pyoracle.OraclePoker(hr_table, **kw), EmuPoker(hr_table, **kw), rng, P, N, 5, 45,
                              vary_ap=vary, rot=rot)


def test_tape_decks_match(hr_table):
    kw = dict(n_players=2, max_players=2, n_games=500)
    o, e = pyoracle.OraclePoker(hr_table, **kw), EmuPoker(hr_table, **kw)
    for hand in range(3):
        o.tape_decks(20260401, 12345, hand)
        e.reset(None, seed=20260401, env_base=12345, hand=hand)
        np.testing.assert_array_equal(o.decks, e.decks)
        assert np.all(np.sort(e.decks, axis=1) == np.arange(1, 53))


def test_reference_live_if_available(hr_table):
    """In the build container the real reference is importable: run it head to head with the oracle."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("/root/reference not present")
    import torch

    kw = dict(n_players=6, max_players=6, n_games=400, starting_bbs=30, w1=.5, w2=.3, K=100, alpha=50)
    ref = ref_loader.make_poker(hr_table, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)

    class RefAdapter(pu.TorchAdapter):
        def step(self, actions):
            obs, rew, done, _, _ = self.env.step(torch.from_numpy(actions))
            return obs, rew, done

    pu.random_rollout_compare(RefAdapter(ref), orc, np.random.default_rng(5), 6, 400, 3, 40, rot=1)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_packed_kernel_body_hostbuild_matches_reference_golden(name, hr_table):
    # the packed layout keeps no equity tensor: its mirror (pulse_poker_unpack_full) rebuilds the reference's cache
    pu.replay_golden(name, EmuPokerPacked(hr_table, **pu.golden_kwargs(name)))


@pytest.mark.parametrize("P,M,N,vary,rot", [(2, 2, 3000, False, 0), (6, 6, 1500, False, 1), (10, 10, 800, True, 3),
                                              (3, 10, 800, True, 2)])
def test_packed_kernel_body_hostbuild_matches_oracle_random(P, M, N, vary, rot, hr_table):
    kw = dict(n_players=P, max_players=M, n_games=N, starting_bbs=50, max_bbs=200, w1=.5, w2=.3, K=100, alpha=50)
    rng = np.random.default_rng(P * 10 + M)
    pu.random_rollout_compare(EmuPokerPacked(hr_table, **kw), pyoracle.OraclePoker(hr_table, **kw), rng, P, N, 5, 45,
                              vary_ap=vary, rot=rot, fields=pu.MIRROR_FIELDS)
