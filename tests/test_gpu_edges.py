"""GPU suite: edge sizes -- empty batches, a single env, sizes around the block/warp boundaries, the widest table
(10 seats observed at max_players 16) -- through the public env classes, against the oracle."""
import numpy as np
import pytest
import torch

from oracle import pyoracle
from tests import poker_util as pu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N", [0, 1, 31, 33, 127, 129])
def test_poker_packed_tiny_batches(N, cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=3, max_players=3, n_games=N, starting_bbs=20, max_bbs=60, w1=.5, w2=.3, K=100, alpha=50)
    env = PokerGPUPacked(cuda_device, seed=1, env_base=2, **kw)
    obs, _ = env.reset()
    assert tuple(obs.shape) == (N, 19)
    if N == 0:
        o, r, d, _, _ = env.step(torch.zeros(0, dtype=torch.int64, device=cuda_device))
        assert o.shape[0] == 0 and r.shape[0] == 0 and d.shape[0] == 0
        return
    orc = pyoracle.OraclePoker(hr_table, **kw)
    orc.tape_decks(1, 2, 0)
    np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(None))
    rng = np.random.default_rng(N)
    for t in range(25):
        a = rng.integers(0, 13, N).astype(np.int64)
        o1, r1, d1, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
        o2, r2, d2 = orc.step(a)
        np.testing.assert_array_equal(o1.cpu().numpy(), o2)
        np.testing.assert_array_equal(d1.cpu().numpy(), d2)
        np.testing.assert_allclose(r1.cpu().numpy(), r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50))


def test_poker_widest_table_both_layouts(cuda_device, hr_table):
    """10 seats, observation sized for 16 (the ABI's maximum): 58 columns, generic (non-specialised) kernels."""
    from pulselib_b200.envs.poker import PokerGPU
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    N = 2049
    kw = dict(n_players=10, max_players=16, n_games=N, starting_bbs=15, max_bbs=40, w1=.5, w2=.3, K=100, alpha=50)
    rng = np.random.default_rng(10)
    for cls in (PokerGPU, PokerGPUPacked):
        env = cls(cuda_device, [], **kw) if cls is PokerGPU else cls(cuda_device, **kw)
        orc = pyoracle.OraclePoker(hr_table, **kw)
        for hand in range(2):
            decks = (np.argsort(rng.random((N, 52)), axis=1) + 1).astype(np.int32)
            obs, _ = env.reset(options={"prefixed_decks": torch.from_numpy(decks), "rotation": hand})
            assert tuple(obs.shape) == (N, 58)
            np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(decks, rotation=hand))
            for t in range(60):
                a = rng.integers(0, 13, N).astype(np.int64)
                o1, r1, d1, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
                o2, r2, d2 = orc.step(a)
                np.testing.assert_array_equal(o1.cpu().numpy(), o2, err_msg=f"{cls.__name__} hand {hand} step {t}")
                np.testing.assert_array_equal(d1.cpu().numpy(), d2)
                np.testing.assert_allclose(r1.cpu().numpy(), r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50))


@pytest.mark.parametrize("N", [0, 1, 1025])
def test_other_envs_tiny_batches(N, cuda_device):
    from pulselib_b200.envs.blackjack import BlackJack
    from pulselib_b200.envs.particle2d import Particle2D
    from pulselib_b200.envs.tetris import TetrisBatched
    from pulselib_b200.envs.tfe import TFEBatched

    tfe = TFEBatched(cuda_device, N, side=4, seed=1)
    obs, _ = tfe.reset()
    assert tuple(obs.shape) == (N, 4, 4)
    o, r, d, _, _ = tfe.step(torch.zeros(N, dtype=torch.int64, device=cuda_device))
    assert o.shape[0] == N and r.shape[0] == N and d.shape[0] == N
    if N:
        assert int((o > 0).sum(dim=(1, 2)).min()) >= 2           # two tiles after reset, one more spawned
    tet = TetrisBatched(cuda_device, N, seed=1)
    tet.reset()
    rows, r, d, _, _ = tet.step(torch.zeros(N, dtype=torch.int64, device=cuda_device))
    assert rows.shape[0] == N
    bj = BlackJack(cuda_device, N, seed=1)
    obs, _ = bj.reset()
    o, r, t, _, _ = bj.step(torch.ones(N, dtype=torch.int64, device=cuda_device))
    assert tuple(o.shape) == (N, 3) and (N == 0 or bool(t.all()))   # stand terminates every env
    p2 = Particle2D(cuda_device, N, seed=1)
    s, _ = p2.reset()
    s2, r, t, _, _ = p2.step(torch.zeros((N, 2), device=cuda_device))
    assert tuple(s2.shape) == (N, 4)
