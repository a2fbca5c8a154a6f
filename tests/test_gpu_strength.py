"""GPU suite for the enumerated hand strength: pulse_poker_packed_enumerated_strength against the
oracle's brute-force enumeration on tables played to every street, and against the reference's river equity."""
import numpy as np
import pytest
import torch

from oracle import pyoracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,N", [(2, 3000), (6, 1500)])
def test_enumerated_strength_matches_bruteforce_on_every_street(P, N, cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    env = PokerGPUPacked(cuda_device, seed=21, env_base=77, n_players=P, max_players=P, n_games=N, starting_bbs=80,
                         max_bbs=300, w1=.5, w2=.3, K=100, alpha=50)
    env.reset()
    rng = np.random.default_rng(P)
    seen = set()
    for t in range(5 * P):
        # mostly calls, a few raises and folds: tables spread over preflop .. river .. finished
        a = rng.choice([1, 1, 1, 1, 2, 0], N).astype(np.int64)
        env.step(torch.from_numpy(a).to(cuda_device))
        if t % P != P - 1:
            continue
        snap = {k: v.cpu().numpy() for k, v in env.unpack().items()}
        got = env.enumerated_strength().cpu().numpy()
        want = pyoracle.enumerated_strength(snap["stages"], snap["hands"], np.maximum(snap["board"], 0), P)
        np.testing.assert_array_equal(got, want, err_msg=f"step {t}")
        seen |= set(np.unique(snap["stages"]).tolist())
        river = snap["stages"] == 3
        if river.any():   # the reference's own river equity (table walk) is the same number
            cards = np.concatenate([snap["hands"][river][:, 0], snap["board"][river]], axis=1).astype(np.int64)
            p = np.full(len(cards), 53, np.int64)
            for i in range(7):
                p = hr_table[p + cards[:, i]]
            ref = np.clip((p.astype(np.float32) - np.float32(4109)) / np.float32(32765), 0, 1)
            np.testing.assert_array_equal(got[river, 0], ref)
    assert {1, 2, 3} <= seen


def test_enumerated_strength_empty_and_null(cuda_device):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    env = PokerGPUPacked(cuda_device, n_players=2, max_players=2, n_games=0, starting_bbs=80, max_bbs=300, w1=.5, w2=.3,
                         K=100, alpha=50)
    assert env.enumerated_strength().shape == (0, 2)
