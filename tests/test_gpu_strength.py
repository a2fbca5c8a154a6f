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


# ---- showdown equity (N4, second half): all hole cards known -----------------------------------------------------
@pytest.mark.parametrize("P,N", [(2, 4000), (6, 2000), (10, 1000)])
def test_showdown_equity_matches_oracle_on_every_street(P, N, cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    env = PokerGPUPacked(cuda_device, seed=33, env_base=1 << 33, n_players=P, max_players=P, n_games=N, starting_bbs=80,
                         max_bbs=300, w1=.5, w2=.3, K=100, alpha=50)
    env.reset()
    rng = np.random.default_rng(40 + P)
    seen = set()

    def check(tag, counter):
        snap = {k: v.cpu().numpy() for k, v in env.unpack().items()}
        got = env.showdown_equity(preflop_samples=96, counter=counter).cpu().numpy()
        want = pyoracle.showdown_equity(snap["hands"], snap["status"], snap["board"], P, samples=96, seed=33,
                                        env_base=1 << 33, counter=counter, hand_ranks=hr_table)
        np.testing.assert_array_equal(got, want, err_msg=tag)
        out = (snap["status"] == 1) | (snap["status"] == 3)
        assert np.all(got[out] == 0.0)
        # a partition of the pot wherever somebody is left to win it (a settled table may show every seat folded)
        np.testing.assert_allclose(got.sum(1), (~out).any(1).astype(np.float32), atol=1e-6)
        return snap

    check("after reset (every table preflop: sampled boards)", 5)
    for t in range(5 * P):
        a = rng.choice([1, 1, 1, 1, 2, 0], N).astype(np.int64)
        env.step(torch.from_numpy(a).to(cuda_device))
        if t % P != P - 1:
            continue
        snap = check(f"step {t}", t)
        seen |= set(np.unique(snap["stages"]).tolist())
    assert {1, 2, 3, 5} <= seen


def test_showdown_equity_default_counter_advances_and_bad_arguments(cuda_device):
    from pulselib_b200 import _lib
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    env = PokerGPUPacked(cuda_device, seed=1, n_players=2, max_players=2, n_games=512)
    env.reset()
    a, b = env.showdown_equity(preflop_samples=64), env.showdown_equity(preflop_samples=64)
    assert not torch.equal(a, b), "successive calls draw fresh preflop boards"
    assert torch.equal(env.showdown_equity(preflop_samples=64, counter=0), a)
    for bad in (0, 65537):
        with pytest.raises(_lib.PulseError):
            env.showdown_equity(preflop_samples=bad)
    empty = PokerGPUPacked(cuda_device, n_players=2, max_players=2, n_games=0)
    assert empty.showdown_equity().shape == (0, 2)


def test_showdown_equity_full_batch_timing(cuda_device):
    """1,048,576 heads-up tables spread over the streets: finishes, partitions the pot, and reports its cost."""
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    N = 1 << 20
    env = PokerGPUPacked(cuda_device, seed=2, n_players=2, max_players=2, n_games=N)
    env.reset()
    g = torch.Generator(device=cuda_device).manual_seed(0)
    for _ in range(3):
        env.step(torch.randint(1, 3, (N,), generator=g, device=cuda_device))
    env.showdown_equity(preflop_samples=256, counter=0)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    eq = env.showdown_equity(preflop_samples=256, counter=1)
    t1.record()
    torch.cuda.synchronize(cuda_device)
    ms = t0.elapsed_time(t1)
    stages = env.unpack(("stages",))["stages"]
    hist = torch.bincount(stages, minlength=6).tolist()
    print(f"showdown_equity: {N} heads-up tables, stage histogram {hist}, {ms:.2f} ms")
    assert torch.allclose(eq.sum(1), torch.ones(N, device=cuda_device), atol=1e-6)
    assert ms < 2000
