"""GPU parity: pulselib_b200.envs.poker.PokerGPU (CUDA, through the C ABI) against the reference's
golden vectors and against the oracle on seeded rollouts.  Integer state, obs, done and equities are
bit-exact; reward is fp32 through tanh, compared at rtol 2e-6 / atol alpha*4e-7."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu

pytestmark = pytest.mark.gpu


def make_env(dev, **kw):
    from pulselib_b200.envs.poker import PokerGPU

    return PokerGPU(device=dev, agents=[], **kw)


@pytest.mark.parametrize("name", ["poker_hu", "poker_6max", "poker_10max_varap"])
def test_cuda_matches_reference_golden(name, cuda_device):
    pu.replay_golden(name, pu.TorchAdapter(make_env(cuda_device, **pu.golden_kwargs(name))))


@pytest.mark.parametrize("P,M,N,vary,rot", [(2, 2, 20000, False, 0), (6, 6, 8000, False, 1), (10, 10, 4000, True, 3),
                                              (3, 10, 4000, True, 2), (5, 8, 4000, True, 1)])
def test_cuda_matches_oracle_random(P, M, N, vary, rot, cuda_device, hr_table):
    kw = dict(n_players=P, max_players=M, n_games=N, starting_bbs=50, max_bbs=200, w1=.5, w2=.3, K=100, alpha=50)
    rng = np.random.default_rng(P * 100 + M)
    pu.random_rollout_compare(pu.TorchAdapter(make_env(cuda_device, **kw)), pyoracle.OraclePoker(hr_table, **kw),
                              rng, P, N, 4, 45, vary_ap=vary, rot=rot)


def test_cuda_tape_decks_match_oracle(cuda_device, hr_table):
    kw = dict(n_players=2, max_players=2, n_games=4096)
    env = make_env(cuda_device, seed=20260401, env_base=777, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    for hand in range(3):
        env.reset()
        orc.tape_decks(20260401, 777, hand)
        np.testing.assert_array_equal(env.decks.cpu().numpy(), orc.decks)


def test_cuda_evaluator_matches_table(cuda_device, hr_table):
    import torch
    from pulselib_b200 import _lib

    lib = _lib.load()
    rng = np.random.default_rng(3)
    for n in (5, 6, 7):
        cards = np.argsort(rng.random((300000, 52)), axis=1)[:, :n].astype(np.int32) + 1
        p = np.full(len(cards), 53, dtype=np.int64)
        for i in range(n):
            p = hr_table[p + cards[:, i]]
        want = p if n == 7 else hr_table[p]
        d = torch.from_numpy(cards).to(cuda_device)
        out = torch.empty(len(cards), dtype=torch.int32, device=cuda_device)
        _lib.check(lib.pulse_poker_eval_batch(_lib.ptr(d), n, _lib.ptr(out), len(cards), _lib.cuda_stream(cuda_device)))
        np.testing.assert_array_equal(out.cpu().numpy(), want)


def test_internal_methods_and_instance_override(cuda_device, hr_table):
    """execute_actions / calculate_equities / resolve_* are callable one at a time, and a per-instance
    calculate_equities replacement is what step() uses (tests/poker/test_poker_gpu_round_progression.py:224-238)."""
    import torch

    kw = dict(n_players=2, max_players=2, n_games=64, starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)
    env = make_env(cuda_device, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    rng = np.random.default_rng(9)
    decks = (np.argsort(rng.random((64, 52)), axis=1) + 1).astype(np.int32)
    env.reset(options={"prefixed_decks": torch.from_numpy(decks)})
    orc.reset(decks)
    a = rng.integers(0, 13, 64)
    env.execute_actions(torch.from_numpy(a).to(cuda_device))
    orc.call("execute_actions", pyoracle.p(np.ascontiguousarray(a, dtype=np.int64)))
    pu.compare_state(env, orc, "execute_actions")
    calls = []

    def fake():
        calls.append(1)
        env.equities[:] = 0.25
        env.equity_dirty[:] = False

    env.calculate_equities = fake
    obs, rew, done, trunc, info = env.step(torch.ones(64, dtype=torch.int64, device=cuda_device))
    assert calls and float(env.equities.max()) == 0.25
    assert set(info) == {"active_players", "stacks", "seat_idx"}
    assert obs is env.obs and trunc is env.is_truncated


def test_bad_inputs_raise(cuda_device):
    import torch

    env = make_env(cuda_device, n_players=2, max_players=2, n_games=8)
    with pytest.raises(ValueError, match="prefixed_decks must have shape"):
        env.reset(options={"prefixed_decks": torch.zeros((4, 52), dtype=torch.int32)})
    env.reset()
    with pytest.raises(ValueError):
        env.step(torch.zeros(3, dtype=torch.int64))
