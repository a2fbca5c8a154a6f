"""GPU parity (through the C ABI) for Blackjack, 2048/TFE and Tetris: the CUDA env classes against the
reference's golden trajectories (tests/golden/*.npz, recorded from the reference's own files) and against
the C oracle on seeded rollouts; plus size-independent properties at the BASELINE.json config sizes.
Everything here is integer work: bit-exact."""
from pathlib import Path

import numpy as np
import pytest

from oracle import pyoracle
from tests.test_env_oracles import BJ_FIELDS

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


def dev_t(x, dev, dtype=None):
    import torch

    t = torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    return t if dtype is None else t.to(dtype)


# ---- Blackjack -------------------------------------------------------------------------------------
def test_blackjack_cuda_matches_reference_golden(cuda_device):
    from pulselib_b200.envs.blackjack import BlackJack

    g = np.load(GOLD / "blackjack.npz")
    N, episodes, steps, seed, env_base = (int(x) for x in g["meta"])
    env = BlackJack(cuda_device, N, seed=seed, env_base=env_base)
    for e in range(episodes):
        obs, info = env.reset()                               # decks from the device-side chance tape
        assert info == {}
        np.testing.assert_array_equal(env.decks.cpu().numpy(), g["decks"][e])
        np.testing.assert_array_equal(obs.cpu().numpy(), g["reset_obs"][e])
        same = None
        for t in range(steps):
            out = env.step(dev_t(g["actions"][e, t], cuda_device))
            obs, rew, term, trunc, info = out
            assert trunc is None and info == {}               # blackjack.py:186
            if same is not None:                              # same tensor objects every call
                assert obs is same[0] and rew is same[1] and term is same[2]
            same = (obs, rew, term)
            for k in BJ_FIELDS:
                np.testing.assert_array_equal(getattr(env, k).cpu().numpy().astype(g[k].dtype), g[k][e, t],
                                              err_msg=f"{k} e{e} t{t}")


def test_blackjack_cuda_matches_oracle_config_size(cuda_device):
    """BASELINE configs[0]: 4096 parallel envs, random policy; whole-batch reset once all are terminated."""
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    N = 4096
    env, orc = BlackJack(cuda_device, N, seed=1, env_base=0), pyoracle.OracleBlackjack(N)
    rng = np.random.default_rng(1)
    live_steps = 0
    for e in range(40):
        orc.tape_decks(1, 0, e)
        np.testing.assert_array_equal(env.reset()[0].cpu().numpy(), orc.reset())
        while not orc.terminated.all():
            live_steps += int((~orc.terminated.astype(bool)).sum())
            a = rng.integers(0, 2, N)
            obs, rew, term, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
            o2, r2, t2 = orc.step(a)
            np.testing.assert_array_equal(obs.cpu().numpy(), o2)
            np.testing.assert_array_equal(rew.cpu().numpy(), r2)
            np.testing.assert_array_equal(term.cpu().numpy(), t2)
        np.testing.assert_array_equal(env.players_cards.cpu().numpy(), orc.players_cards)
        np.testing.assert_array_equal(env.dealer_cards.cpu().numpy(), orc.dealer_cards)
    assert live_steps > 200_000


def test_blackjack_injected_decks_and_errors(cuda_device):
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    N = 1000
    rng = np.random.default_rng(3)
    decks = np.argsort(rng.random((N, 52)), axis=1)
    env, orc = BlackJack(cuda_device, N), pyoracle.OracleBlackjack(N)
    obs, _ = env.reset(options={"decks": torch.from_numpy(decks)})
    np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(decks))
    with pytest.raises(ValueError):
        env.reset(options={"decks": torch.zeros((N, 51))})
    with pytest.raises(AssertionError):
        env.step(torch.zeros(N + 1, dtype=torch.int64, device=cuda_device))  # blackjack.py:114


# ---- 2048 / TFE ------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["tfe_4x4", "tfe_3x3"])
def test_tfe_cuda_matches_reference_golden(name, cuda_device):
    from pulselib_b200.envs.tfe import TFEBatched

    g = np.load(GOLD / f"{name}.npz")
    side, N, steps, seed, env_base = (int(x) for x in g["meta"])
    env = TFEBatched(cuda_device, N, side=side, seed=seed, env_base=env_base, auto_reset=False)
    obs, info = env.reset()                                   # tape counter 0
    np.testing.assert_array_equal(obs.cpu().numpy(), g["reset_boards"])
    env._t = 0                                                # the fixture numbers step counters from 0
    for t in range(steps):
        obs, rew, done, trunc, info = env.step(dev_t(g["actions"][t], cuda_device))
        np.testing.assert_array_equal(obs.cpu().numpy(), g["boards"][t], err_msg=f"t{t}")
        np.testing.assert_array_equal(rew.cpu().numpy(), g["rewards"][t])
        np.testing.assert_array_equal(done.cpu().numpy(), g["done"][t].astype(bool))
        np.testing.assert_array_equal(info["score"].cpu().numpy(), g["scores"][t])
        assert trunc is False


@pytest.mark.parametrize("side,act_dtype", [(4, "int64"), (4, "uint8"), (3, "int64"), (2, "uint8")])
def test_tfe_cuda_matches_oracle_autoreset(side, act_dtype, cuda_device):
    import torch
    from pulselib_b200.envs.tfe import TFEBatched

    N, steps = 50_001, 300                                    # ragged: not a multiple of the block size
    rng = np.random.default_rng(side)
    env = TFEBatched(cuda_device, N, side=side, seed=9, env_base=77, auto_reset=True)
    orc = pyoracle.OracleTFE(N, side)
    np.testing.assert_array_equal(env.reset()[0].cpu().numpy(), orc.reset(orc.tape(9, 77, pyoracle.STREAM_TFE_RESET, 0, 4)))
    for t in range(1, steps):
        a = rng.integers(0, 4, N)
        b2, r2, d2 = orc.step(a, orc.tape(9, 77, pyoracle.STREAM_TFE_SPAWN, t, 2), auto_reset=True,
                              reset_chance4=orc.tape(9, 77, pyoracle.STREAM_TFE_RESET, t, 4))
        obs, rew, done, _, info = env.step(torch.from_numpy(a).to(cuda_device).to(getattr(torch, act_dtype)))
        if t % 10 == 0 or t == steps - 1:
            np.testing.assert_array_equal(obs.cpu().numpy(), b2, err_msg=f"t{t}")
            np.testing.assert_array_equal(rew.cpu().numpy(), r2)
            np.testing.assert_array_equal(done.cpu().numpy(), d2)
            np.testing.assert_array_equal(info["score"].cpu().numpy(), orc.scores)
    np.testing.assert_array_equal(env.board_values().cpu().numpy(), b2)


def test_tfe_injected_chance_and_large_tiles(cuda_device):
    import torch
    from pulselib_b200.envs.tfe import TFEBatched

    rng = np.random.default_rng(0)
    N = 4096
    exps = rng.integers(0, 15, (N, 4, 4))
    exps[rng.random((N, 4, 4)) < 0.3] = 0
    vals = np.where(exps > 0, 1 << exps, 0).astype(np.int32)
    for a in range(4):
        env, orc = TFEBatched(cuda_device, N, auto_reset=False), pyoracle.OracleTFE(N, 4)
        env.set_board_values(vals)
        orc.boards[...] = vals
        ch = rng.integers(0, 2**32, (N, 2), dtype=np.uint64).astype(np.uint32)
        obs, rew, done, _, info = env.step(torch.full((N,), a, dtype=torch.int64, device=cuda_device), chance=ch)
        b2, r2, d2 = orc.step(np.full(N, a), ch)
        np.testing.assert_array_equal(obs.cpu().numpy(), b2)
        np.testing.assert_array_equal(rew.cpu().numpy(), r2)
        np.testing.assert_array_equal(done.cpu().numpy(), d2)
        np.testing.assert_array_equal(info["score"].cpu().numpy(), orc.scores)
    with pytest.raises(ValueError):
        TFEBatched(cuda_device, 2).set_board_values(np.full((2, 4, 4), 3))


def test_tfe_single_board_env_is_reference_shaped(cuda_device):
    from pulselib_b200 import registration
    from pulselib_b200.gym_compat import gym

    registration.register_all()
    env = gym.make("Pulse-2048-v2", board_height=4, board_width=4)
    board, info = env.reset(seed=3)
    assert isinstance(board, np.ndarray) and board.dtype == np.int32 and board.shape == (4, 4)
    assert (board > 0).sum() == 2 and info == {"score": 0}
    total = 0
    for t in range(50):
        b2, rew, done, trunc, info = env.step(t % 4)
        assert b2 is board and isinstance(rew, int) and isinstance(done, bool) and trunc is False
        total = info["score"]
    assert total > 0


def test_tfe_full_size_properties(cuda_device):
    """BASELINE configs[1]: 1M boards, random moves, seeded spawns, auto-reset.  Properties that hold for any
    size: the tile sum grows by exactly 2 or 4 per step while a board has room (merging conserves the sum),
    scores never decrease within a game, the same seed reproduces the run, and sharding the env range over two
    engines (env_base) gives the same boards as one engine."""
    import torch
    from pulselib_b200.envs.tfe import TFEBatched

    N = 1 << 20
    def run(n, base):
        env = TFEBatched(cuda_device, n, seed=2, env_base=base, auto_reset=True)
        env.reset()
        gen = torch.Generator(device=cuda_device)
        gen.manual_seed(base)
        chk = torch.zeros((), dtype=torch.int64, device=cuda_device)
        prev_sum = env.obs.sum(dim=(1, 2)).clone()
        prev_score = env.total_score.clone()
        prev_done = env.done.clone()
        for t in range(64):
            a = torch.randint(0, 4, (n,), dtype=torch.int64, device=cuda_device, generator=gen)
            obs, rew, done, _, info = env.step(a)
            s = obs.sum(dim=(1, 2))
            grew = s - prev_sum
            live = ~prev_done
            full_before = False
            assert bool(((grew[live] == 2) | (grew[live] == 4) | (grew[live] == 0)).all())
            assert bool((info["score"][live] >= prev_score[live]).all())
            assert bool((s[prev_done] <= 8).all())        # auto-reset boards restart with two tiles
            chk += (obs.long().sum() * (t + 1)) + rew.long().sum()
            prev_sum, prev_score, prev_done = s.clone(), info["score"].clone(), done.clone()
        return env, int(chk)
    e1, c1 = run(N, 0)
    e2, c2 = run(N, 0)
    assert c1 == c2
    half = N // 2
    ea = TFEBatched(cuda_device, half, seed=2, env_base=0)
    eb = TFEBatched(cuda_device, half, seed=2, env_base=half)
    ea.reset(); eb.reset()
    whole = TFEBatched(cuda_device, N, seed=2, env_base=0)
    whole.reset()
    gen = torch.Generator(device=cuda_device); gen.manual_seed(1)
    for t in range(32):
        a = torch.randint(0, 4, (N,), dtype=torch.int64, device=cuda_device, generator=gen)
        whole.step(a); ea.step(a[:half].contiguous()); eb.step(a[half:].contiguous())
    assert torch.equal(whole.boards[:half], ea.boards) and torch.equal(whole.boards[half:], eb.boards)


# ---- Tetris ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("act_dtype", ["int64", "uint8"])
def test_tetris_cuda_matches_spec_oracle(act_dtype, cuda_device):
    import torch
    from pulselib_b200.envs.tetris import TetrisBatched

    N, steps = 20_011, 300
    rng = np.random.default_rng(4)
    env, orc = TetrisBatched(cuda_device, N, seed=4, env_base=31, auto_reset=True), pyoracle.OracleTetris(N)
    rows, info = env.reset()
    orc.reset(orc.tape_pieces(4, 31, 0, 1))
    np.testing.assert_array_equal(info["piece"].cpu().numpy(), orc.piece)
    for t in range(1, steps):
        a = rng.integers(0, 40, N)
        rows2, r2, d2 = orc.step(a, orc.tape_pieces(4, 31, t, 0), auto_reset=True, reset_piece=orc.tape_pieces(4, 31, t, 1))
        rows, rew, done, trunc, info = env.step(torch.from_numpy(a).to(cuda_device).to(getattr(torch, act_dtype)))
        if t % 10 == 0 or t == steps - 1:
            np.testing.assert_array_equal(rows.cpu().numpy().view(np.uint16), rows2, err_msg=f"t{t}")
            np.testing.assert_array_equal(rew.cpu().numpy(), r2)
            np.testing.assert_array_equal(done.cpu().numpy(), d2)
            np.testing.assert_array_equal(info["piece"].cpu().numpy(), orc.piece)
            np.testing.assert_array_equal(info["lines"].cpu().numpy(), orc.lines)


def test_tetris_guided_line_clears_and_injected_pieces(cuda_device):
    import torch
    from pulselib_b200.envs.tetris import TetrisBatched

    N = 300
    env, orc = TetrisBatched(cuda_device, N, auto_reset=False), pyoracle.OracleTetris(N)
    first = np.zeros(N, np.uint8)
    env.reset(options={"first_piece": first})
    orc.reset(first)
    seq = [(0, 0), (0, 4), (1, 8)]
    total = 0
    for t in range(60):
        nxt = np.full(N, seq[(t + 1) % 3][0], np.uint8)
        a = np.full(N, seq[t % 3][1])
        rows, rew, done, _, info = env.step(torch.from_numpy(a).to(cuda_device), next_piece=nxt, reset_piece=nxt)
        rows2, r2, d2 = orc.step(a, nxt)
        np.testing.assert_array_equal(rows.cpu().numpy().view(np.uint16), rows2)
        np.testing.assert_array_equal(rew.cpu().numpy(), r2)
        total += int(r2[0])
    assert total == 20 and int(info["lines"][0]) == 20
    assert env.cells().shape == (N, 20, 10)


def test_tetris_full_size_properties(cuda_device):
    """BASELINE configs[3]: 2M envs.  Cell count: +4 per placement, -10 per cleared row; reset boards are
    empty; no row is ever full after a step; same seed, same run."""
    import torch
    from pulselib_b200.envs.tetris import TetrisBatched

    N = 1 << 21
    def popcount(rows):
        x = rows.to(torch.int32) & 0x3FF
        x = x - ((x >> 1) & 0x555)
        x = (x & 0x333) + ((x >> 2) & 0x333)
        x = (x + (x >> 4)) & 0xF0F
        return ((x & 0xFF) + (x >> 8)).sum(dim=1)
    sums = []
    for rep in range(2):
        env = TetrisBatched(cuda_device, N, seed=4, auto_reset=True)
        env.reset()
        gen = torch.Generator(device=cuda_device); gen.manual_seed(7)
        filled_prev = torch.zeros(N, dtype=torch.int64, device=cuda_device)
        prev_done = torch.zeros(N, dtype=torch.bool, device=cuda_device)
        chk = torch.zeros((), dtype=torch.int64, device=cuda_device)
        for t in range(48):
            a = torch.randint(0, 40, (N,), dtype=torch.int64, device=cuda_device, generator=gen)
            rows, rew, done, _, info = env.step(a)
            filled = popcount(rows)
            live = ~prev_done & ~done
            assert torch.equal(filled[live], (filled_prev + 4 - 10 * rew.long())[live])
            assert bool((filled[prev_done] == 0).all()) and not bool(done[prev_done].any())
            newly = done & ~prev_done
            assert torch.equal(filled[newly], filled_prev[newly])
            assert not bool(((rows.to(torch.int32) & 0x3FF) == 0x3FF).any())
            chk += (rows.long().sum() * (t + 1)) + rew.long().sum()
            filled_prev, prev_done = filled, done.clone()
        assert bool(prev_done.any())
        sums.append(int(chk))
    assert sums[0] == sums[1]


def test_tfe_domain_check_flags_32768_tiles(cuda_device):
    """tiles above 32768 are outside the packed board's domain (4 bits per cell): the audit must say so"""
    import torch
    from pulselib_b200 import _lib
    from pulselib_b200.envs.tfe import TFEBatched

    env = TFEBatched(cuda_device, 64, side=4, seed=1)
    env.reset()
    assert env.domain_check() == (0, 0)
    vals = env.board_values().clone()
    vals[3, 0, 0] = 32768
    vals[5, 0, 0] = 32768; vals[5, 0, 1] = 32768
    env.set_board_values(vals)
    assert env.domain_check(raise_on_overflow=False) == (2, 1)
    with pytest.raises(_lib.PulseError):
        env.domain_check()


def test_blackjack_rollout_equals_k_steps(cuda_device):
    """pulse_blackjack_rollout: K steps in one launch == K step() calls, against the oracle too"""
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    n, K = 4096, 9
    a = BlackJack(device=cuda_device, batch_size=n, seed=3, env_base=10)
    b = BlackJack(device=cuda_device, batch_size=n, seed=3, env_base=10)
    orc = pyoracle.OracleBlackjack(n)
    rng = np.random.default_rng(8)
    for episode in range(3):
        acts = rng.integers(0, 3, (K, n)).astype(np.int64)      # 2 = neither hit nor stand: a no-op
        t_acts = torch.from_numpy(acts).to(cuda_device)
        if episode == 1:
            t_acts = t_acts.to(torch.uint8)                      # byte-wide actions
        obs_k, rew_k, term_k, first = a.rollout(t_acts, reset_first=True)
        o0, _ = b.reset()
        orc.tape_decks(3, 10, episode)
        np.testing.assert_array_equal(first.cpu().numpy(), orc.reset(None))
        np.testing.assert_array_equal(first.cpu().numpy(), o0.cpu().numpy())
        for k in range(K):
            o, r, t, _, _ = b.step(torch.from_numpy(acts[k]).to(cuda_device))
            o2, r2, t2 = orc.step(acts[k])
            np.testing.assert_array_equal(obs_k[k].cpu().numpy(), o.cpu().numpy(), err_msg=f"ep {episode} step {k} obs")
            np.testing.assert_array_equal(rew_k[k].cpu().numpy(), r.cpu().numpy())
            np.testing.assert_array_equal(term_k[k].cpu().numpy(), t.cpu().numpy())
            np.testing.assert_array_equal(obs_k[k].cpu().numpy(), o2)
            np.testing.assert_array_equal(rew_k[k].cpu().numpy(), r2)
        # a second rollout without reset carries on from the state the first one left
        more = rng.integers(0, 2, (2, n)).astype(np.int64)
        ok, rk, tk = a.rollout(torch.from_numpy(more).to(cuda_device))
        for k in range(2):
            o, r, t, _, _ = b.step(torch.from_numpy(more[k]).to(cuda_device))
            np.testing.assert_array_equal(ok[k].cpu().numpy(), o.cpu().numpy())
            np.testing.assert_array_equal(tk[k].cpu().numpy(), t.cpu().numpy())


@pytest.mark.parametrize("side,n", [(4, 4096), (4, 1 << 18), (3, 1000)])
def test_tfe_rollout_equals_k_steps(side, n, cuda_device):
    """pulse_tfe_rollout: K steps in one launch == K step() calls (same tape counters), auto-reset included"""
    import torch
    from pulselib_b200.envs.tfe import TFEBatched

    K = 24
    a = TFEBatched(cuda_device, n, side=side, seed=7, env_base=5, auto_reset=True)
    b = TFEBatched(cuda_device, n, side=side, seed=7, env_base=5, auto_reset=True)
    a.reset(); b.reset()
    rng = np.random.default_rng(side)
    for rnd in range(4 if side == 3 else 2):              # 3x3 boards finish within tens of moves: resets happen
        acts = torch.from_numpy(rng.integers(0, 4, (K, n)).astype(np.uint8 if rnd % 2 else np.int64)).to(cuda_device)
        obs_k, rew_k, done_k = a.rollout(acts)
        for k in range(K):
            o, r, d, _, info = b.step(acts[k])
            np.testing.assert_array_equal(obs_k[k].cpu().numpy(), o.cpu().numpy(), err_msg=f"round {rnd} step {k} boards")
            np.testing.assert_array_equal(rew_k[k].cpu().numpy(), r.cpu().numpy())
            np.testing.assert_array_equal(done_k[k].cpu().numpy(), d.cpu().numpy())
        np.testing.assert_array_equal(a.total_score.cpu().numpy(), b.total_score.cpu().numpy())
        np.testing.assert_array_equal(a.boards.cpu().numpy(), b.boards.cpu().numpy())
    if side == 3:
        assert int(done_k.sum()) > 0


def test_blackjack_rollout_many_hands_per_launch(cuda_device):
    """n_hands hands of K steps in one launch == n_hands x (reset(), K x step())"""
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    n, K, H = 4096, 8, 5
    a = BlackJack(device=cuda_device, batch_size=n, seed=9, env_base=3)
    b = BlackJack(device=cuda_device, batch_size=n, seed=9, env_base=3)
    rng = np.random.default_rng(2)
    acts = rng.integers(0, 2, (H * K, n)).astype(np.uint8)
    obs, rew, term, deals = a.rollout(torch.from_numpy(acts).to(cuda_device), reset_first=True, n_hands=H)
    for h in range(H):
        o0, _ = b.reset()
        np.testing.assert_array_equal(deals[h].cpu().numpy(), o0.cpu().numpy(), err_msg=f"hand {h} deal")
        for k in range(K):
            o, r, t, _, _ = b.step(torch.from_numpy(acts[h * K + k].astype(np.int64)).to(cuda_device))
            i = h * K + k
            np.testing.assert_array_equal(obs[i].cpu().numpy(), o.cpu().numpy(), err_msg=f"hand {h} step {k}")
            np.testing.assert_array_equal(rew[i].cpu().numpy(), r.cpu().numpy())
            np.testing.assert_array_equal(term[i].cpu().numpy(), t.cpu().numpy())
