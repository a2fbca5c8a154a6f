"""GPU suite: 1,048,576 seeded trajectories per integer env, CUDA vs the oracle, bit for bit
(BASELINE.json north_star: "bit-exact with the reference on 1e6 seeded trajectories per integer env").

The oracle is the C restatement that the CPU suite pins on the reference's golden vectors and on the
reference's own unit tests; chance comes from the shared Philox tape on both sides (the device draws it,
the oracle re-derives it on the host), actions from one seeded numpy stream."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu

pytestmark = pytest.mark.gpu
N = 1 << 20
FIELDS = pu.INT_N + pu.INT_NP + ("board", "hands")


@pytest.mark.parametrize("P,steps", [(2, 16), (6, 30)])
def test_poker_packed_million_tables(P, steps, cuda_device, hr_table):
    import torch
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=P, max_players=P, n_games=N, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50)
    env = PokerGPUPacked(cuda_device, seed=20260401, env_base=123456789, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    rng = np.random.default_rng(P)
    for hand in range(2):                                   # second hand: persistent stacks, moved button
        obs, _ = env.reset()
        orc.tape_decks(20260401, 123456789, hand)
        np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(None), err_msg=f"reset obs hand {hand}")
        for t in range(steps):
            a = rng.integers(0, 13, N).astype(np.int64)
            o1, r1, d1, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
            o2, r2, d2 = orc.step(a)
            tag = f"P{P} hand {hand} step {t}"
            np.testing.assert_array_equal(o1.cpu().numpy(), o2, err_msg=tag + " obs")
            np.testing.assert_array_equal(d1.cpu().numpy(), d2, err_msg=tag + " done")
            np.testing.assert_allclose(r1.cpu().numpy(), r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50), err_msg=tag)
            if t % 5 == 4 or t == steps - 1:
                snap = env.unpack()
                for f in FIELDS:
                    assert np.array_equal(snap[f].cpu().numpy(), getattr(orc, f)), f"{tag} field {f}"
        assert d2.mean() > 0.99                            # the trajectories really ran to the end of the hand


def test_tfe_million_boards(cuda_device):
    import torch
    from pulselib_b200.envs.tfe import TFEBatched

    steps = 200
    rng = np.random.default_rng(2)
    env = TFEBatched(cuda_device, N, side=4, seed=2, env_base=1 << 33, auto_reset=True)
    orc = pyoracle.OracleTFE(N, 4)
    base = 1 << 33
    np.testing.assert_array_equal(env.reset()[0].cpu().numpy(), orc.reset(orc.tape(2, base, pyoracle.STREAM_TFE_RESET, 0, 4)))
    for t in range(1, steps):
        a = rng.integers(0, 4, N)
        b2, r2, d2 = orc.step(a, orc.tape(2, base, pyoracle.STREAM_TFE_SPAWN, t, 2), auto_reset=True,
                              reset_chance4=orc.tape(2, base, pyoracle.STREAM_TFE_RESET, t, 4))
        obs, rew, done, _, info = env.step(torch.from_numpy(a).to(cuda_device).to(torch.uint8))
        if t % 20 == 0 or t == steps - 1:
            np.testing.assert_array_equal(obs.cpu().numpy(), b2, err_msg=f"t{t} boards")
            np.testing.assert_array_equal(rew.cpu().numpy(), r2, err_msg=f"t{t} reward")
            np.testing.assert_array_equal(done.cpu().numpy(), d2, err_msg=f"t{t} done")
            np.testing.assert_array_equal(info["score"].cpu().numpy(), orc.scores, err_msg=f"t{t} score")


def test_tetris_million_envs(cuda_device):
    import torch
    from pulselib_b200.envs.tetris import TetrisBatched

    steps = 120
    rng = np.random.default_rng(4)
    env, orc = TetrisBatched(cuda_device, N, seed=4, env_base=977, auto_reset=True), pyoracle.OracleTetris(N)
    env.reset()
    orc.reset(orc.tape_pieces(4, 977, 0, 1))
    for t in range(1, steps):
        a = rng.integers(0, 40, N)
        rows2, r2, d2 = orc.step(a, orc.tape_pieces(4, 977, t, 0), auto_reset=True, reset_piece=orc.tape_pieces(4, 977, t, 1))
        rows, rew, done, _, info = env.step(torch.from_numpy(a).to(cuda_device).to(torch.uint8))
        if t % 20 == 0 or t == steps - 1:
            np.testing.assert_array_equal(rows.cpu().numpy().view(np.uint16), rows2, err_msg=f"t{t} rows")
            np.testing.assert_array_equal(rew.cpu().numpy(), r2)
            np.testing.assert_array_equal(done.cpu().numpy(), d2)
            np.testing.assert_array_equal(info["lines"].cpu().numpy(), orc.lines)
    assert int(orc.lines.sum()) >= 0 and d2.sum() >= 0


def test_blackjack_million_envs(cuda_device):
    import torch
    from pulselib_b200.envs.blackjack import BlackJack

    env, orc = BlackJack(device=cuda_device, batch_size=N, seed=5, env_base=31), pyoracle.OracleBlackjack(N)
    rng = np.random.default_rng(5)
    for episode in range(2):
        obs, _ = env.reset()
        orc.tape_decks(5, 31, episode)
        np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(None))
        for t in range(8):
            a = rng.integers(0, 2, N)
            o1, r1, d1, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
            o2, r2, d2 = orc.step(a)
            np.testing.assert_array_equal(o1.cpu().numpy(), o2, err_msg=f"ep {episode} t{t} obs")
            np.testing.assert_array_equal(r1.cpu().numpy(), r2, err_msg=f"ep {episode} t{t} reward")
            np.testing.assert_array_equal(d1.cpu().numpy().astype(bool), np.asarray(d2).astype(bool))


# ---- the same trajectories the REFERENCE stepped (tests/golden/make_digests_1m.py), digest for digest -------------
class _CudaPokerEngine:
    def __init__(self, dev, name):
        from pulselib_b200.envs.poker_packed import PokerGPUPacked
        from tests import digest_util as du

        c = du.CASES[name]
        self.dev, self.c = dev, c
        self.env = PokerGPUPacked(dev, n_players=c["P"], max_players=c["M"], n_games=c["runs"] * c["n_run"],
                                  seed=c["seed"], env_base=0, **c["cfg"])

    def reset(self, hand, ap, rot):
        import torch
        from unittest import mock

        opts = {"rotation": rot}
        if ap != self.c["P"]:
            opts["active_players"] = True
            with mock.patch("torch.randint", lambda *a, **k: torch.tensor([ap])):
                obs, _ = self.env.reset(options=opts)
        else:
            obs, _ = self.env.reset(options=opts)
        self._snap = None
        return obs.cpu().numpy()

    def step(self, a):
        import torch

        o, r, d, _, _ = self.env.step(torch.from_numpy(a).to(self.dev))
        self._snap = None
        return o.cpu().numpy(), r.cpu().numpy(), d.cpu().numpy()

    def get(self, f):
        if self._snap is None:
            self._snap = {k: v.cpu().numpy() for k, v in self.env.unpack().items()}
        return self._snap[f]


@pytest.mark.parametrize("name", ["poker_hu", "poker_6max", "poker_10max"])
def test_poker_digests_from_reference(name, cuda_device):
    from tests import digest_util as du

    assert du.check_poker_case(name, _CudaPokerEngine(cuda_device, name)) >= 1 << 20


def test_blackjack_digests_from_reference(cuda_device):
    import torch
    from pulselib_b200.envs.blackjack import BlackJack
    from tests import digest_util as du

    c, want = du.CASES["blackjack"], du.load_digests()["blackjack"]["episode"]
    n = c["runs"] * c["n_run"]
    env = BlackJack(device=cuda_device, batch_size=n, seed=c["seed"], env_base=0)
    obs, _ = env.reset()
    assert du.hexd(du.digest(obs.cpu().numpy())) == want["reset_obs"]
    acts = du.case_actions("blackjack")
    for t in range(c["steps"]):
        o, r, d, _, _ = env.step(torch.from_numpy(acts[t]).to(cuda_device))
        w = want["steps"][t]
        assert du.hexd(du.digest(o.cpu().numpy())) == w["obs"] and du.hexd(du.digest(r.cpu().numpy())) == w["rewards"], f"step {t}"
        dn = d.cpu().numpy().astype(np.uint8)
        assert du.hexd(du.digest(dn)) == w["terminated"] and int(dn.sum()) == w["terminated_count"]


def test_tfe_digests_from_reference(cuda_device):
    import torch
    from pulselib_b200.envs.tfe import TFEBatched
    from tests import digest_util as du

    c, want = du.CASES["tfe_4x4"], du.load_digests()["tfe_4x4"]["episode"]
    env = TFEBatched(cuda_device, c["n"], side=c["side"], seed=c["seed"], env_base=0, auto_reset=False)
    assert du.hexd(du.digest(env.reset()[0].cpu().numpy())) == want["reset_boards"]
    acts = du.case_actions("tfe_4x4")
    for t in range(c["steps"]):
        obs, rew, done, _, info = env.step(torch.from_numpy(acts[t]).to(cuda_device).to(torch.uint8))
        w = want["steps"][t]
        assert du.hexd(du.digest(obs.cpu().numpy())) == w["boards"] and du.hexd(du.digest(rew.cpu().numpy())) == w["rewards"], f"step {t}"
        assert du.hexd(du.digest(done.cpu().numpy().astype(np.uint8))) == w["done"]
        assert du.hexd(du.digest(info["score"].cpu().numpy())) == w["scores"]
