"""GPU parity of the packed-layout PokerGPU (the throughput path) against the oracle: full integer
state via unpack(), obs/done bit-exact, reward within the fp32 tanh tolerance; plus size-independent
properties at the benchmark's table count."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu

pytestmark = pytest.mark.gpu

FIELDS = pu.INT_N + pu.INT_NP + ("board", "hands")


class PackedAdapter(pu.TorchAdapter):
    def __getattr__(self, name):
        if name == "equities":
            raise AttributeError(name)
        return getattr(self.env, name)


def rollout(a, b, rng, P, N, hands, steps, vary, rot):
    import torch

    for h in range(hands):
        ap = int(rng.integers(2, P + 1)) if vary else P
        decks = (np.argsort(rng.random((N, 52)), axis=1) + 1).astype(np.int32)
        r = (rot * h) % P
        oa = a.reset(decks, active_players=ap, rotation=r)
        ob = b.reset(decks, active_players=ap, rotation=r)
        np.testing.assert_array_equal(pu.to_np(oa), ob)
        for t in range(steps):
            acts = rng.integers(0, 13, N).astype(np.int64)
            oa, ra, da = a.step(acts)
            ob, rb, db = b.step(acts)
            snap = a.env.unpack()
            for f in FIELDS:
                x = snap[f].cpu().numpy()
                assert np.array_equal(x, getattr(b, f)), f"hand {h} step {t} field {f}"
            np.testing.assert_array_equal(pu.to_np(da), db)
            np.testing.assert_array_equal(pu.to_np(oa), ob)
            np.testing.assert_allclose(pu.to_np(ra), rb, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50))
            if db.all():
                break


@pytest.mark.parametrize("P,M,N,vary,rot", [(2, 2, 20011, False, 0), (6, 6, 6007, False, 1), (10, 10, 3001, True, 3),
                                              (3, 10, 3001, True, 2)])
def test_packed_matches_oracle(P, M, N, vary, rot, cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=P, max_players=M, n_games=N, starting_bbs=50, max_bbs=200, w1=.5, w2=.3, K=100, alpha=50)
    rollout(PackedAdapter(PokerGPUPacked(cuda_device, **kw)), pyoracle.OraclePoker(hr_table, **kw),
            np.random.default_rng(P + M), P, N, 4, 45, vary, rot)


def test_packed_tape_reset_matches_oracle(cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=6, max_players=6, n_games=5000, starting_bbs=100)
    env = PokerGPUPacked(cuda_device, seed=20260401, env_base=99, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    for hand in range(2):
        obs, _ = env.reset()
        orc.tape_decks(20260401, 99, hand)
        want = orc.reset(None)
        np.testing.assert_array_equal(obs.cpu().numpy(), want)
        np.testing.assert_array_equal(env.unpack()["hands"].cpu().numpy(), orc.hands)


def test_full_size_properties(cuda_device):
    """4M heads-up tables (the benchmark size): chips are conserved per table, every table is done
    after 16 random steps, done tables are no-ops, and the same seed gives the same trajectory."""
    import torch
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    N = 4_194_304
    kw = dict(n_players=2, max_players=2, n_games=N, starting_bbs=100, w1=.5, w2=.3, K=100, alpha=50)
    sums = []
    for rep in range(2):
        env = PokerGPUPacked(cuda_device, seed=7, **kw)
        env.reset()
        gen = torch.Generator(device=cuda_device)
        gen.manual_seed(1)
        chk = torch.zeros((), dtype=torch.float64, device=cuda_device)
        for t in range(16):
            a = torch.randint(0, 13, (N,), dtype=torch.int64, device=cuda_device, generator=gen)
            obs, rew, done, _, _ = env.step(a)
            chk += obs.double().sum() * (t + 1) + rew.double().sum()
            snap = env.unpack()
            total = snap["stacks"].sum(dim=1) + snap["pots"]
            lost = (snap["is_done"] & (total != 200))
            # chips only vanish through the reference's quirks (pot unpaid when nobody is live,
            # unclaimed side-pot layers): never created
            assert int((total > 200).sum()) == 0
            if t == 15:
                assert float(done.float().mean()) > 0.999  # heads-up hands rarely outlast 16 random steps
                assert float(lost.float().mean()) < 0.2
        before = env.unpack()
        obs2, rew2, done2, _, _ = env.step(torch.zeros(N, dtype=torch.int64, device=cuda_device))
        after = env.unpack()
        was_done = before["is_done"].bool()                 # tables still live may legitimately change
        for k in ("stacks", "pots", "stages", "idx", "board"):
            assert torch.equal(before[k][was_done], after[k][was_done]), k
        assert float(rew2[was_done].abs().max()) == 0.0
        sums.append(float(chk))
    assert sums[0] == sums[1]


def test_auto_reset_deals_next_hand(cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    import torch

    N = 4096
    kw = dict(n_players=2, max_players=2, n_games=N, starting_bbs=100)
    env = PokerGPUPacked(cuda_device, seed=11, env_base=5, auto_reset=True, **kw)
    env.reset()
    fold = torch.zeros(N, dtype=torch.int64, device=cuda_device)
    _, _, done, _, _ = env.step(fold)            # everyone folds: all tables done
    assert bool(done.all())
    obs, rew, done, _, _ = env.step(fold)        # auto-reset step: new hand, action ignored
    assert not bool(done.any()) and float(rew.abs().max()) == 0.0
    snap = env.unpack()
    assert int(snap["button"].min()) == 1 and int(snap["stages"].max()) == 0
    orc = pyoracle.OraclePoker(hr_table, **kw)
    orc.tape_decks(11, 5, 1)                     # hand counter 1 of each table's tape
    np.testing.assert_array_equal(snap["hands"].cpu().numpy()[:, :, :], orc.decks[:, :4].reshape(N, 2, 2))


@pytest.mark.parametrize("P,N", [(2, 30_011), (6, 8_009)])
def test_auto_reset_rollout_matches_inline_semantics(P, N, cuda_device, hr_table):
    """auto_reset on the device is two launches (the step marks finished tables, the redeal pass deals them);
    the kernel body compiled for the host deals inline.  Both must give the same trajectories: state, obs,
    reward, done, step after step, with tables finishing and redealing at different times."""
    import torch
    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from tests.hostemu.emu import EmuPokerPacked

    kw = dict(n_players=P, max_players=P, n_games=N, starting_bbs=40, max_bbs=150, w1=.5, w2=.3, K=100, alpha=50)
    env = PokerGPUPacked(cuda_device, seed=3, env_base=17, auto_reset=True, **kw)
    emu = EmuPokerPacked(hr_table, auto_reset=True, seed=3, env_base=17, **kw)
    rng = np.random.default_rng(P)
    np.testing.assert_array_equal(env.reset()[0].cpu().numpy(), emu.reset(None, hand=0))
    redeals = 0
    for t in range(60):
        a = rng.integers(0, 13, N).astype(np.int64)
        was_done = emu.is_done.astype(bool).copy()
        o1, r1, d1, _, _ = env.step(torch.from_numpy(a).to(cuda_device))
        o2, r2, d2 = emu.step(a)
        redeals += int(was_done.sum())
        np.testing.assert_array_equal(o1.cpu().numpy(), o2, err_msg=f"step {t} obs")
        np.testing.assert_array_equal(d1.cpu().numpy(), d2, err_msg=f"step {t} done")
        np.testing.assert_allclose(r1.cpu().numpy(), r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50))
        if t % 10 == 9:
            snap = env.unpack()
            for f in FIELDS:
                assert np.array_equal(snap[f].cpu().numpy(), getattr(emu, f)), f"step {t} field {f}"
    assert redeals > N  # every table was redealt more than once on average
