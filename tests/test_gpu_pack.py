"""GPU suite for pulse_poker_pack: reference-shaped tensors -> packed layout, and the drop-in class going
back onto the throughput layout after white-box edits (PokerGPU.repack)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(dev, P, N, layout="packed"):
    from pulselib_b200.envs.poker import PokerGPU

    return PokerGPU(dev, [], n_players=P, max_players=P, n_games=N, starting_bbs=40, max_bbs=200, w1=.5, w2=.3, K=100,
                    alpha=50, seed=4, env_base=50, layout=layout)


@pytest.mark.parametrize("P,N", [(2, 5003), (6, 2001), (10, 777)])
def test_unpack_then_pack_is_the_identity_on_device(P, N, cuda_device):
    """the packed engine's arrays after unpack_full -> pulse_poker_pack, on tables spread over every street"""
    from pulselib_b200.envs.poker import _State, _STATE_FIELDS
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    env = PokerGPUPacked(cuda_device, n_players=P, max_players=P, n_games=N, starting_bbs=40, max_bbs=200, seed=4,
                         env_base=50, track_prev_invested=True)
    env.reset(options={"rotation": 1})
    g = torch.Generator(device=cuda_device).manual_seed(P)
    for t in range(3 * P + 2):
        a = torch.randint(0, 13, (N,), generator=g, device=cuda_device)
        a[torch.rand(N, generator=g, device=cuda_device) < 0.6] = 1
        env.step(a)
        if t % P:
            continue
        mirror = env.unpack_into(env.alloc_mirror(None, full=True), full=True)
        want = {n: getattr(env, n).clone() for n in ("_tbl", "_tbl2", "_stack", "_crb", "_inv", "_hole", "_idx", "_prev_inv")}
        want_ev = env._ev.clone()
        for n in want:
            getattr(env, n).fill_(0x5A)
        env._ev.zero_()
        s = _State()
        for name in _STATE_FIELDS:
            tns = env.hand_ranks if name == "hand_ranks" else mirror.get(name)
            setattr(s, name, None if tns is None else tns.data_ptr())
        s.n_hand_ranks = env.hand_ranks.numel()
        assert env.pack_from(s) == 0
        for n in want:
            assert torch.equal(getattr(env, n), want[n]), f"step {t}: {n}"
        assert torch.equal(env._ev[: env.active_players], want_ev[: env.active_players]), f"step {t}: hand-value words"


@pytest.mark.parametrize("P", [2, 6])
def test_dropin_repack_after_white_box_edit(P, cuda_device):
    """edit through the reference's attributes -> reference-shaped mode; repack() -> packed mode again, and from there
    the env returns exactly what an env that stayed reference-shaped returns"""
    N = 4096
    a, b = _mk(cuda_device, P, N), _mk(cuda_device, P, N, layout="compat")
    rng = np.random.default_rng(P)
    a.reset(); b.reset()

    def both(k, tag):
        for t in range(k):
            act = torch.from_numpy(rng.integers(0, 13, N)).to(cuda_device)
            oa, ra, da, _, _ = a.step(act)
            ob, rb, db, _, _ = b.step(act)
            assert torch.equal(oa, ob), f"{tag} step {t}: obs"
            assert torch.equal(da, db), f"{tag} step {t}: done"
            torch.testing.assert_close(ra, rb, rtol=2e-6, atol=50 * 4e-7, msg=f"{tag} step {t}: reward")

    both(2, "before")
    assert a._mode == "packed"
    for e in (a, b):
        e.stacks[::7, 0] += 9           # the reference's tests do exactly this
        e.pots[::5] += 3
    both(1, "edited")
    assert a._mode == "compat"
    assert a.repack() is True and a._mode == "packed"
    both(3 * P, "repacked")
    for f in ("stacks", "pots", "stages", "status", "board", "hands", "decks", "deck_positions", "total_invested"):
        assert torch.equal(getattr(a, f).to(torch.int64), getattr(b, f).to(torch.int64)), f
    assert a._mode == "packed"          # reading the mirror does not leave the packed layout
    # a state the packed layout cannot hold: repack declines and nothing changes
    a.deck_positions[0] += 1
    b.deck_positions[0] += 1
    both(1, "unrepresentable")
    assert a._mode == "compat" and a.repack() is False and a._mode == "compat"
    both(2, "still reference-shaped")
