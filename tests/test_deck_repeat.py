"""No (table, deck) pair may repeat when whole-batch reset() calls are mixed with per-table auto_reset redeals.

The chance tape deals the deck of (seed, global table id, hand index) -- DESIGN.md section 3.  A whole-batch reset
deals hand max(hand_ctr[table] + 1, hand_counter), a per-table redeal hand_ctr[table] + 1, so the hand index of a
table only ever grows and a trajectory never sees the same shuffle twice (the reference draws a fresh torch.rand
shuffle per reset, /root/reference/environments/Poker/PokerGPU.py:84-90; a repeated deck would be a correlation it
cannot have).  CPU: the kernel bodies on the host (tests/hostemu); GPU: the product class."""
import numpy as np
import pytest


class _Ledger:
    """every deck a table has been dealt, keyed by table; a new deal is recognised by the table's hand index moving"""

    def __init__(self, n):
        self.seen = [set() for _ in range(n)]
        self.ctr = np.full(n, -1, np.int64)
        self.last = [None] * n
        self.deals = 0

    def note(self, hand_ctr, decks, tag):
        hand_ctr = np.asarray(hand_ctr).astype(np.int64)
        decks = np.asarray(decks)
        moved = np.nonzero(hand_ctr != self.ctr)[0]
        assert np.all(hand_ctr[moved] > self.ctr[moved]), f"{tag}: a table's hand index went backwards"
        same = np.nonzero(hand_ctr == self.ctr)[0]
        for g in same[:: max(1, len(same) // 64)]:            # a sample: the deck of an undealt table stays put
            assert decks[g].tobytes() == self.last[g], f"{tag}: table {g}'s deck changed without a deal"
        for g in moved:
            d = decks[g].tobytes()
            assert sorted(decks[g].tolist()) == list(range(1, 53)), f"{tag}: table {g} holds no permutation of 52 cards"
            assert d not in self.seen[g], f"{tag}: table {g} was dealt a deck it has already played (hand {hand_ctr[g]})"
            self.seen[g].add(d)
            self.last[g] = d
            self.deals += 1
        self.ctr = hand_ctr.copy()
        return len(moved)


def _drive(n, steps, reset_every, reset_fn, step_fn, seed):
    rng = np.random.default_rng(seed)
    led = _Ledger(n)
    resets = redeals = 0
    for t in range(steps):
        if t % reset_every == 0:
            ctr, decks = reset_fn(resets)
            assert led.note(ctr, decks, f"reset {resets}") == n      # a whole-batch reset deals every table
            resets += 1
        acts = rng.integers(0, 13, n)
        acts[rng.random(n) < 0.35] = 0                                # plenty of folds: hands end, tables redeal
        ctr, decks = step_fn(acts)
        redeals += led.note(ctr, decks, f"step {t}")
    assert redeals > n, "the run must contain per-table redeals between the whole-batch resets"
    assert led.deals == resets * n + redeals
    assert min(len(s) for s in led.seen) >= resets
    return led


def test_no_deck_repeats_across_reset_and_auto_reset_host_bodies(hr_table):
    from tests.hostemu.emu import EmuPokerPacked

    n = 256
    env = EmuPokerPacked(hr_table, auto_reset=True, seed=77, env_base=1000, n_players=2, max_players=2, n_games=n)

    def reset(k):
        env.reset(hand=k, rotation=k % 2)
        return env.hand_ctr.copy(), env.decks.copy()

    def step(a):
        env.step(a)
        return env.hand_ctr.copy(), env.decks.copy()

    _drive(n, 60, 11, reset, step, 1)


@pytest.mark.gpu
@pytest.mark.parametrize("P", [2, 6])
def test_no_deck_repeats_across_reset_and_auto_reset_gpu(P, cuda_device):
    import torch

    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    n = 4096
    env = PokerGPUPacked(cuda_device, n_players=P, max_players=P, n_games=n, auto_reset=True, seed=77)
    mirror = env.alloc_mirror(fields=("decks",), full=True)

    def snap():
        env.unpack_into(mirror, full=True)
        torch.cuda.synchronize(cuda_device)
        return env._hand_ctr.cpu().numpy(), mirror["decks"].cpu().numpy()

    def reset(k):
        env.reset(options={"rotation": k % P})
        return snap()

    def step(a):
        env.step(torch.from_numpy(a).to(cuda_device))
        return snap()

    _drive(n, 90, 17, reset, step, 2)
