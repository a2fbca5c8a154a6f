"""GPU suite for the hand-level statistics row: the device accumulator against the reference's own
metric numbers (golden trace) and against the numpy restatement on live packed-engine rollouts."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import pyoracle
from tests import poker_util as pu
from tests.test_handstats_cpu import Z, check_summary, replay

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["six", "hu"])
def test_device_hand_stats_match_reference_metrics(tag, cuda_device):
    from pulselib_b200 import _lib
    from pulselib_b200._lib import c_i, ptr

    lib = _lib.load()
    dev = cuda_device
    N, P = int(Z[f"{tag}/meta"][0]), int(Z[f"{tag}/meta"][1])
    vec = torch.zeros(67, dtype=torch.float64, device=dev)
    want = np.zeros(67)
    term_dev = torch.zeros(N, dtype=torch.bool, device=dev)

    def update(stacks, initial, stages, button, done, terminated, q_seat, AP, first):
        if first:
            term_dev.zero_()
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (stacks, stages, button, done, initial)]
        with _lib.device_guard(dev):
            _lib.check(lib.pulse_poker_hand_stats(ptr(t[0]), c_i(P), ptr(t[1]), ptr(t[2]), ptr(t[3]), ptr(term_dev), ptr(t[4]),
                                                  c_i(q_seat), c_i(AP), ptr(vec), C.c_longlong(N), _lib.cuda_stream(dev)))
        pyoracle.hand_stats_update(want, stacks[:, q_seat], initial, stages, button, done, terminated, q_seat, AP)
        np.testing.assert_array_equal(term_dev.cpu().numpy(), terminated)

    AP = replay(tag, update)
    np.testing.assert_array_equal(vec.cpu().numpy(), want)     # integer-valued sums: exact, whatever the atomic order
    check_summary(tag, vec.cpu().numpy(), AP)


@pytest.mark.parametrize("P,q_seat", [(2, 1), (6, 3)])
def test_hand_stats_on_packed_rollout(P, q_seat, cuda_device):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from pulselib_b200.hand_stats import HandStats

    N = 100_003
    env = PokerGPUPacked(cuda_device, n_players=P, max_players=P, n_games=N, starting_bbs=100, seed=9, env_base=4)
    hs = HandStats(env, q_seat=q_seat)
    want = np.zeros(67)
    gen = torch.Generator(device=cuda_device)
    gen.manual_seed(3)
    for hand in range(3):
        env.reset()
        hs.begin_hand()
        initial = env.unpack()["stacks"][:, q_seat].cpu().numpy()
        terminated = np.zeros(N, bool)
        for t in range(14 if P == 2 else 26):
            a = torch.randint(0, 13, (N,), dtype=torch.int64, device=cuda_device, generator=gen)
            _, _, done, _, _ = env.step(a)
            hs.update(done)
            snap = env.unpack()
            pyoracle.hand_stats_update(want, snap["stacks"][:, q_seat].cpu().numpy(), initial, snap["stages"].cpu().numpy(),
                                       snap["button"].cpu().numpy(), done.cpu().numpy(), terminated, q_seat, P)
    np.testing.assert_array_equal(hs.vec.cpu().numpy(), want)
    s = hs.summary()
    assert s["hands_completed"] == int(want[0]) > 2 * N
    assert sum(m["hands"] for m in s["by_position"].values()) == s["hands_completed"]
    assert sum(m["hands"] for m in s["by_stage"].values()) == s["hands_completed"]


def test_hand_stats_rearm_with_auto_reset(cuda_device):
    """Per-table auto-reset: every done flag is one finished hand, measured from that hand's own starting stack
    (the stack the deal left in the seat, refilled when busted or above max_bbs)."""
    from pulselib_b200.envs.poker_packed import PokerGPUPacked
    from pulselib_b200.hand_stats import HandStats

    N, P, q = 60_001, 2, 0
    env = PokerGPUPacked(cuda_device, n_players=P, max_players=P, n_games=N, starting_bbs=30, max_bbs=45, seed=2,
                         env_base=8, auto_reset=True)
    hs = HandStats(env, q_seat=q)
    env.reset()
    hs.begin_hand()
    initial = env.unpack()["stacks"][:, q].cpu().numpy().astype(np.int64)
    want = np.zeros(67)
    gen = torch.Generator(device=cuda_device)
    gen.manual_seed(5)
    refills = 0
    for t in range(40):
        a = torch.randint(0, 13, (N,), dtype=torch.int64, device=cuda_device, generator=gen)
        _, _, done, _, _ = env.step(a)
        hs.update(done)
        snap = env.unpack()
        d = done.cpu().numpy()
        fin = snap["stacks"][:, q].cpu().numpy().astype(np.int64)
        pyoracle.hand_stats_update(want, fin, initial, snap["stages"].cpu().numpy(), snap["button"].cpu().numpy(), d,
                                   np.zeros(N, bool), q, P)
        refill = (fin == 0) | (fin > 45)
        refills += int((d & refill).sum())
        initial = np.where(d, np.where(refill, 30, fin), initial)
    np.testing.assert_array_equal(hs.vec.cpu().numpy(), want)
    assert want[0] > 3 * N and refills > 0
