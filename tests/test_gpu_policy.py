"""GPU suite for the scripted opponents / build_actions row: pulse_poker_policy and the fused
packed step against the reference-generated golden vectors (tests/golden/poker_policy.npz) and the oracle."""
import numpy as np
import pytest
import torch

from oracle import pyoracle
from tests import poker_util as pu

pytestmark = pytest.mark.gpu
KINDS = {"heuristic_hands": 2, "tight_aggressive": 3, "loose_passive": 4, "small_ball": 5}
Z = np.load(pu.GOLDEN / "poker_policy.npz")


@pytest.mark.parametrize("name", sorted(KINDS))
def test_each_scripted_agent_matches_reference(name, cuda_device):
    from pulselib_b200 import policy

    obs = torch.from_numpy(Z["direct/obs"]).to(cuda_device)
    n = obs.shape[0]
    acts = torch.full((n,), -7, dtype=torch.int64, device=cuda_device)
    policy.scripted_actions(obs, torch.zeros(n, dtype=torch.int32, device=cuda_device), bytes([KINDS[name]]), acts,
                            chance2=torch.from_numpy(Z["direct/chance2"].astype(np.int64)))
    np.testing.assert_array_equal(acts.cpu().numpy(), Z[f"direct/{name}"])


@pytest.mark.parametrize("tag,names", [
    ("mixed6", ["qlearning", "random", "heuristic_hands", "tight_aggressive", "loose_passive", "small_ball"]),
    ("hu", ["qlearning", "tight_aggressive"]),
    ("scripted4", ["small_ball", "small_ball", "random", "loose_passive"])])
def test_build_actions_drop_in_matches_reference(tag, names, cuda_device):
    """Same call as the reference's build_actions(state, actions, curr_players, agents, agent_types, device)."""
    from pulselib_b200 import policy

    class FakeQ:  # as in make_policy_golden.py
        def get_actions(self, states):
            return (states[:, 9].long() + states[:, 5].long()) % 13

    state = torch.from_numpy(Z[f"{tag}/obs"]).to(cuda_device)
    seat = torch.from_numpy(Z[f"{tag}/seat_idx"]).to(cuda_device)
    actions = torch.zeros(state.shape[0], dtype=torch.long, device=cuda_device)
    agents = [FakeQ() if x == "qlearning" else None for x in names]
    policy.build_actions(state, actions, seat, agents, names, cuda_device,
                         chance2=torch.from_numpy(Z[f"{tag}/chance2"].astype(np.int64)))
    np.testing.assert_array_equal(actions.cpu().numpy(), Z[f"{tag}/actions"])


def test_policy_bad_inputs_raise(cuda_device):
    from pulselib_b200 import policy
    from pulselib_b200._lib import PulseError

    obs = torch.zeros((8, 16), device=cuda_device)
    seat = torch.zeros(8, dtype=torch.int32, device=cuda_device)
    with pytest.raises(ValueError):
        policy.scripted_actions(obs, seat, bytes([1]), torch.zeros(7, dtype=torch.int64, device=cuda_device))
    with pytest.raises(PulseError):
        policy.scripted_actions(obs, seat, bytes([9]), torch.zeros(8, dtype=torch.int64, device=cuda_device))
    with pytest.raises(PulseError):
        policy.scripted_actions(torch.zeros(8, 16), seat, bytes([1]), torch.zeros(8, dtype=torch.int64))


@pytest.mark.parametrize("P,kinds,N", [(2, [3, 5], 200_003), (6, [0, 1, 2, 3, 4, 5], 50_001), (10, [4] * 10, 20_000)])
def test_fused_packed_step_equals_oracle_policy_then_step(P, kinds, N, cuda_device, hr_table):
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=P, max_players=P, n_games=N, starting_bbs=60, max_bbs=300, w1=.5, w2=.3, K=100, alpha=50)
    env = PokerGPUPacked(cuda_device, seed=5, env_base=1000, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    rng = np.random.default_rng(P)
    kb = bytes(kinds)
    step = 0
    for hand in range(2):
        obs, _ = env.reset()
        orc.tape_decks(5, 1000, hand)
        np.testing.assert_array_equal(obs.cpu().numpy(), orc.reset(None))
        for t in range(12 if P == 2 else 30):
            ext = rng.integers(0, 13, N).astype(np.int64)
            inject = t % 2 == 0
            chance = rng.integers(0, 2 ** 32, (N, 2), dtype=np.uint64).astype(np.uint32) if inject else \
                pyoracle.policy_tape(5, 1000, step, N)
            want = pyoracle.poker_policy(orc.obs, orc.idx, np.frombuffer(kb, np.uint8), chance, ext.copy())
            seat_before = orc.idx.copy()
            o2, r2, d2 = orc.step(want)
            o1, r1, d1, _, info = env.step_policy(kb, actions=torch.from_numpy(ext).to(cuda_device),
                                                  chance2=torch.from_numpy(chance.astype(np.int64)) if inject else None,
                                                  counter=step, return_seat=True)
            np.testing.assert_array_equal(info["seat_idx_before"].cpu().numpy(), seat_before)
            tag = f"P{P} hand {hand} step {t}"
            np.testing.assert_array_equal(info["actions"].cpu().numpy(), want, err_msg=tag + " played actions")
            np.testing.assert_array_equal(o1.cpu().numpy(), o2, err_msg=tag + " obs")
            np.testing.assert_array_equal(d1.cpu().numpy(), d2, err_msg=tag + " done")
            np.testing.assert_allclose(r1.cpu().numpy(), r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50), err_msg=tag)
            step += 1
        snap = env.unpack()
        for f in pu.INT_N + pu.INT_NP + ("board", "hands"):
            assert np.array_equal(snap[f].cpu().numpy(), getattr(orc, f)), f"hand {hand} field {f}"


@pytest.mark.parametrize("P,auto", [(2, True), (6, False), (6, True)])
def test_rollout_equals_k_policy_steps(P, auto, cuda_device):
    """pulse_poker_packed_rollout: K steps in one launch == K step_policy() calls with the same tape counters"""
    from pulselib_b200 import policy
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=P, max_players=P, n_games=256, starting_bbs=100, max_bbs=1000, w1=.5, w2=.3, K=100, alpha=50,
              seed=11, env_base=5, auto_reset=auto)
    kinds = bytes(([policy.TIGHT_AGGRESSIVE, policy.RANDOM, policy.SMALL_BALL, policy.LOOSE_PASSIVE, policy.HEURISTIC_HANDS,
                    policy.RANDOM])[:P])
    a, b = PokerGPUPacked(cuda_device, **kw), PokerGPUPacked(cuda_device, **kw)
    a.reset(); b.reset()
    K = 40
    obs, rew, done, acts = a.rollout_policy(kinds, K, counter=100)
    for k in range(K):
        o, r, d, _, info = b.step_policy(kinds, counter=100 + k)
        np.testing.assert_array_equal(acts[k].cpu().numpy(), info["actions"].cpu().numpy(), err_msg=f"step {k} actions")
        np.testing.assert_array_equal(rew[k].cpu().numpy(), r.cpu().numpy(), err_msg=f"step {k} rewards")
        np.testing.assert_array_equal(done[k].cpu().numpy(), d.cpu().numpy(), err_msg=f"step {k} dones")
    np.testing.assert_array_equal(obs.cpu().numpy(), o.cpu().numpy())
    sa, sb = a.unpack(), b.unpack()
    for f in sa:
        np.testing.assert_array_equal(sa[f].cpu().numpy(), sb[f].cpu().numpy(), err_msg=f)


@pytest.mark.parametrize("P,N", [(2, 10_007), (6, 3_001), (10, 1_025)])
def test_rollout_on_supplied_actions_equals_k_steps(P, N, cuda_device):
    """pulse_poker_packed_rollout_actions: the table is loaded once, stepped K times in registers and stored once; rewards,
    done flags, the final observation and the whole final state must equal K single-step launches on the same actions
    (negative and oversized action ids included)."""
    from pulselib_b200.envs.poker_packed import PokerGPUPacked

    kw = dict(n_players=P, max_players=P, n_games=N, starting_bbs=60, max_bbs=400, w1=.5, w2=.3, K=100, alpha=50, seed=9,
              env_base=123, track_prev_invested=True)
    a, b = PokerGPUPacked(cuda_device, **kw), PokerGPUPacked(cuda_device, **kw)
    K = 4 * P + 3
    g = torch.Generator(device=cuda_device).manual_seed(P)
    for hand in range(2):
        a.reset(); b.reset()
        acts = torch.randint(-1, 15, (K, N), generator=g, device=cuda_device)
        obs, rew, done = a.rollout_actions(acts)
        for k in range(K):
            o2, r2, d2, _, _ = b.step(acts[k])
            assert torch.equal(rew[k], r2), f"hand {hand} step {k}: rewards"
            assert torch.equal(done[k], d2), f"hand {hand} step {k}: done flags"
        assert torch.equal(obs, o2)
        sa, sb = a.unpack(), b.unpack()
        for f in sa:
            assert torch.equal(sa[f], sb[f]), f"hand {hand}: state field {f}"
        assert torch.equal(a._prev_inv, b._prev_inv) and torch.equal(a._idx, b._idx)
        assert torch.equal(a._ev[:, :], b._ev[:, :])
    assert bool(done.any()) and not bool(done[0].all())
