"""CPU suite for the scripted opponents / build_actions row: the oracle and the kernel
rule (compiled for the host) against golden vectors produced by the reference's own Player.py agents and
utils.build_actions (tests/golden/make_policy_golden.py), and the fused packed step against step-after-policy."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import poker_util as pu
from tests.hostemu.emu import EmuPokerPacked, emu_policy

KINDS = {"heuristic_hands": 2, "tight_aggressive": 3, "loose_passive": 4, "small_ball": 5}
Z = np.load(pu.GOLDEN / "poker_policy.npz")


@pytest.mark.parametrize("engine", ["oracle", "kernel_body"])
@pytest.mark.parametrize("name", sorted(KINDS))
def test_each_scripted_agent_matches_reference(name, engine):
    obs, chance = Z["direct/obs"], Z["direct/chance2"]
    n = obs.shape[0]
    seat = np.zeros(n, np.int32)
    acts = np.full(n, -7, np.int64)
    kinds = np.array([KINDS[name]], np.uint8)
    if engine == "oracle":
        pyoracle.poker_policy(obs, seat, kinds, chance, acts)
    else:
        emu_policy(obs, seat, kinds, acts, chance2=chance)
    np.testing.assert_array_equal(acts, Z[f"direct/{name}"])


@pytest.mark.parametrize("engine", ["oracle", "kernel_body"])
@pytest.mark.parametrize("tag", ["mixed6", "hu", "scripted4"])
def test_build_actions_matches_reference(tag, engine):
    obs, chance, seat, kinds = (Z[f"{tag}/{k}"] for k in ("obs", "chance2", "seat_idx", "kinds"))
    acts = np.zeros(obs.shape[0], np.int64)
    ext = kinds[seat] == 0
    acts[ext] = Z[f"{tag}/q_actions"][ext]            # what the Q agent's get_actions put there
    if engine == "oracle":
        pyoracle.poker_policy(obs, seat, kinds, chance, acts)
    else:
        emu_policy(obs, seat, kinds, acts, chance2=chance)
    np.testing.assert_array_equal(acts, Z[f"{tag}/actions"])


def test_tape_draws_agree_between_oracle_and_kernel_body():
    obs, seat, kinds = Z["mixed6/obs"], Z["mixed6/seat_idx"], Z["mixed6/kinds"].copy()
    kinds[0] = 1                                         # no external seat: every action comes from the tape
    n = obs.shape[0]
    a = emu_policy(obs, seat, kinds, np.zeros(n, np.int64), chance2=None, seed=77, env_base=1 << 40, counter=9)
    b = pyoracle.poker_policy(obs, seat, kinds, pyoracle.policy_tape(77, 1 << 40, 9, n), np.zeros(n, np.int64))
    np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("P,kinds", [(2, [3, 5]), (6, [0, 1, 2, 3, 4, 5]), (4, [5, 5, 1, 4])])
def test_fused_packed_step_equals_policy_then_step(P, kinds, hr_table):
    """One launch (rule evaluated inside the step on the packed state) == the oracle's policy on the observation
    followed by the oracle's step, for whole hands."""
    N = 3000
    kw = dict(n_players=P, max_players=P, n_games=N, starting_bbs=60, max_bbs=300, w1=.5, w2=.3, K=100, alpha=50)
    emu = EmuPokerPacked(hr_table, seed=5, env_base=1000, **kw)
    orc = pyoracle.OraclePoker(hr_table, **kw)
    rng = np.random.default_rng(P)
    kinds = np.array(kinds, np.uint8)
    step = 0
    for hand in range(3):
        decks = (np.argsort(rng.random((N, 52)), axis=1) + 1).astype(np.int32)
        np.testing.assert_array_equal(emu.reset(decks, rotation=hand % P), orc.reset(decks, rotation=hand % P))
        for t in range(12 if P == 2 else 30):
            ext = rng.integers(0, 13, N).astype(np.int64)            # the external (Q) seat's actions
            inject = t % 2 == 0
            chance = rng.integers(0, 2 ** 32, (N, 2), dtype=np.uint64).astype(np.uint32) if inject else \
                pyoracle.policy_tape(5, 1000, step, N)
            want = pyoracle.poker_policy(orc.obs, orc.idx, kinds, chance, ext.copy())
            o2, r2, d2 = orc.step(want)
            o1, r1, d1, chosen = emu.step_policy(kinds, chance2=chance if inject else None, counter=step, actions=ext)
            tag = f"hand {hand} step {t}"
            np.testing.assert_array_equal(chosen, want, err_msg=tag + " played actions")
            np.testing.assert_array_equal(o1, o2, err_msg=tag + " obs")
            np.testing.assert_array_equal(d1, d2, err_msg=tag + " done")
            np.testing.assert_allclose(r1, r2, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(50), err_msg=tag)
            pu.compare_state(emu, orc, tag)
            step += 1
