"""Replays of the traces tests/golden/record_reference_loops.py recorded from the reference's own training loop
(/root/reference/scripts/Poker/trainGPU.py:37-110) and benchmark cases (/root/reference/benchmarking/Poker/cases.py:66-170).
Shared by the CPU suite (product classes on the host-compiled kernel bodies) and the GPU suite (the B200)."""
from pathlib import Path
from unittest import mock

import numpy as np
import torch

from tests import poker_util as pu

GOLDEN = Path(__file__).resolve().parent / "golden" / "poker_loops.npz"


def load():
    return np.load(GOLDEN)


def calls_of(z, prefix):
    n = int(z[f"{prefix}/n_calls"])
    kinds = [str(k) for k in z[f"{prefix}/kinds"]]
    out = []
    for i in range(n):
        pre = f"{prefix}/c{i}/"
        out.append(dict(kind=kinds[i], **{k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}))
    return out


def env_kwargs(z, prefix):
    P, M, N = (int(v) for v in z[f"{prefix}/meta"][:3])
    sb, mb, K, alpha = (int(v) for v in z[f"{prefix}/cfg"])
    return dict(n_players=P, n_games=N, starting_bbs=sb, w1=float(z[f"{prefix}/w"][0]), w2=float(z[f"{prefix}/w"][1]),
                K=K, alpha=alpha)


def t2n(t):
    return t.detach().cpu().numpy()


def do_reset(env, c):
    """the trainer's / benchmark's reset call, with the recorded table-size draw standing in for torch.randint (:76-80)"""
    options = {"rotation": int(c["rotation"]), "active_players": bool(int(c["active_players_flag"])),
               "q_agent_seat": int(c["q_agent_seat"])}
    draw = int(c["ap_draw"])
    if draw >= 0:
        with mock.patch("torch.randint", lambda *a, **k: torch.tensor([draw])):
            return env.reset(options=options)
    return env.reset(options=options)


def check_reset(env, c, obs, info, tag):
    np.testing.assert_array_equal(t2n(obs), c["obs"], err_msg=tag + " obs")
    assert int(info["active_players"]) == int(c["active_players"]), tag + " active_players"
    np.testing.assert_array_equal(t2n(info["seat_idx"]).astype(np.int64), c["seat_idx"].astype(np.int64), err_msg=tag + " seat_idx")
    np.testing.assert_array_equal(t2n(info["stacks"]).astype(np.int64), c["stacks"].astype(np.int64), err_msg=tag + " stacks")
    np.testing.assert_array_equal(t2n(env.button).astype(np.int64), c["button"].astype(np.int64), err_msg=tag + " button")
    np.testing.assert_array_equal(t2n(env.decks).astype(np.int64), c["decks"].astype(np.int64), err_msg=tag + " decks")


def check_step(c, obs, rew, done, info, alpha, tag):
    np.testing.assert_array_equal(t2n(obs), c["obs"], err_msg=tag + " obs")
    np.testing.assert_array_equal(t2n(done).astype(np.uint8), c["dones"], err_msg=tag + " dones")
    np.testing.assert_allclose(t2n(rew), c["rewards"], rtol=pu.REWARD_RTOL, atol=pu.reward_atol(alpha), err_msg=tag + " rewards")
    np.testing.assert_array_equal(t2n(info["seat_idx"]).astype(np.int64), c["seat_idx"].astype(np.int64), err_msg=tag + " seat_idx")
    np.testing.assert_array_equal(t2n(info["stacks"]).astype(np.int64), c["stacks"].astype(np.int64), err_msg=tag + " stacks")
    assert int(info["active_players"]) == int(c["active_players"]), tag


def replay_calls(env, calls, device, alpha, tag):
    """Every recorded env-facing call, in order, answers compared."""
    n = 0
    for i, c in enumerate(calls):
        if c["kind"] == "reset":
            obs, info = do_reset(env, c)
            check_reset(env, c, obs, info, f"{tag} call {i} reset")
        else:
            obs, rew, done, trunc, info = env.step(torch.from_numpy(c["actions"]).to(device))
            assert trunc.dtype == torch.bool and not bool(trunc.any())
            check_step(c, obs, rew, done, info, alpha, f"{tag} call {i} step")
        n += 1
    return n


def run_train_loop_body(env, z, device):
    """The body of train_agent (/root/reference/scripts/Poker/trainGPU.py:56-108) restated against `env`, with the
    actions the reference's own build_actions + agents produced standing in for build_actions (the agents are out of
    scope; the env-facing behaviour of the loop is what is checked): same masks from info["seat_idx"], same
    termination rule, same bookkeeping.  Must reproduce the step count of every episode, the Q agent's transition
    batches, the episode reward and the episode chip profit the reference run ended with."""
    calls = calls_of(z, "train")
    P, M, N, episodes, seed = (int(v) for v in z["train/meta"])
    alpha = int(z["train/cfg"][3])
    actions = torch.zeros(N, dtype=torch.long, device=device)
    q_agent_idx = 0
    batches, reward_scores, scores = [], [], []
    ci = 0
    for episode in range(episodes):
        target_seat = episode % P                                  # get_rotated_agents (utils.py:187-197)
        rotations = (target_seat - q_agent_idx) % P
        q_seat = target_seat
        c = calls[ci]; ci += 1
        assert c["kind"] == "reset" and int(c["rotation"]) == rotations and int(c["q_agent_seat"]) == q_seat
        state, info = do_reset(env, c)
        initial_stacks = info["stacks"][:, q_seat].clone()
        terminated = torch.zeros(N, dtype=torch.bool, device=device)
        episode_reward_tensor = torch.tensor(0.0, device=device)
        termination_threshold = torch.tensor(0.8, device=device)
        idx = 0
        while True:
            actions.fill_(0)
            q_mask = info["seat_idx"] == q_seat
            assert torch.is_tensor(q_mask) and q_mask.dtype == torch.bool and tuple(q_mask.shape) == (N,)
            c = calls[ci]; ci += 1
            assert c["kind"] == "step", "the reference's loop ended this episode earlier"
            actions.copy_(torch.from_numpy(c["actions"]).to(device))       # build_actions (utils.py:108-123)
            next_state, rewards, dones, truncated, info = env.step(actions)
            del truncated
            active_games = q_mask & ~terminated
            terminated |= dones
            if active_games.any():
                batches.append((int(active_games.sum()), float(rewards[active_games].double().sum()),
                                int(dones[active_games].sum()), float(state[active_games].double().sum()),
                                float(next_state[active_games].double().sum()), int(actions[active_games].sum())))
            episode_reward_tensor += rewards[active_games].sum()
            state = next_state
            if idx % 5 == 0 and bool(terminated.float().mean() > termination_threshold):
                break
            idx += 1
        assert ci == len(calls) or calls[ci]["kind"] == "reset", "the reference's loop ran this episode longer"
        final_stacks = info["stacks"][:, q_seat]
        scores.append((final_stacks - initial_stacks).sum().item())
        reward_scores.append(episode_reward_tensor.item())
    assert ci == len(calls)
    np.testing.assert_array_equal(np.array(scores, np.float64), z["train/profit_scores"])
    np.testing.assert_allclose(np.array(reward_scores), z["train/reward_scores"], rtol=1e-4, atol=N * pu.reward_atol(alpha))
    want = z["train/q_batches"]
    got = np.array(batches, np.float64)
    assert got.shape == want.shape
    np.testing.assert_array_equal(got[:, [0, 2, 5]], want[:, [0, 2, 5]])            # counts, dones, action sums
    np.testing.assert_array_equal(got[:, 3:5], want[:, 3:5])                        # observation sums (integers in fp32)
    np.testing.assert_allclose(got[:, 1], want[:, 1], rtol=1e-4, atol=N * pu.reward_atol(alpha))
    return len(calls)


BENCH_STATE = ("deck_positions", "pots", "stages", "button", "sb", "bb", "idx", "highest", "agg", "acted",
               "last_raise_size", "stacks", "current_round_bet", "total_invested", "status", "board", "hands",
               "equity_dirty", "decks")


def check_final_state(env, z, name):
    poker_env = env.unwrapped
    ap = int(z[f"bench/{name}/final/active_players"])
    assert int(poker_env.active_players) == ap
    for f in BENCH_STATE:
        want = z[f"bench/{name}/final/{f}"]
        got = t2n(getattr(poker_env, f))
        assert np.array_equal(got.astype(np.int64).reshape(want.shape), want.astype(np.int64)), f"{name}: final {f} differs"
    np.testing.assert_array_equal(t2n(poker_env.equities)[:, :ap], z[f"bench/{name}/final/equities"][:, :ap],
                                  err_msg=f"{name}: final equities")


def run_bench_cases(make_env, z, device):
    """benchmarking/Poker/cases.py:66-170 restated against make_env(seed) (warmup 1, measure 2, as recorded)."""
    alpha = int(z["bench/cfg"][3])
    N = int(z["bench/meta"][2])
    reset_opts_calls = {n: calls_of(z, f"bench/{n}") for n in ("run_env_reset", "run_env_calculate_equities",
                                                                "run_env_execute_actions", "run_env_step")}
    # run_env_reset (:66-84)
    name = "run_env_reset"
    env = make_env(int(z[f"bench/{name}/seed"]))
    for i, c in enumerate(reset_opts_calls[name]):
        obs, info = do_reset(env, c)
        check_reset(env, c, obs, info, f"{name} {i}")
    check_final_state(env, z, name)
    env.close()

    # run_env_calculate_equities (:87-121)
    name = "run_env_calculate_equities"
    env = make_env(int(z[f"bench/{name}/seed"]))
    calls = reset_opts_calls[name]
    obs, info = do_reset(env, calls[0])
    check_reset(env, calls[0], obs, info, f"{name} 0")
    poker_env = env.unwrapped
    for i, c in enumerate(calls[1:]):
        obs, info = do_reset(env, c)                                   # prepare_stage(3)
        check_reset(env, c, obs, info, f"{name} {i + 1}")
        poker_env.stages.fill_(3)
        poker_env.board[:, 0:3] = poker_env.deal_cards(poker_env.g, 3)
        poker_env.board[:, 3] = poker_env.deal_cards(poker_env.g, 1).squeeze(1)
        poker_env.board[:, 4] = poker_env.deal_cards(poker_env.g, 1).squeeze(1)
        poker_env.equities.fill_(0.5)
        poker_env.calculate_equities()
    check_final_state(env, z, name)
    env.close()

    # run_env_execute_actions (:124-146)
    name = "run_env_execute_actions"
    env = make_env(int(z[f"bench/{name}/seed"]))
    calls = reset_opts_calls[name]
    obs, info = do_reset(env, calls[0])
    poker_env = env.unwrapped
    fixed_actions = torch.ones(N, dtype=torch.long, device=device)
    poker_env.execute_actions(fixed_actions)                           # warmup 1
    for c in calls[1:]:
        obs, info = do_reset(env, c)
        check_reset(env, c, obs, info, name)
        poker_env.execute_actions(fixed_actions)
    check_final_state(env, z, name)
    env.close()

    # run_env_step (:149-170)
    name = "run_env_step"
    env = make_env(int(z[f"bench/{name}/seed"]))
    n = replay_calls(env, reset_opts_calls[name], device, alpha, name)
    check_final_state(env, z, name)
    env.close()
    return n
