"""Chains the "bit-exact on 1e6 seeded trajectories" claim to the reference itself (build container only).

Runs the reference's OWN env files, imported in place from /root/reference on CPU, head to head with the C oracle
on >= 1,048,576 seeded trajectories per integer env (cases in tests/digest_util.py::CASES): every step, every
table, every integer field and the observation must be equal (rewards within the fp32 tolerance).  While doing so
it writes one 64-bit digest per array per step to tests/golden/digests_1m.json.  The CPU suite re-derives that
file from the oracle alone (tests/test_digests_1m.py), the GPU suite from the CUDA kernels
(tests/test_gpu_million.py::test_*_digests_from_reference) -- same seeds, same digests, on the B200.

    python tests/golden/make_digests_1m.py [poker_hu poker_6max blackjack tfe_4x4]      (~25 min on 8 cores)
"""
import json
import sys
import time
from pathlib import Path
from unittest import mock

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import pyoracle, ref_loader  # noqa: E402
from pulselib_b200 import handranks  # noqa: E402
from tests import digest_util as du  # noqa: E402
from tests import poker_util as pu  # noqa: E402

OUT = Path(__file__).resolve().parent / "digests_1m.json"


def poker(name):
    c = du.CASES[name]
    P, M, n, runs, hands, steps = c["P"], c["M"], c["n_run"], c["runs"], c["hands"], c["steps"]
    hr = handranks.load_table()
    out = [dict(reset_obs=0, steps=[dict(obs=0, done=0, state=0, done_count=0, reward_sum=0.0) for _ in range(steps)])
           for _ in range(hands)]
    acts = [du.case_actions(name, h) for h in range(hands)]
    t0 = time.time()
    for run in range(runs):
        base = run * n
        ref = ref_loader.make_poker(hr, n_players=P, max_players=M, n_games=n, **c["cfg"])
        orc = pyoracle.OraclePoker(hr, n_players=P, max_players=M, n_games=n, **c["cfg"])
        for h in range(hands):
            rot = (c["rot"] * h) % P
            ap = c["aps"][h] if "aps" in c else P
            orc.tape_decks(c["seed"], base, h)
            opts = {"prefixed_decks": torch.from_numpy(orc.decks.copy()), "rotation": rot}
            if ap != P:   # the table-size draw of PokerGPU.py:76-80, pinned the way the reference's own tests pin it
                opts["active_players"] = True
                with mock.patch("torch.randint", lambda *a, **k: torch.tensor([ap])):
                    obs_r, _ = ref.reset(options=opts)
            else:
                obs_r, _ = ref.reset(options=opts)
            obs_o = orc.reset(None, active_players=ap, rotation=rot)
            np.testing.assert_array_equal(obs_r.numpy(), obs_o, err_msg=f"{name} run {run} hand {h} reset obs")
            out[h]["reset_obs"] = du.add(out[h]["reset_obs"], du.digest(obs_o, base))
            for t in range(steps):
                a = acts[h][t, base:base + n]
                o_r, r_r, d_r, _, _ = ref.step(torch.from_numpy(a))
                o_o, r_o, d_o = orc.step(a)
                tag = f"{name} run {run} hand {h} step {t}"
                np.testing.assert_array_equal(o_r.numpy(), o_o, err_msg=tag + " obs")
                np.testing.assert_array_equal(d_r.numpy(), d_o, err_msg=tag + " done")
                np.testing.assert_array_equal(ref.equities.numpy(), orc.equities, err_msg=tag + " equities")
                for f in du.POKER_STATE:
                    x, y = getattr(ref, f).numpy().astype(np.int64), getattr(orc, f).astype(np.int64)
                    assert np.array_equal(x, y), f"{tag}: field {f} differs"
                np.testing.assert_allclose(r_r.numpy(), r_o, rtol=pu.REWARD_RTOL, atol=pu.reward_atol(c["cfg"]["alpha"]),
                                           err_msg=tag + " reward")
                s = out[h]["steps"][t]
                s["obs"] = du.add(s["obs"], du.digest(o_o, base))
                s["done"] = du.add(s["done"], du.digest(np.asarray(d_o).astype(np.uint8), base))
                s["state"] = du.add(s["state"], du.poker_state_digest(lambda f: getattr(orc, f), base))
                s["done_count"] += int(np.asarray(d_o).sum())
                s["reward_sum"] += float(r_r.numpy().astype(np.float64).sum())
        print(f"  {name}: run {run + 1}/{runs} equal, {time.time() - t0:.0f} s", flush=True)
    for h in out:
        h["reset_obs"] = du.hexd(h["reset_obs"])
        for s in h["steps"]:
            for k in ("obs", "done", "state"):
                s[k] = du.hexd(s[k])
    return dict(trajectories=runs * n * hands, env_steps_compared=runs * n * hands * steps, hands=out)


def blackjack(name):
    c = du.CASES[name]
    n, runs, steps = c["n_run"], c["runs"], c["steps"]
    mod = ref_loader.blackjack_module()
    acts = du.case_actions(name)
    out = dict(reset_obs=0, steps=[dict(obs=0, rewards=0, terminated=0, terminated_count=0) for _ in range(steps)])
    for run in range(runs):
        base = run * n
        env = mod.BlackJack(device=torch.device("cpu"), batch_size=n)
        orc = pyoracle.OracleBlackjack(n)
        perm = orc.tape_decks(c["seed"], base, 0).copy()
        keys = np.empty((n, 52), np.float32)
        np.put_along_axis(keys, perm.astype(np.int64), np.arange(52, dtype=np.float32)[None, :] / 52.0, axis=1)
        with mock.patch("torch.rand", lambda *a, **k: torch.from_numpy(keys)):
            obs_r, _ = env.reset()
        assert np.array_equal(env.decks.numpy(), perm)
        obs_o = orc.reset(None)
        np.testing.assert_array_equal(obs_r.numpy(), obs_o)
        out["reset_obs"] = du.add(out["reset_obs"], du.digest(obs_o, base))
        for t in range(steps):
            a = acts[t, base:base + n]
            o_r, r_r, d_r, _, _ = env.step(torch.from_numpy(a))
            o_o, r_o, d_o = orc.step(a)
            tag = f"blackjack run {run} step {t}"
            np.testing.assert_array_equal(o_r.numpy(), o_o, err_msg=tag + " obs")
            np.testing.assert_array_equal(r_r.numpy(), r_o, err_msg=tag + " rewards")
            np.testing.assert_array_equal(d_r.numpy(), d_o, err_msg=tag + " terminated")
            for f in ("deck_positions", "player_card_sums", "dealer_card_sums", "players_card_idx", "dealer_card_idx"):
                assert np.array_equal(getattr(env, f).numpy().astype(np.int64), getattr(orc, f).astype(np.int64)), tag + f
            s = out["steps"][t]
            s["obs"] = du.add(s["obs"], du.digest(o_o, base))
            s["rewards"] = du.add(s["rewards"], du.digest(r_o, base))
            s["terminated"] = du.add(s["terminated"], du.digest(np.asarray(d_o).astype(np.uint8), base))
            s["terminated_count"] += int(np.asarray(d_o).sum())
        print(f"  blackjack: run {run + 1}/{runs} equal", flush=True)
    out["reset_obs"] = du.hexd(out["reset_obs"])
    for s in out["steps"]:
        for k in ("obs", "rewards", "terminated"):
            s[k] = du.hexd(s[k])
    return dict(trajectories=runs * n, env_steps_compared=runs * n * steps, episode=out)


def tfe(name):
    c = du.CASES[name]
    n, steps, side, seed = c["n"], c["steps"], c["side"], c["seed"]
    mod = ref_loader.tfe_module()
    orc = pyoracle.OracleTFE(n, side)
    acts = du.case_actions(name)
    draws = []

    def add_tile(board):      # the chance seam of TFE.py:17-34, fed from the tape (rule of oracle_tfe.c)
        u0, u1 = draws.pop(0)
        empties = [(r, c_) for r in range(board.shape[0]) for c_ in range(board.shape[1]) if board[r, c_] == 0]
        if not empties:
            return
        r, c_ = empties[(int(u0) * len(empties)) >> 32]
        board[r, c_] = 4 if int(u1) > 3865470566 else 2

    reset_chance = orc.tape(seed, 0, pyoracle.STREAM_TFE_RESET, 0, 4)
    step_chance = np.stack([orc.tape(seed, 0, pyoracle.STREAM_TFE_SPAWN, t + 1, 2) for t in range(steps)])
    boards = np.zeros((steps, n, side, side), np.int32)
    rewards = np.zeros((steps, n), np.int32)
    done = np.zeros((steps, n), np.uint8)
    scores = np.zeros((steps, n), np.int64)
    reset_boards = np.zeros((n, side, side), np.int32)
    t0 = time.time()
    with mock.patch.object(mod, "add_tile_numba", add_tile):
        for g in range(n):
            env = mod.TFE(side, side)
            draws[:] = [tuple(reset_chance[g, :2]), tuple(reset_chance[g, 2:])]
            obs, _ = env.reset()
            reset_boards[g] = obs
            for t in range(steps):
                draws[:] = [tuple(step_chance[t, g])]
                obs, rew, d, _, info = env.step(int(acts[t, g]))
                boards[t, g] = obs; rewards[t, g] = rew; done[t, g] = d; scores[t, g] = info["score"]
            if g % 16384 == 0:
                print(f"  tfe: board {g}/{n}, {time.time() - t0:.0f} s", flush=True)
    np.testing.assert_array_equal(orc.reset(reset_chance), reset_boards)
    out = dict(reset_boards=du.hexd(du.digest(reset_boards)), steps=[])
    for t in range(steps):
        b, r, d = orc.step(acts[t], step_chance[t])
        np.testing.assert_array_equal(b, boards[t], err_msg=f"tfe step {t} boards")
        np.testing.assert_array_equal(r, rewards[t], err_msg=f"tfe step {t} reward")
        np.testing.assert_array_equal(d.astype(np.uint8), done[t], err_msg=f"tfe step {t} done")
        np.testing.assert_array_equal(orc.scores, scores[t], err_msg=f"tfe step {t} score")
        out["steps"].append(dict(boards=du.hexd(du.digest(boards[t])), rewards=du.hexd(du.digest(rewards[t])),
                                 done=du.hexd(du.digest(done[t])), scores=du.hexd(du.digest(scores[t])),
                                 done_count=int(done[t].sum())))
    return dict(trajectories=n, env_steps_compared=n * steps, episode=out)


def main():
    names = sys.argv[1:] or list(du.CASES)
    res = json.loads(OUT.read_text()) if OUT.exists() else {}
    for name in names:
        t0 = time.time()
        print(f"{name}: reference vs oracle ...", flush=True)
        r = poker(name) if name.startswith("poker") else blackjack(name) if name == "blackjack" else tfe(name)
        r["case"] = {k: v for k, v in du.CASES[name].items()}
        r["seconds"] = round(time.time() - t0, 1)
        r["reference"] = "cerredz/Pulselib imported in place from /root/reference, torch CPU / numba"
        res[name] = r
        OUT.write_text(json.dumps(res, indent=1))
        print(f"{name}: {r['trajectories']} trajectories, {r['env_steps_compared']} env-steps equal, {r['seconds']} s", flush=True)


if __name__ == "__main__":
    main()
